"""Condenses an `ncu --set full` report into the text summary kept under profiles/.

usage: ncu_summary.py <report.ncu-rep> [<object with the kernel> <mangled-name substring>] [--title "..."]
Prints, per captured launch, the metrics the roofline argument rests on (duration, DRAM bytes, fp64-pipe
utilisation, issue-slot utilisation, registers / occupancy limits, stall-sample histogram); with an object
file it appends tools/ncu_lines.py's per-source-line table of the first launch.
"""
import csv
import io
import os
import re
import subprocess
import sys

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|launch__(block_size|grid_size|registers_per_thread|"
    r"waves_per_multiprocessor|occupancy_limit_\w+|shared_mem_per_block_\w+)|sm__cycles_elapsed\.avg(\.per_second)?|"
    r"sm__inst_executed_pipe_(fp64|alu|fma|fmaheavy|xu|lsu|uniform)\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_fp64_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
    r"sm__throughput\.avg\.pct_of_peak_sustained_elapsed|gpu__compute_memory_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"smsp__inst_executed\.sum|smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.avg\.per_cycle_active|"
    r"sm__inst_executed\.avg\.per_cycle_(active|elapsed)|smsp__pcsamp_warps_issue_stalled_\w+|"
    r"smsp__sass_thread_inst_executed_op_(dfma|dmul|dadd)_pred_on\.sum\.per_cycle_elapsed|"
    r"smsp__thread_inst_executed_per_inst_executed\.ratio|lts__t_sector_hit_rate\.pct|l1tex__t_sector_hit_rate\.pct|"
    r"lts__t_bytes\.sum(\.per_second)?|smsp__sass_inst_executed_op_local_(ld|st)\.sum|sm__sass_inst_executed_op_local\w*\.sum)$")


def main():
    args = sys.argv[1:]
    title = None
    if "--title" in args:
        i = args.index("--title")
        title = args[i + 1]
        del args[i:i + 2]
    rep = args[0]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    ik = hdr.index("Kernel Name")
    if title:
        print(f"# {title}")
    print(f"# source: {os.path.basename(rep)} (ncu --set full --clock-control none --import-source on); one block per captured launch")
    for r in rows[2:]:
        if len(r) != len(hdr):
            continue
        name = r[ik]
        print(f"## launch {r[hdr.index('ID')]}: {name[:140]}")
        for h, u, v in zip(hdr, units, r):
            if KEEP.match(h) and v not in ("", "n/a"):
                if h.startswith("smsp__pcsamp") and (v == "0" or h.endswith("_not_issued")):
                    continue
                print(f"{h},{u},{v}")
    if len(args) >= 3:
        here = os.path.dirname(os.path.abspath(__file__))
        out = subprocess.run([sys.executable, os.path.join(here, "ncu_lines.py"), rep, args[1], args[2]],
                             capture_output=True, text=True).stdout
        print("# --- per source line / opcode (tools/ncu_lines.py, first launch)")
        sys.stdout.write(out)


if __name__ == "__main__":
    main()
