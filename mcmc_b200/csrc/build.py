"""Builds libmcmc_b200.so (fast) and libmcmc_b200_strict.so (reference operation order,
no FMA contraction) in-tree with nvcc for sm_100a.  Called by __graft_entry__.build().

One translation unit per (sampler family, target) so the targets compile in parallel;
objects are cached under build/<variant>/ and rebuilt when any source or flag changes.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
NUM_TARGETS = 9
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
BASE = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
        "-Xcompiler", "-fPIC"]
VARIANTS = {
    "fast": (["-fmad=true"], "libmcmc_b200.so"),
    "strict": (["-DMCMC_FAITHFUL", "-fmad=false", "-prec-div=true", "-prec-sqrt=true"], "libmcmc_b200_strict.so"),
}
HEADERS = ["common.cuh", "philox.cuh", "targets.cuh", "rowstore.cuh", "hmc_kernel.cuh", "rwmh_kernel.cuh",
           "nuts_kernel.cuh", os.path.join("..", "..", "include", "mcmc_b200.h")]


def _units():
    yield "abi.cu", None
    yield "diag.cu", None
    for fam in ("hmc_inst.cu", "rwmh_inst.cu", "nuts_inst.cu"):
        for t in range(NUM_TARGETS):
            yield fam, t


def _digest(flags) -> str:
    h = hashlib.sha256(" ".join(flags).encode())
    for f in sorted(os.listdir(HERE)):
        if f.endswith((".cu", ".cuh", ".inc")):
            h.update(open(os.path.join(HERE, f), "rb").read())
    h.update(open(os.path.join(HERE, "..", "..", "include", "mcmc_b200.h"), "rb").read())
    return h.hexdigest()[:16]


def build(variants=("fast", "strict"), verbose=False, jobs=None, extra_flags=(), suffix="") -> list[str]:
    """`extra_flags` + `suffix` build an experiment next to the product libraries (e.g. extra_flags=["-DMCMC_STEP_UNROLL=2"],
    suffix="_u2" -> libmcmc_b200_u2.so, objects under build/fast_u2/); load it with MCMC_B200_LIB=<path>."""
    if not os.path.exists(NVCC):
        raise RuntimeError(f"nvcc not found at {NVCC}")
    jobs = jobs or max(1, (os.cpu_count() or 4))
    outs = []
    for var in variants:
        vflags, libname = VARIANTS[var]
        flags = BASE + vflags + list(extra_flags)
        tag = _digest(flags)
        odir = os.path.join(HERE, "build", var + suffix)
        libname = libname.replace(".so", suffix + ".so")
        os.makedirs(odir, exist_ok=True)
        lib = os.path.join(HERE, libname)
        stamp = os.path.join(odir, "digest.txt")
        if os.path.exists(lib) and os.path.exists(stamp) and open(stamp).read().strip() == tag:
            outs.append(lib)
            continue

        def compile_one(unit):
            src, t = unit
            obj = os.path.join(odir, src.replace(".cu", "") + ("" if t is None else f"_t{t}") + ".o")
            cmd = [NVCC] + flags + (["-DMCMC_TID=%d" % t] if t is not None else []) + ["-c", os.path.join(HERE, src), "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src} t={t} [{var}]:\n{r.stdout}\n{r.stderr}")
            if verbose and r.stderr.strip():
                print(r.stderr, file=sys.stderr)
            return obj

        with ThreadPoolExecutor(max_workers=jobs) as ex:
            objs = list(ex.map(compile_one, list(_units())))
        cmd = [NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed [{var}]:\n{r.stdout}\n{r.stderr}")
        open(stamp, "w").write(tag)
        outs.append(lib)
    return outs


def build_xla_ffi(out_dir: str | None = None, include_dir: str | None = None, verbose: bool = False) -> str | None:
    """Compiles mcmc_xla_ffi.cc (the jax.ffi custom-call handlers over the C ABI) into libmcmc_b200_xla_ffi.so, linked
    against libmcmc_b200.so.  The XLA FFI header ships inside jaxlib: with `import jax` unavailable (this image) the
    shim is SKIPPED with a message and None is returned — unless `include_dir` is given (any directory: without the
    header in it the source compiles to its stub, which only exports mcmc_xla_ffi_available() == 0)."""
    if include_dir is None:
        try:
            import jax  # noqa: F401
            include_dir = jax.ffi.include_dir()
        except Exception as e:
            if verbose:
                print(f"xla ffi shim skipped: jax is not importable ({type(e).__name__}); "
                      "mcmc_b200/csrc/mcmc_xla_ffi.cc builds where jaxlib is installed", file=sys.stderr)
            return None
    out_dir = out_dir or HERE
    out = os.path.join(out_dir, "libmcmc_b200_xla_ffi.so")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", os.path.join(HERE, "mcmc_xla_ffi.cc"), "-I", include_dir,
           "-I", "/usr/local/cuda/include", "-o", out]
    main_lib = os.path.join(HERE, VARIANTS["fast"][1])
    if os.path.exists(main_lib):
        cmd += [main_lib, "-Wl,-rpath," + HERE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"xla ffi shim failed to compile:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    # python -m mcmc_b200.csrc.build [fast|strict ...] [--suffix _x -DFOO=1 ...]
    argv = sys.argv[1:]
    sfx = ""
    if "--suffix" in argv:
        i = argv.index("--suffix")
        sfx = argv[i + 1]
        del argv[i:i + 2]
    extra = [a for a in argv if a.startswith("-")]
    vs = tuple(a for a in argv if not a.startswith("-")) or (("fast",) if sfx else ("fast", "strict"))
    for p in build(vs, verbose=True, extra_flags=extra, suffix=sfx):
        print(p)
