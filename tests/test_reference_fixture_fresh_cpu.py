"""Where the reference tree is present (the build container; not the GPU box), re-run two fixture cases from the
reference's own source and check that the committed tests/golden/reference_golden.npz still holds exactly those outputs —
the fixtures are generated, not edited.  Runs in a subprocess: the jax stand-in sets torch's default dtype to float64."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("MCMC_REFERENCE", "/root/reference")

SNIPPET = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1])
import make_reference_golden as G
blob = np.load(sys.argv[2])
for c in G.CASES:
    if c["id"] in ("hmc/rosenbrock", "grahmc/tanh/neals_funnel", "nuts/standard_normal/depth_cap2"):
        ins, out = G.run_case(c)
        for k, v in {**ins, **out}.items():
            assert np.array_equal(np.asarray(v), blob[c["id"] + "|" + k], equal_nan=True), (c["id"], k)
        print("fresh", c["id"])
"""


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "samplers")), reason="reference tree not present on this machine")
def test_committed_fixtures_are_what_the_reference_source_produces():
    golden = os.path.join(ROOT, "tests", "golden")
    r = subprocess.run([sys.executable, "-c", SNIPPET, golden, os.path.join(golden, "reference_golden.npz")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("fresh") == 3


SNIPPET_LEAPFROG = r"""
import sys, numpy as np
sys.path.insert(0, sys.argv[1])
import make_reference_leapfrog_golden as G
blob = np.load(sys.argv[2])
for c in G.CASES:
    if c["id"] in ("leapfrog/rosenbrock", "leapfrog/log_gamma", "trajectory/sigmoid/neals_funnel", "trajectory/constant/rosenbrock_L5"):
        ins, out = G.run_case(c)
        for k, v in {**ins, **out}.items():
            assert np.array_equal(np.asarray(v), blob[c["id"] + "|" + k], equal_nan=True), (c["id"], k)
        print("fresh", c["id"])
"""


@pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "samplers")), reason="reference tree not present on this machine")
def test_committed_leapfrog_fixtures_are_what_the_reference_source_produces():
    golden = os.path.join(ROOT, "tests", "golden")
    r = subprocess.run([sys.executable, "-c", SNIPPET_LEAPFROG, golden, os.path.join(golden, "reference_leapfrog_golden.npz")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert r.stdout.count("fresh") == 4
