"""Opt-in: the GPU parity suite re-run under an experimental kernel variant.

The forward kernel of bf16 d = 128 is chosen once per process from B200_FWD_VARIANT (csrc/fwd_sm100.cu, dispatch_fwd) and
the backward's issue loop from B200_BWD_VARIANT (csrc/bwd_sm100.cu, bwd_launch_d128), so every variant gets its own pytest
subprocess over the modules that reach it: plain / causal / strided and grouped-head layouts / host-buffer pipeline /
single-GPU ring emulation.  This is the gate a variant has to pass before it may become the default:

    B200_TEST_VARIANTS=1 python -m pytest tests/test_gpu_kernel_variants.py -m gpu -q

It is skipped otherwise, so the default `-m gpu` run only exercises the shipped configuration.
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FWD_MODULES = ["tests/test_gpu_fwd_tc.py", "tests/test_gpu_causal.py", "tests/test_gpu_layouts.py", "tests/test_gpu_host.py",
               "tests/test_gpu_ring.py"]
BWD_MODULES = ["tests/test_gpu_bwd_tc.py", "tests/test_gpu_causal.py", "tests/test_gpu_layouts.py", "tests/test_gpu_host.py",
               "tests/test_gpu_ring.py"]
# (environment, modules): see the tables in DESIGN.md 4.1 / 7
VARIANTS = [
    # 64-key tiles, double-buffered scores, one issuer per tile, 4/16 exp2 on the FMA pipe: the promotion candidate
    ({"B200_FWD_VARIANT": "4", "B200_DBUF_POLY": "4"}, FWD_MODULES),
    ({"B200_FWD_VARIANT": "4", "B200_DBUF_POLY": "0"}, FWD_MODULES),
    # + branch-free steady-state issue loops (never run on a GPU yet)
    ({"B200_FWD_VARIANT": "4", "B200_DBUF_POLY": "4", "B200_DBUF_PEEL": "1"}, FWD_MODULES),
    # one shared score buffer, P in its own TMEM columns
    ({"B200_FWD_VARIANT": "6", "B200_SHS_POLY": "0"}, FWD_MODULES),
    # CTA pairs (full tiles only; other shapes fall back to the default kernel)
    ({"B200_FWD_VARIANT": "3"}, FWD_MODULES),
    # default forward kernel with one issuer per tile
    ({"B200_FWD_VARIANT": "5"}, FWD_MODULES),
    # backward: steady state of the issue loop unrolled by six, dQ-free wait merged into the dS wait (never run on a GPU yet)
    ({"B200_BWD_VARIANT": "1"}, BWD_MODULES),
]


@pytest.mark.skipif(os.environ.get("B200_TEST_VARIANTS") != "1", reason="opt-in: set B200_TEST_VARIANTS=1")
@pytest.mark.parametrize("extra,modules", VARIANTS, ids=["-".join(f"{k[5:].lower()}{v}" for k, v in e.items()) for e, _ in VARIANTS])
def test_suite_under_variant(extra, modules):
    env = dict(os.environ, **extra)
    env.pop("B200_TEST_VARIANTS", None)                       # no recursion
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", *modules], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
