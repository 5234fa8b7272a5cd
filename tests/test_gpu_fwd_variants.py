"""Opt-in: the whole forward-facing GPU parity suite, re-run under an experimental forward kernel.

The forward kernel of bf16 d = 128 is chosen once per process from B200_FWD_VARIANT (csrc/fwd_sm100.cu, dispatch_fwd),
so every variant gets its own pytest subprocess over the modules that reach the forward: plain / causal / strided and
grouped-head layouts / host-buffer pipeline / single-GPU ring emulation.  This is the gate a variant has to pass before
it may become the default: run it with

    B200_TEST_VARIANTS=1 python -m pytest tests/test_gpu_fwd_variants.py -m gpu -q

It is skipped otherwise, so the default `-m gpu` run only exercises the shipped configuration.
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MODULES = ["tests/test_gpu_fwd_tc.py", "tests/test_gpu_causal.py", "tests/test_gpu_layouts.py", "tests/test_gpu_host.py",
           "tests/test_gpu_ring.py"]
# (B200_FWD_VARIANT, extra environment): see the table in DESIGN.md 7
VARIANTS = [
    ("4", {"B200_DBUF_POLY": "4"}),     # 64-key tiles, double-buffered scores, one issuer per tile, 4/16 exp2 on the FMA pipe
    ("4", {"B200_DBUF_POLY": "0"}),
    ("4", {"B200_DBUF_POLY": "4", "B200_DBUF_PEEL": "1"}),   # + branch-free steady-state issue loops (never run on a GPU yet)
    ("6", {"B200_SHS_POLY": "0"}),      # one shared score buffer, P in its own TMEM columns
    ("3", {}),                          # CTA pairs (full tiles only; other shapes fall back to the default kernel)
    ("5", {}),                          # default kernel with one issuer per tile
]


@pytest.mark.skipif(os.environ.get("B200_TEST_VARIANTS") != "1", reason="opt-in: set B200_TEST_VARIANTS=1")
@pytest.mark.parametrize("variant,extra", VARIANTS, ids=[f"v{v}-{'-'.join(e.values()) or 'x'}" for v, e in VARIANTS])
def test_forward_suite_under_variant(variant, extra):
    env = dict(os.environ, B200_FWD_VARIANT=variant, **extra)
    env.pop("B200_TEST_VARIANTS", None)                       # no recursion
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", *MODULES], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
