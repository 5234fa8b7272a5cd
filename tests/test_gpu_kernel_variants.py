"""Opt-in: the GPU parity suite re-run under an experimental kernel variant.

The product library selects its kernels on shape and dtype only.  The measured-negative / unpromoted variants
(csrc/experiments/, POLY sweeps) are compiled into libb200attn_exp.so (`make -C cuda_flashattention_b200 EXPERIMENTS=1`),
where B200_FWD_VARIANT / B200_DBUF_POLY pick one per process; every variant gets its own pytest subprocess (with
B200ATTN_LIB pointing at the side library) over the modules that reach it.  This is the gate a variant has to pass
before it may become the default -- fwd_dbuf_kernel<4> went through it in r2:

    B200_TEST_VARIANTS=1 python -m pytest tests/test_gpu_kernel_variants.py -m gpu -q

Skipped otherwise (and when the side library is not built), so the default `-m gpu` run only exercises what ships.
"""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FWD_MODULES = ["tests/test_gpu_fwd_tc.py", "tests/test_gpu_causal.py", "tests/test_gpu_layouts.py", "tests/test_gpu_host.py",
               "tests/test_gpu_ring.py"]
EXP_LIB = os.path.join(ROOT, "cuda_flashattention_b200", "libb200attn_exp.so")
# (environment, modules): see the table in DESIGN.md 4.1
VARIANTS = [
    ({"B200_FWD_VARIANT": "1"}, FWD_MODULES),                              # the r1 default: 128-key tiles, P in place
    ({"B200_FWD_VARIANT": "4", "B200_DBUF_POLY": "0"}, FWD_MODULES),      # shipped kernel without the FMA-pipe exp2
    ({"B200_FWD_VARIANT": "4", "B200_DBUF_POLY": "4", "B200_DBUF_PEEL": "1"}, FWD_MODULES),   # branch-free issue loops (-1.5 %)
    ({"B200_FWD_VARIANT": "6", "B200_SHS_POLY": "0"}, FWD_MODULES),       # one shared score buffer, P in its own TMEM columns
    ({"B200_FWD_VARIANT": "3"}, FWD_MODULES),                              # CTA pairs (full tiles only)
    ({"B200_FWD_VARIANT": "5"}, FWD_MODULES),                              # r1 kernel with one issuer per tile
]


@pytest.mark.skipif(os.environ.get("B200_TEST_VARIANTS") != "1" or not os.path.exists(EXP_LIB),
                    reason="opt-in: set B200_TEST_VARIANTS=1 and build libb200attn_exp.so")
@pytest.mark.parametrize("extra,modules", VARIANTS, ids=["-".join(f"{k[5:].lower()}{v}" for k, v in e.items()) for e, _ in VARIANTS])
def test_suite_under_variant(extra, modules):
    env = dict(os.environ, B200ATTN_LIB=EXP_LIB, **extra)
    env.pop("B200_TEST_VARIANTS", None)                       # no recursion
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-m", "gpu", *modules], cwd=ROOT, env=env,
                       capture_output=True, text=True, timeout=1800)
    assert r.returncode == 0, r.stdout[-4000:] + r.stderr[-2000:]
