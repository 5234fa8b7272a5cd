"""b200_fa2_bwd_static (include/b200attn.h): the batched backward whose scratch is a fixed device array inside the
library, for callers that must not allocate.  No counterpart in the reference.  Checked through the C-ABI against the
existing entry point b200_fa2_bwd on the same inputs, on consecutive calls that share the array (the second and third
problems are smaller / of another dtype than the first, so stale contents of the array would show).
Tolerance: the suite's own bounds -- bf16: max-abs <= 2e-2 and rel-L2 <= 1e-2; fp32 (tf32 operands): max-abs <= 2e-3.
Named test_zz_* so that pytest collects it last."""
import numpy as np
import pytest
import torch

import cuda_flashattention_b200 as fa
from cuda_flashattention_b200 import _lib
from gpu_util import dev, host, rel_l2

lib = fa.lib


def test_static_scratch_rejects_sizes_beyond_the_maximum_without_a_device():
    # pure host check, before anything touches a device: C3's backward needs ~4.3 GB of scratch
    assert lib.b200_fa2_bwd_workspace_bytes(8, 16, 8192, 8192, 128, 0) > _lib.B200_STATIC_SCRATCH_BYTES
    st = lib.b200_fa2_bwd_static(8, 8, 8, 8, 8, 8, 8, 8, 8, 8, 16, 8192, 8192, 128, 0.0, 0, None)
    assert st == -6, st                                              # B200_ERR_WORKSPACE
    assert b"workspace" in lib.b200_strerror(st)
    # argument validation and the empty problem behave as in b200_fa2_bwd
    assert lib.b200_fa2_bwd_static(None, None, None, None, None, None, None, None, None, 1, 1, 64, 64, 128, 0.0, 7, None) == -1
    assert lib.b200_fa2_bwd_static(None, None, None, None, None, None, None, None, None, 0, 4, 64, 64, 128, 0.0, 0, None) == 0
    assert lib.b200_fa2_bwd_static(None, None, None, None, None, None, None, None, None, -1, 4, 64, 64, 128, 0.0, 0, None) == -1


def _problem(B, H, Nq, Nkv, d, dtype, seed):
    rng = np.random.default_rng(seed)
    mk = (lambda s: dev(rng.standard_normal(s).astype(np.float32), dtype)) if dtype == torch.bfloat16 else \
         (lambda s: dev((rng.random(s) - 0.5).astype(np.float32), dtype))
    q, do = mk((B, H, Nq, d)), mk((B, H, Nq, d))
    k, v = mk((B, H, Nkv, d)), mk((B, H, Nkv, d))
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0)
    return q, k, v, O, L, do


def _both(B, H, Nq, Nkv, d, dtype, seed):
    q, k, v, O, L, do = _problem(B, H, Nq, Nkv, d, dtype, seed)
    code = _lib.B200_DTYPE_BF16 if dtype == torch.bfloat16 else _lib.B200_DTYPE_F32
    need = lib.b200_fa2_bwd_workspace_bytes(B, H, Nq, Nkv, d, code)
    assert 0 < need <= _lib.B200_STATIC_SCRATCH_BYTES
    st = torch.cuda.current_stream().cuda_stream
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")        # the existing entry point with a caller workspace
    want = [torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)]
    got = [torch.full_like(q, float("nan")), torch.full_like(k, float("nan")), torch.full_like(v, float("nan"))]
    p = lambda t: t.data_ptr()
    _lib.check(lib.b200_fa2_bwd(p(q), p(k), p(v), p(O), p(L), p(do), p(want[0]), p(want[1]), p(want[2]), B, H, Nq, Nkv, d, 0.0, code,
                              p(ws), need, st), "b200_fa2_bwd")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    _lib.check(lib.b200_fa2_bwd_static(p(q), p(k), p(v), p(O), p(L), p(do), p(got[0]), p(got[1]), p(got[2]), B, H, Nq, Nkv, d, 0.0,
                                     code, st), "b200_fa2_bwd_static")
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    tol_abs, tol_rel = (2e-2, 1e-2) if dtype == torch.bfloat16 else (2e-3, 2e-3)
    for name, g, w in zip(("dQ", "dK", "dV"), got, want):
        g, w = host(g), host(w)
        assert np.isfinite(g).all(), name
        mx = float(np.abs(g - w).max()) / max(1.0, float(np.abs(w).max()))
        rl = rel_l2(g, w)
        assert mx <= tol_abs and rl <= tol_rel, (name, mx, rl)
    return free0, free1


@pytest.mark.gpu
def test_static_scratch_backward_matches_the_workspace_entry_point_on_consecutive_calls():
    assert fa.uses_tensor_cores(torch.bfloat16, 128, 1024, 1024, backward=True)
    _both(1, 4, 1024, 1024, 128, torch.bfloat16, 1)                  # call 1: 2.1 MB of fp32 dQ accumulator + statistics
    _both(2, 2, 200, 333, 128, torch.bfloat16, 2)                    # call 2: smaller, ragged -- finds call 1's contents
    _both(1, 2, 320, 320, 64, torch.float32, 3)                      # call 3: the tf32 kernel's workspace layout on top
    _both(1, 1, 48, 48, 32, torch.float32, 4)                        # call 4: generic CUDA-core path (only D in the scratch)
    # every kernel is loaded by now: a repeat of call 1 must leave the device's free memory exactly where it was
    free0, free1 = _both(1, 4, 1024, 1024, 128, torch.bfloat16, 5)
    assert free1 == free0, (free0, free1)
