"""GPU parity of the causal extension (B200_FLAG_CAUSAL, include/b200attn.h) against the oracle's
causal forward / backward.  The reference is non-causal (its README lists the mask as the next
step); the oracle's causal form is pinned in tests/test_oracle.py as the reference's row arithmetic
on a truncated key range.  Tolerances are the north_star ones: bf16 max-abs <= 2e-2, rel-L2 <= 1e-2;
tf32 / fp32 <= 2e-3."""
import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa
from cuda_flashattention_b200._lib import lib
from gpu_util import dev, host, rel_l2

pytestmark = pytest.mark.gpu


def _rand(shape, rng, dtype, s=1.0):
    return dev(s * rng.standard_normal(shape).astype(np.float32), dtype)


@pytest.mark.parametrize("dtype,d,N", [
    (torch.bfloat16, 128, 64), (torch.bfloat16, 128, 128), (torch.bfloat16, 128, 256), (torch.bfloat16, 128, 384),
    (torch.bfloat16, 128, 1000), (torch.bfloat16, 128, 2048), (torch.bfloat16, 64, 777), (torch.bfloat16, 64, 1024),
    (torch.float32, 64, 300), (torch.float32, 64, 1024), (torch.float32, 128, 200), (torch.float32, 128, 1100),
])
def test_causal_forward_tensor_core(dtype, d, N):
    rng = np.random.default_rng(1000 + d + N)
    q, k, v = [_rand((N, d), rng, dtype) for _ in range(3)]
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=True)
    Oo, Lo = oracle.forward(host(q), host(k), host(v), 0.0, causal=True)
    # The first rows average over a handful of keys, so |O| reaches |V|max (~4) instead of the ~0.1 of
    # a full row: the absolute bound is scaled by that magnitude (tf32 rounds V to 2^-11 relative).
    tol = (2e-2 if dtype == torch.bfloat16 else 2e-3) * max(1.0, float(np.abs(Oo).max()))
    assert np.abs(host(O) - Oo).max() <= tol
    assert rel_l2(host(O), Oo) <= (1e-2 if dtype == torch.bfloat16 else 2e-3)
    assert np.abs(host(L) - Lo).max() <= (2e-2 if dtype == torch.bfloat16 else 2e-3)
    # row 0 sees key 0 only: O[0] = V[0]
    assert np.abs(host(O)[0] - host(v)[0]).max() <= tol


@pytest.mark.parametrize("dtype,d,Nq,Nkv", [
    (torch.float32, 32, 50, 50), (torch.float32, 96, 100, 260), (torch.bfloat16, 128, 200, 333),
    (torch.float32, 64, 1, 129), (torch.bfloat16, 64, 128, 640),
    # r2e: cross-length causal runs on the tensor-core kernels too (chunked prefill / decode shapes)
    (torch.bfloat16, 128, 256, 1024), (torch.bfloat16, 128, 1000, 1500), (torch.bfloat16, 128, 64, 2048), (torch.bfloat16, 128, 130, 131),
    (torch.bfloat16, 64, 333, 1000), (torch.float32, 64, 300, 700), (torch.float32, 128, 100, 400),
])
def test_causal_generic_and_bottom_right_alignment(dtype, d, Nq, Nkv):
    # Nq < Nkv: query i sees keys 0 .. i + (Nkv - Nq) (bottom-right aligned).  d in {64, 128}: tensor-core kernels (fp32:
    # tf32 forward; tf32 backward at d = 64, generic backward at d = 128); other d: the CUDA-core kernels
    rng = np.random.default_rng(Nq * 3 + Nkv)
    q = _rand((Nq, d), rng, dtype)
    k, v = _rand((Nkv, d), rng, dtype), _rand((Nkv, d), rng, dtype)
    do = _rand((Nq, d), rng, dtype)
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=True)
    Oo, Lo = oracle.forward(host(q), host(k), host(v), 0.0, causal=True)
    tc = d in (64, 128)
    assert lib.b200_fa2_fwd_route(q.data_ptr(), k.data_ptr(), v.data_ptr(), O.data_ptr(), Nq, Nkv, d, 0.0,
                                  0 if dtype == torch.bfloat16 else 1, 1) == (1 if tc else 0)
    tol = 2e-2 if dtype == torch.bfloat16 else (2e-3 if tc else 1e-4)      # bf16 / tf32 / exact fp32
    assert np.abs(host(O) - Oo).max() <= tol and np.abs(host(L) - Lo).max() <= tol
    dQ, dK, dV = fa.flash_attention_2_backward(q, k, v, O, L, do, 0.0, causal=True)
    ref = oracle.backward(host(q), host(k), host(v), host(do), 0.0, causal=True)
    for got, want in zip((dQ, dK, dV), ref):
        assert np.abs(host(got) - want).max() <= tol * max(1.0, float(np.abs(want).max()))
        assert rel_l2(host(got), want) <= (1e-2 if dtype == torch.bfloat16 else (2e-3 if tc else 1e-4))


def test_causal_cross_lengths_batched_tensor_core():
    """Batched, several heads, Nq < Nkv with both ragged: every (b, h) slab against the oracle's causal rows."""
    rng = np.random.default_rng(99)
    B, H, Nq, Nkv, d = 2, 3, 300, 777, 128
    q, do = _rand((B, H, Nq, d), rng, torch.bfloat16), _rand((B, H, Nq, d), rng, torch.bfloat16)
    k, v = _rand((B, H, Nkv, d), rng, torch.bfloat16), _rand((B, H, Nkv, d), rng, torch.bfloat16)
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=True)
    dQ, dK, dV = fa.flash_attention_2_backward(q, k, v, O, L, do, 0.0, causal=True)
    for b in range(B):
        for h in range(H):
            Oo, Lo = oracle.forward(host(q[b, h]), host(k[b, h]), host(v[b, h]), 0.0, causal=True)
            assert np.abs(host(O[b, h]) - Oo).max() <= 2e-2 and np.abs(host(L[b, h]) - Lo).max() <= 2e-2
            ref = oracle.backward(host(q[b, h]), host(k[b, h]), host(v[b, h]), host(do[b, h]), 0.0, causal=True)
            for name, got, want in zip(("dQ", "dK", "dV"), (dQ, dK, dV), ref):
                g = host(got[b, h])
                assert np.abs(g - want).max() <= 2e-2 * max(1.0, float(np.abs(want).max())), name
                assert rel_l2(g, want) <= 1e-2, name


def test_causal_needs_nq_le_nkv():
    q = torch.zeros(8, 64, device="cuda")
    k = torch.zeros(4, 64, device="cuda")
    with pytest.raises(fa.B200Error):
        fa.flash_attention_2_forward(q, k, k, 0.0, causal=True)


@pytest.mark.parametrize("N,B,H,d", [(64, 1, 1, 128), (128, 1, 1, 128), (192, 1, 1, 128), (256, 1, 1, 128), (1000, 1, 1, 128),
                                     (2048, 1, 1, 128), (640, 2, 3, 128), (1000, 1, 2, 64)])
def test_causal_backward_tensor_core(N, B, H, d):
    rng = np.random.default_rng(77 + N)
    q, k, v, do = [_rand((B, H, N, d), rng, torch.bfloat16) for _ in range(4)]
    assert fa.uses_tensor_cores(torch.bfloat16, d, N, N, backward=True)
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=True)
    dQ, dK, dV = fa.flash_attention_2_backward(q, k, v, O, L, do, 0.0, causal=True)
    for b in range(B):
        for h in range(H):
            ref = oracle.backward(host(q[b, h]), host(k[b, h]), host(v[b, h]), host(do[b, h]), 0.0, causal=True)
            for name, got, want in zip(("dQ", "dK", "dV"), (dQ, dK, dV), ref):
                g = host(got[b, h])
                assert np.abs(g - want).max() <= 2e-2 * max(1.0, float(np.abs(want).max())), name
                assert rel_l2(g, want) <= 1e-2, name


def test_causal_prefix_property_full_size():
    # size-independent property at a bench-sized sequence: under the causal mask the first n rows of
    # the output depend on the first n keys only, so they equal the causal result of the length-n
    # prefix problem (same tiles in the same order; the tolerance only covers the bf16 store).
    rng = np.random.default_rng(5)
    N, n, d = 8192, 1024, 128
    q, k, v = [_rand((1, 2, N, d), rng, torch.bfloat16) for _ in range(3)]
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=True)
    Op, Lp = fa.flash_attention_2_forward(q[:, :, :n].contiguous(), k[:, :, :n].contiguous(), v[:, :, :n].contiguous(),
                                          0.0, causal=True)
    assert np.abs(host(O[:, :, :n]) - host(Op)).max() <= 1e-2
    assert np.abs(host(L[:, :, :n]) - host(Lp)).max() <= 1e-3
    # and the last row sees everything: it is the non-causal last row
    On, Ln = fa.flash_attention_2_forward(q[:, :, -128:].contiguous(), k, v, 0.0)
    assert np.abs(host(O[:, :, -1]) - host(On[:, :, -1])).max() <= 1e-2
    assert np.abs(host(L[:, :, -1]) - host(Ln[:, :, -1])).max() <= 1e-3
