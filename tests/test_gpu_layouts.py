"""GPU parity of the descriptor entry points (b200_fa2_fwd_desc / b200_fa2_bwd_desc): token-major and
otherwise strided tensors, grouped K/V heads (GQA / MQA), and the autograd function built on them.
Every (b, h) slab is checked against the oracle's single-head forward / backward; for grouped heads the
oracle's dK / dV of the query heads of a group are summed.  Tolerances: north_star (bf16 2e-2 / 1e-2;
tf32 2e-3)."""
import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa
from gpu_util import host, rel_l2

pytestmark = pytest.mark.gpu


def _rand(shape, seed, dtype=torch.bfloat16, s=1.0):
    rng = np.random.default_rng(seed)
    return torch.from_numpy((s * rng.standard_normal(shape)).astype(np.float32)).to("cuda").to(dtype)


def _bhnd(t, layout):
    return t if layout == "bhnd" else t.permute(0, 2, 1, 3)


def _check_fwd_bwd(q, k, v, do, layout, causal, dtype, bwd=True):
    O, L = fa.attention_forward(q, k, v, causal=causal, layout=layout)
    qh, kh, vh, Oh = [_bhnd(t, layout) for t in (q, k, v, O)]
    doh = _bhnd(do, layout) if do is not None else None
    B, H, Nq, d = qh.shape
    Hkv = kh.shape[1]
    G = H // Hkv
    tol = 2e-2 if dtype == torch.bfloat16 else 2e-3
    rl = 1e-2 if dtype == torch.bfloat16 else 2e-3
    if bwd:
        dQ, dK, dV = fa.attention_backward(q, k, v, O, L, do, causal=causal, layout=layout)
        dQh, dKh, dVh = [_bhnd(t, layout) for t in (dQ, dK, dV)]
    for b in range(B):
        for hk in range(Hkv):
            dK_ref = dV_ref = None
            for g in range(G):
                h = hk * G + g
                args = (host(qh[b, h]), host(kh[b, hk]), host(vh[b, hk]))
                Oo, Lo = oracle.forward(*args, 0.0, causal=causal)
                scl = max(1.0, float(np.abs(Oo).max())) if causal else 1.0
                assert np.abs(host(Oh[b, h]) - Oo).max() <= tol * scl, ("O", b, h)
                assert np.abs(host(L[b, h]) - Lo).max() <= tol, ("L", b, h)
                if bwd:
                    rq, rk, rv = oracle.backward(*args, host(doh[b, h]), 0.0, causal=causal)
                    assert np.abs(host(dQh[b, h]) - rq).max() <= tol * max(1.0, float(np.abs(rq).max())), ("dQ", b, h)
                    assert rel_l2(host(dQh[b, h]), rq) <= rl, ("dQ", b, h)
                    dK_ref = rk if dK_ref is None else dK_ref + rk
                    dV_ref = rv if dV_ref is None else dV_ref + rv
            if bwd:
                for name, got, want in (("dK", dKh[b, hk], dK_ref), ("dV", dVh[b, hk], dV_ref)):
                    assert np.abs(host(got) - want).max() <= tol * max(1.0, float(np.abs(want).max())), (name, b, hk)
                    assert rel_l2(host(got), want) <= rl, (name, b, hk)


@pytest.mark.parametrize("d,N,causal", [(128, 300, False), (128, 512, True), (64, 200, False)])
def test_token_major_bnhd(d, N, causal):
    B, H = 2, 3
    q, k, v, do = [_rand((B, N, H, d), 10 + i + N) for i in range(4)]
    _check_fwd_bwd(q, k, v, do, "bnhd", causal, torch.bfloat16)


@pytest.mark.parametrize("H,Hkv,causal", [(4, 2, False), (4, 1, False), (6, 2, True)])
def test_grouped_kv_heads(H, Hkv, causal):
    B, N, d = 2, 384, 128
    q, do = _rand((B, N, H, d), 1), _rand((B, N, H, d), 2)
    k, v = _rand((B, N, Hkv, d), 3), _rand((B, N, Hkv, d), 4)
    _check_fwd_bwd(q, k, v, do, "bnhd", causal, torch.bfloat16)


def test_slices_of_a_packed_qkv_projection():
    # q, k, v are views into one [B, N, 3, H, d] tensor: row stride 3*H*d, nothing is copied
    B, N, H, d = 1, 640, 2, 128
    qkv = _rand((B, N, 3, H, d), 7)
    q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
    assert not q.is_contiguous() and q.stride(-1) == 1
    _check_fwd_bwd(q, k, v, _rand((B, N, H, d), 8), "bnhd", False, torch.bfloat16)


def test_cross_lengths_bhnd_grouped():
    B, H, Hkv, Nq, Nkv, d = 1, 4, 2, 200, 333, 128
    q, do = _rand((B, H, Nq, d), 1), _rand((B, H, Nq, d), 2)
    k, v = _rand((B, Hkv, Nkv, d), 3), _rand((B, Hkv, Nkv, d), 4)
    _check_fwd_bwd(q, k, v, do, "bhnd", False, torch.bfloat16)


def test_fp32_forward_token_major_tf32():
    B, N, H, d = 1, 400, 2, 64
    q, k, v = [_rand((B, N, H, d), 20 + i, torch.float32) for i in range(3)]
    _check_fwd_bwd(q, k, v, None, "bnhd", False, torch.float32, bwd=False)


def test_unsupported_strided_shapes_fail_loudly():
    q = _rand((1, 64, 2, 32), 1)                  # d = 32 has no tensor-core kernel and token-major has no generic one
    with pytest.raises(fa.B200Error):
        fa.attention_forward(q, q, q, layout="bnhd")


def test_autograd_function_matches_a_float32_torch_reference():
    B, N, H, Hkv, d = 2, 256, 4, 2, 128
    q, k, v = _rand((B, N, H, d), 1), _rand((B, N, Hkv, d), 2), _rand((B, N, Hkv, d), 3)
    w = _rand((B, N, H, d), 4, torch.float32)
    qa, ka, va = [t.clone().requires_grad_(True) for t in (q, k, v)]
    out = fa.flash_attn_func(qa, ka, va, causal=True)
    (out.float() * w).sum().backward()
    # plain fp32 torch math on the same (bf16-rounded) inputs
    qr, kr, vr = [t.float().clone().requires_grad_(True) for t in (q, k, v)]
    G = H // Hkv
    kk, vv = kr.repeat_interleave(G, dim=2), vr.repeat_interleave(G, dim=2)
    s = torch.einsum("bqhd,bkhd->bhqk", qr, kk) / d ** 0.5
    mask = torch.ones(N, N, device="cuda", dtype=torch.bool).tril()
    p = torch.softmax(s.masked_fill(~mask, float("-inf")), dim=-1)
    ref = torch.einsum("bhqk,bkhd->bqhd", p, vv)
    (ref * w).sum().backward()
    assert (out.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    for got, want in ((qa.grad, qr.grad), (ka.grad, kr.grad), (va.grad, vr.grad)):
        assert rel_l2(host(got), host(want)) <= 1e-2
        assert (got.float() - want).abs().max().item() <= 2e-2 * max(1.0, want.abs().max().item())


def test_autograd_expanded_gradient_is_not_read_as_contiguous():
    """O.sum().backward() hands the backward a gradient expanded over every dimension (all strides 0).  The all-zero
    layout means "contiguous [B,H,N,d]" in the C-ABI, so the host mirror has to materialise it."""
    B, N, H, d = 1, 192, 2, 128
    q, k, v = _rand((B, N, H, d), 11), _rand((B, N, H, d), 12), _rand((B, N, H, d), 13)
    qa, ka, va = [t.clone().requires_grad_(True) for t in (q, k, v)]
    fa.flash_attn_func(qa, ka, va).sum().backward()
    qr, kr, vr = [t.float().clone().requires_grad_(True) for t in (q, k, v)]
    s = torch.einsum("bqhd,bkhd->bhqk", qr, kr) / d ** 0.5
    torch.einsum("bhqk,bkhd->bqhd", torch.softmax(s, dim=-1), vr).sum().backward()
    for got, want in ((qa.grad, qr.grad), (ka.grad, kr.grad), (va.grad, vr.grad)):
        assert (got.float() - want).abs().max().item() <= 2e-2
    # a descriptor the tensor maps cannot express is UNSUPPORTED, not a CUDA error (row stride not a multiple of 16 bytes)
    big = _rand((1, 2, 64, 132), 14)
    with pytest.raises(fa.B200Error) as ei:
        fa.attention_forward(big[..., 2:130], big[..., 2:130], big[..., 2:130], layout="bhnd")
    assert ei.value.status == -8


def test_registered_op_traces_under_torch_compile():
    """flash_attn_func is a torch.library custom op (b200attn::fwd / b200attn::bwd): it has to pass the operator checks and
    trace without graph breaks (fullgraph=True), forward and backward, with the same numbers as the eager call."""
    B, N, H, d = 1, 256, 2, 128
    q, k, v = _rand((B, N, H, d), 21), _rand((B, N, H, d), 22), _rand((B, N, H, d), 23)
    torch.library.opcheck(torch.ops.b200attn.fwd.default, (q, k, v, True, 0.0, "bnhd"), test_utils=("test_schema", "test_faketensor"))

    def f(q_, k_, v_):
        return (fa.flash_attn_func(q_ * 1.0, k_, v_, causal=True).float() ** 2).sum()

    qa, ka, va = [t.clone().requires_grad_(True) for t in (q, k, v)]
    f(qa, ka, va).backward()
    qc, kc, vc = [t.clone().requires_grad_(True) for t in (q, k, v)]
    torch.compile(f, backend="aot_eager", fullgraph=True)(qc, kc, vc).backward()
    for a, b in ((qa.grad, qc.grad), (ka.grad, kc.grad), (va.grad, vc.grad)):
        assert torch.equal(a, b)


@pytest.mark.parametrize("causal,Hkv", [(False, 2), (True, 4), (False, 1)])
def test_packed_variable_length_batch(causal, Hkv):
    # three sequences of different lengths in one packed [total, H, d] batch (cross lengths when not causal)
    H, d = 4, 128
    lens_q = [200, 64, 333]
    lens_k = lens_q if causal else [260, 64, 130]
    cu_q, cu_k = np.concatenate([[0], np.cumsum(lens_q)]), np.concatenate([[0], np.cumsum(lens_k)])
    q, do = _rand((cu_q[-1], H, d), 1), _rand((cu_q[-1], H, d), 2)
    k, v = _rand((cu_k[-1], Hkv, d), 3), _rand((cu_k[-1], Hkv, d), 4)
    O, L = fa.attention_varlen_forward(q, k, v, cu_q, cu_k, causal=causal)
    dQ, dK, dV = fa.attention_varlen_backward(q, k, v, O, L, do, cu_q, cu_k, causal=causal)
    G = H // Hkv
    for i in range(len(lens_q)):
        sq, sk = slice(cu_q[i], cu_q[i + 1]), slice(cu_k[i], cu_k[i + 1])
        for hk in range(Hkv):
            dK_ref = dV_ref = None
            for g in range(G):
                h = hk * G + g
                args = (host(q[sq, h]), host(k[sk, hk]), host(v[sk, hk]))
                Oo, Lo = oracle.forward(*args, 0.0, causal=causal)
                scl = max(1.0, float(np.abs(Oo).max()))
                assert np.abs(host(O[sq, h]) - Oo).max() <= 2e-2 * scl and np.abs(host(L[h, sq]) - Lo).max() <= 2e-2
                rq, rk, rv = oracle.backward(*args, host(do[sq, h]), 0.0, causal=causal)
                assert rel_l2(host(dQ[sq, h]), rq) <= 1e-2
                dK_ref = rk if dK_ref is None else dK_ref + rk
                dV_ref = rv if dV_ref is None else dV_ref + rv
            assert rel_l2(host(dK[sk, hk]), dK_ref) <= 1e-2 and rel_l2(host(dV[sk, hk]), dV_ref) <= 1e-2
            assert np.abs(host(dV[sk, hk]) - dV_ref).max() <= 2e-2 * max(1.0, float(np.abs(dV_ref).max()))


def test_packed_variable_length_batch_more_sequences_than_side_streams():
    # r2h: the sequences of a packed batch run on 4 internal streams (round-robin), each with its own workspace slot; nine
    # sequences (one of them empty) make every stream and slot serve two or three of them, in both directions
    H, Hkv, d = 4, 2, 128
    lens = [130, 64, 0, 257, 96, 333, 64, 200, 150]
    cu = np.concatenate([[0], np.cumsum(lens)])
    q, do = _rand((cu[-1], H, d), 11), _rand((cu[-1], H, d), 12)
    k, v = _rand((cu[-1], Hkv, d), 13), _rand((cu[-1], Hkv, d), 14)
    O, L = fa.attention_varlen_forward(q, k, v, cu, cu, causal=True)
    dQ, dK, dV = fa.attention_varlen_backward(q, k, v, O, L, do, cu, cu, causal=True)
    torch.cuda.synchronize()
    G = H // Hkv
    for i in range(len(lens)):
        if lens[i] == 0:
            continue
        s = slice(cu[i], cu[i + 1])
        for hk in range(Hkv):
            dK_ref = dV_ref = None
            for g in range(G):
                h = hk * G + g
                args = (host(q[s, h]), host(k[s, hk]), host(v[s, hk]))
                Oo, Lo = oracle.forward(*args, 0.0, causal=True)
                assert np.abs(host(O[s, h]) - Oo).max() <= 2e-2 * max(1.0, float(np.abs(Oo).max())) and np.abs(host(L[h, s]) - Lo).max() <= 2e-2
                rq, rk, rv = oracle.backward(*args, host(do[s, h]), 0.0, causal=True)
                assert rel_l2(host(dQ[s, h]), rq) <= 1e-2, (i, h)
                dK_ref = rk if dK_ref is None else dK_ref + rk
                dV_ref = rv if dV_ref is None else dV_ref + rv
            assert rel_l2(host(dK[s, hk]), dK_ref) <= 1e-2 and rel_l2(host(dV[s, hk]), dV_ref) <= 1e-2, (i, hk)
