"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/b200attn.h declares, and fails loudly (never falls back) when there is no CUDA device."""
import ctypes as C
import os
import re

import pytest
import torch

import cuda_flashattention_b200 as fa
from conftest import ROOT


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "b200attn.h")).read()
    return sorted(set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", hdr)))


def test_header_symbols_exported():
    names = _declared_symbols()
    assert len(names) >= 20
    raw = C.CDLL(fa.LIB_PATH)
    for n in names:
        assert hasattr(raw, n), f"{n} declared in include/b200attn.h but not exported"
    assert set(names) == set(fa.EXPORTS), "python binding table out of sync with the header"


def test_version_and_strerror():
    assert fa.lib.b200_version() >= 100
    assert b"no CPU fallback" in fa.lib.b200_strerror(-4)
    assert fa.lib.b200_strerror(0) == b"ok"


def test_ring_schedule_is_reference_schedule():
    # kv_block_idx = (rank - step + nranks) % nranks, common/ring_attention_kernel.cu:198
    for P in (1, 2, 3, 4, 8):
        for r in range(P):
            seen = [fa.ring_kv_block(r, s, P) for s in range(P)]
            assert seen == [(r - s + P) % P for s in range(P)]
            assert sorted(seen) == list(range(P))          # every block visited exactly once


def test_invalid_arguments_rejected_without_device():
    assert fa.lib.b200_fa2_fwd_f32(None, None, None, None, None, -1, 4, 1.0, None) == -1
    assert fa.lib.b200_fa2_fwd(None, None, None, None, None, 1, 1, 4, 4, 4, 1.0, 7, None) == -1      # bad dtype
    assert fa.lib.b200_fa2_fwd_f32(None, None, None, None, None, 0, 4, 1.0, None) == 0                 # empty: no-op
    assert fa.lib.b200_fa2_bwd_workspace_bytes(1, 1, 128, 128, 64, 0) >= 128 * 4


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    # dummy non-null pointers: the call must stop at the device check, before touching them
    st = fa.lib.b200_fa2_fwd_f32(8, 8, 8, 8, None, 4, 4, 1.0, None)
    assert st in (-4, -3), st
    q = torch.zeros(4, 4)
    with pytest.raises(ValueError):
        fa.flash_attention_2_forward(q, q, q)


def test_zigzag_causal_ring_schedule_covers_the_causal_triangle_evenly():
    # host logic of b200_ring_fwd_ex(B200_FLAG_CAUSAL): step 0 = local causal; a block from rank b < r = all local
    # queries x first half of the block; from b > r = second half of the local queries x the whole block.
    # Check on chunk granularity that the rectangles are exactly the visible pairs, once each, same area every step.
    for P in (1, 2, 3, 4, 8):
        for r in range(P):
            mine = [fa.lib.b200_ring_zigzag_chunk(r, 0, P), fa.lib.b200_ring_zigzag_chunk(r, 1, P)]
            assert mine == [r, 2 * P - 1 - r]
            seen = set()
            for s in range(P):
                b = fa.ring_kv_block(r, s, P)
                theirs = [b, 2 * P - 1 - b]
                if s == 0:
                    pairs = {(mine[0], mine[0], "diag"), (mine[1], mine[0], "full"), (mine[1], mine[1], "diag")}
                elif b < r:
                    pairs = {(mine[0], theirs[0], "full"), (mine[1], theirs[0], "full")}
                else:
                    pairs = {(mine[1], theirs[0], "full"), (mine[1], theirs[1], "full")}
                area = sum(1.0 if kind == "full" else 0.5 for _, _, kind in pairs)
                assert area == 2.0                                   # every step costs the same on every rank
                for q, k, kind in pairs:
                    assert (k < q and kind == "full") or (k == q and kind == "diag")
                    assert (q, k) not in seen
                    seen.add((q, k))
            want = {(q, k) for q in mine for k in range(2 * P) if k <= q}
            assert seen == want


def test_descriptor_and_varlen_argument_checks_without_device():
    # argument validation of the extended entry points happens before any device call
    import ctypes as C
    from cuda_flashattention_b200._lib import AttnDesc
    d = AttnDesc(B=1, H=6, Hkv=4, Nq=64, Nkv=64, d=128, dtype=0, flags=0, scale=0.0)     # 6 % 4 != 0
    assert fa.lib.b200_fa2_fwd_desc(d, 8, 8, 8, 8, 8, None) == -1
    d.Hkv = 2
    d.flags = 2                                                                          # unknown flag bit
    assert fa.lib.b200_fa2_fwd_desc(d, 8, 8, 8, 8, 8, None) == -1
    d.flags, d.Nq, d.Nkv = 1, 128, 64                                                    # causal needs Nq <= Nkv
    assert fa.lib.b200_fa2_fwd_desc(d, 8, 8, 8, 8, 8, None) == -1
    assert fa.lib.b200_fa2_bwd_desc(d, *([8] * 9), None, 0, None) == -1
    d.B = 0                                                                              # empty batch: no-op
    d.flags, d.Nq = 0, 64
    assert fa.lib.b200_fa2_fwd_desc(d, None, None, None, None, None, None) == 0
    # packed batches: lengths must be non-negative; causal needs equal q / k lengths
    d = AttnDesc(B=1, H=4, Hkv=4, Nq=0, Nkv=0, d=128, dtype=0, flags=0, scale=0.0)
    bad = (C.c_int * 3)(0, 10, 5)
    ok = (C.c_int * 3)(0, 10, 30)
    ok_k = (C.c_int * 3)(0, 10, 40)
    assert fa.lib.b200_fa2_fwd_varlen(d, 2, bad, bad, 8, 8, 8, 8, 8, None) == -1
    d.flags = 1
    assert fa.lib.b200_fa2_fwd_varlen(d, 2, ok_k, ok, 8, 8, 8, 8, 8, None) == -1         # causal with a sequence of Nq > Nkv
    assert fa.lib.b200_fa2_fwd_varlen(d, 0, None, None, None, None, None, None, None, None) == 0
    # ring: unknown flags / bad ranks are rejected before anything else
    assert fa.lib.b200_ring_fwd_ex(8, 8, 8, 8, 8, 1, 1, 64, 128, 0.0, 0, 4, None, 0, 2, None) == -1
    assert fa.lib.b200_ring_zigzag_chunk(3, 0, 2) == -1


def test_route_query_reports_the_demotions():
    """b200_fa2_fwd_route / b200_fa2_bwd_route tell a caller which kernel family ONE call runs on, including the demotions
    b200_*_uses_tensor_cores cannot see: unaligned pointers, scale <= 0, fp32 backward at d = 128 (pure host functions)."""
    R = fa.lib.b200_fa2_fwd_route
    BF, F32 = 0, 1
    A = 4096                                   # a 16-byte aligned dummy address (never dereferenced)
    assert R(A, A, A, A, 8192, 8192, 128, 0.0, BF, 0) == 1
    assert R(A, A, A, A, 4096, 4096, 64, 0.0, F32, 0) == 1           # fp32 I/O: tf32 tensor cores
    assert R(A, A, A, A, 17, 17, 40, 0.0, BF, 0) == 0                 # head dim
    assert R(A + 2, A, A, A, 8192, 8192, 128, 0.0, BF, 0) == 0        # unaligned Q
    assert R(A, A, A, A, 8192, 8192, 128, -1.0, BF, 0) == 0           # scale <= 0
    assert R(A, A, A, A, 1024, 1024, 128, 0.0, BF, 1) == 1            # causal, square
    assert R(A, A, A, A, 512, 1024, 128, 0.0, BF, 1) == 1             # causal, cross lengths (tensor cores since r2e)
    assert R(A, A, A, A, 1024, 512, 128, 0.0, BF, 1) == -1            # causal needs Nq <= Nkv
    Rb = fa.lib.b200_fa2_bwd_route
    assert Rb(A, A, A, A, A, A, A, 8192, 8192, 128, 0.0, BF, 0) == 1
    assert Rb(A, A, A, A, A, A, A + 4, 8192, 8192, 128, 0.0, BF, 0) == 0
    assert Rb(A, A, A, A, A, A, A, 512, 1024, 128, 0.0, BF, 1) == 1
    assert Rb(A, A, A, A, A, A, A, 512, 1024, 64, 0.0, F32, 1) == 1         # fp32 causal backward, d = 64: tf32 tensor cores (r2h)
    assert Rb(A, A, A, A, A, A, A, 512, 1024, 128, 0.0, F32, 1) == 0        # fp32 d = 128 backward: generic
    assert Rb(A, A, A, A, A, A, A, 128, 128, 64, 0.0, F32, 0) == fa.lib.b200_bwd_uses_tensor_cores(F32, 64, 128, 128)
