"""GPU parity of the fp8 (e4m3) forward (SURVEY 8f-4 stretch: kind::f8f6f4 MMAs; fp8 Q/K/V in, bf16 O out, d = 128) against
the oracle evaluated on the DEQUANTISED inputs -- so the only errors are the kernel's own: P rounded to e4m3 (3 mantissa
bits, relative 2^-4 per probability, averaging down over the keys of a row) and the bf16 store.  Stated tolerance:
max-abs <= 5e-2 and rel-L2 <= 3e-2 on O for unit-normal data, LSE <= 1e-3 (the scores and the row sums are fp32)."""
import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa
from cuda_flashattention_b200._lib import lib
from gpu_util import host, rel_l2

pytestmark = pytest.mark.gpu


def _fp8(shape, rng, s=1.0):
    return torch.from_numpy((s * rng.standard_normal(shape)).astype(np.float32)).cuda().to(torch.float8_e4m3fn)


@pytest.mark.parametrize("Nq,Nkv,B,H,causal", [(128, 128, 1, 1, False), (256, 256, 1, 2, False), (1000, 1000, 1, 2, False),
                                               (2048, 2048, 1, 1, False), (333, 700, 2, 2, False), (1024, 1024, 1, 2, True),
                                               (200, 500, 1, 1, True)])
def test_fp8_forward_vs_oracle(Nq, Nkv, B, H, causal):
    d = 128
    rng = np.random.default_rng(Nq + 3 * Nkv)
    q, k, v = _fp8((B, H, Nq, d), rng), _fp8((B, H, Nkv, d), rng), _fp8((B, H, Nkv, d), rng)
    assert lib.b200_fwd_uses_tensor_cores(2, d, Nq, Nkv) == 1
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0, causal=causal)
    assert O.dtype == torch.bfloat16
    worst = (0.0, 0.0, 0.0)
    for b in range(B):
        for h in range(H):
            Oo, Lo = oracle.forward(host(q[b, h].float()), host(k[b, h].float()), host(v[b, h].float()), 0.0, causal=causal)
            g = host(O[b, h])
            worst = max(worst, (float(np.abs(g - Oo).max()), rel_l2(g, Oo), float(np.abs(host(L[b, h]) - Lo).max())))
    print(Nq, Nkv, causal, worst)
    assert worst[0] <= 5e-2 and worst[1] <= 3e-2 and worst[2] <= 1e-3


def test_fp8_is_forward_only_and_d128_only():
    rng = np.random.default_rng(0)
    q = _fp8((1, 1, 128, 64), rng)
    with pytest.raises(fa.B200Error) as ei:
        fa.flash_attention_2_forward(q, q, q, 0.0)
    assert ei.value.status == -8                                   # B200_ERR_UNSUPPORTED: no generic fp8 path
    q = _fp8((1, 1, 128, 128), rng)
    O, L = fa.flash_attention_2_forward(q, q, q, 0.0)
    P = lambda t: t.data_ptr()
    st = lib.b200_fa2_bwd(P(q), P(q), P(q), P(O), P(L), P(O), P(O), P(O), P(O), 1, 1, 128, 128, 128, 0.0, 2, None, 0, None)
    assert st == -8                                                # the fp8 dtype is forward-only
