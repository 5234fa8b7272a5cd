"""Host-buffer entry points (b200_fa2_fwd_host / b200_fa2_fwd_bwd_host): pipelined H2D / attention / D2H must give
exactly what the device-resident calls give, and match the oracle."""
import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,H,N,d,dtype", [(2, 3, 384, 128, torch.bfloat16), (1, 2, 200, 64, torch.float32),
                                           (3, 40, 2048, 128, torch.bfloat16)])   # last one spans several chunks
def test_host_pipeline_matches_device_calls(B, H, N, d, dtype):
    g = torch.Generator().manual_seed(B * 1000 + N)
    q, k, v, do = [torch.randn(B, H, N, d, generator=g).to(dtype).pin_memory() for _ in range(4)]
    O, L, dQ, dK, dV = fa.flash_attention_2_forward_backward_host(q, k, v, do, 0.0)
    Of, Lf = fa.flash_attention_2_forward_host(q, k, v, 0.0)
    torch.cuda.synchronize()
    qd, kd, vd, dod = q.cuda(), k.cuda(), v.cuda(), do.cuda()
    Od, Ld = fa.flash_attention_2_forward(qd, kd, vd, 0.0)
    gq, gk, gv = fa.flash_attention_2_backward(qd, kd, vd, Od, Ld, dod, 0.0)
    torch.cuda.synchronize()
    assert torch.equal(O, Od.cpu()) and torch.equal(L, Ld.cpu()) and torch.equal(Of, Od.cpu()) and torch.equal(Lf, Ld.cpu())
    assert torch.equal(dK, gk.cpu()) and torch.equal(dV, gv.cpu())
    tol = 2e-2 if dtype == torch.bfloat16 else 2e-3
    assert (dQ.float() - gq.cpu().float()).abs().max() <= tol          # dQ is reduced with fp32 atomics: order varies
    # one head against the oracle
    h = lambda t: t[0, 0].float().numpy()
    Oo, Lo = oracle.forward(h(q), h(k), h(v), 0.0)
    oq, ok_, ov = oracle.backward(h(q), h(k), h(v), h(do), 0.0)
    for got, want in ((O, Oo), (dQ, oq), (dK, ok_), (dV, ov)):
        assert np.abs(h(got) - want).max() <= tol
    assert np.abs(L[0, 0].numpy() - Lo).max() <= tol


def test_host_pipeline_rejects_device_tensors():
    q = torch.zeros(1, 1, 128, 128, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(ValueError):
        fa.flash_attention_2_forward_host(q, q, q)
