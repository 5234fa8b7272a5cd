"""GPU parity against the reference's OWN CUDA kernels (north_star: "... and its CUDA FA2/ring kernels").

oracle/_ref/libref_cuda_fwd.so / libref_cuda_bwd.so are the reference's unmodified flash_attention_2_forward
(src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:133-172) and flash_attention_2_backward
(src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:249-299), compiled for sm_100a by
`make -C oracle refcuda` in the build container (they travel to the GPU box prebuilt).  Both implementations run on the
reference's own fixtures G3 / G4 and on a ragged random case; ours must agree with theirs within the tf32 bound (2e-3),
and both must agree with the golden vectors the reference CPU code produced (its own thresholds: 1e-3 fwd, 5e-3 bwd)."""
import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa
from gpu_util import dev, host

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not oracle.ref_cuda_available(), reason="oracle/_ref/libref_cuda_*.so not built")]


def _ref_forward(Q, K, V, sc):
    O, L = torch.empty_like(Q), torch.empty(Q.shape[0], device="cuda")
    torch.cuda.synchronize()
    assert oracle.ref_cuda_forward(Q, K, V, O, L, sc) == 0
    torch.cuda.synchronize()
    return O, L


def test_forward_G3_ours_vs_reference_cuda(golden):
    g = golden("G3")
    sc = float(g["scale"])
    Q, K, V = dev(g["Q"]), dev(g["K"]), dev(g["V"])
    Or, Lr = _ref_forward(Q, K, V, sc)
    O, L = fa.flash_attention_2_forward(Q, K, V, sc)
    assert np.abs(host(Or) - g["O"]).max() < 1e-3 and np.abs(host(Lr) - g["L"]).max() < 1e-3      # the reference kernel itself
    assert np.abs(host(O) - g["O"]).max() < 1e-3 and np.abs(host(L) - g["L"]).max() < 1e-3        # 02_*_forward/main.cu threshold
    assert (O - Or).abs().max().item() <= 2e-3 and (L - Lr).abs().max().item() <= 2e-3


@pytest.mark.parametrize("N,d", [(1000, 64), (2048, 64), (300, 32)])
def test_forward_random_ours_vs_reference_cuda(N, d):
    rng = np.random.default_rng(N + d)
    Q, K, V = [dev(rng.standard_normal((N, d)).astype(np.float32)) for _ in range(3)]
    sc = 1.0 / d ** 0.5
    Or, Lr = _ref_forward(Q, K, V, sc)
    O, L = fa.flash_attention_2_forward(Q, K, V, sc)
    print(N, d, (O - Or).abs().max().item(), (L - Lr).abs().max().item())
    assert (O - Or).abs().max().item() <= 2e-3 and (L - Lr).abs().max().item() <= 2e-3


def test_backward_G4_ours_vs_reference_cuda(golden):
    g = golden("G4")
    sc = float(g["scale"])
    t = {k: dev(g[k]) for k in ("Q", "K", "V", "O", "L", "dO")}
    ref = [torch.empty_like(t["Q"]) for _ in range(3)]
    torch.cuda.synchronize()
    assert oracle.ref_cuda_backward(t["Q"], t["K"], t["V"], t["O"], t["L"], t["dO"], *ref, sc) == 0
    torch.cuda.synchronize()
    ours = fa.flash_attention_2_backward(t["Q"], t["K"], t["V"], t["O"], t["L"], t["dO"], sc)
    for name, a, b in zip(("dQ", "dK", "dV"), ours, ref):
        print(name, (a - b).abs().max().item(), np.abs(host(b) - g[name]).max())
        assert np.abs(host(b) - g[name]).max() < 5e-3            # the reference kernel vs its CPU code (main.cu:253-258)
        assert np.abs(host(a) - g[name]).max() < 5e-3
        assert (a - b).abs().max().item() <= 2e-3


# ---- the reference's CUDA RING kernel (oracle/_ref/libref_cuda_ring.so) ---------------------------------------------
# ring_attention_forward_kernel (src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:13-140) with its (m, l, O)
# carry, launched for the P ranks of a ring one after the other with the reference's own schedule; our side is the same
# emulation with the library's per-block forward + LSE merge (what b200_ring_fwd enqueues per step, tests/test_gpu_ring.py).
def _ours_ring_emulated(q, k, v, P, scale):
    n = q.shape[0] // P
    outs, lses = [], []
    for r in range(P):
        qs = q[r * n:(r + 1) * n].contiguous()
        O = L = None
        for step in range(P):
            b = fa.ring_kv_block(r, step, P)
            Ob, Lb = fa.flash_attention_2_forward(qs, k[b * n:(b + 1) * n].contiguous(), v[b * n:(b + 1) * n].contiguous(), scale)
            if O is None:
                O, L = Ob, Lb
            else:
                fa.lse_merge_(O, L, Ob, Lb)
        outs.append(O)
        lses.append(L)
    return torch.cat(outs), torch.cat(lses)


ring_only = pytest.mark.skipif(not oracle.ref_cuda_ring_available(), reason="oracle/_ref/libref_cuda_ring.so not built")


@ring_only
@pytest.mark.parametrize("P", [1, 2, 4, 8])
def test_ring_G5_ours_vs_reference_cuda_ring_kernel(P, golden):
    from conftest import simple_test_data
    g = golden("G5")
    Q, K, V = simple_test_data(5096, 64)                   # the reference ring fixture (04_ring_attention.cu:19-21), scale 1
    q, k, v = dev(Q), dev(K), dev(V)
    Or, Lr = oracle.ref_cuda_ring_forward(q, k, v, P, 1.0)
    O, L = _ours_ring_emulated(q, k, v, P, 1.0)
    mag = float(np.abs(g["O"]).max())                      # ~1e4: judged relative, like the reference's own rule
    assert oracle.compare_outputs(g["O"], host(Or), 5e-3, 1.0) == 0          # the reference kernel itself (attention_helper.h:184)
    assert oracle.compare_outputs(g["O"], host(O), 5e-3, 1.0) == 0
    print(P, (O - Or).abs().max().item() / mag, (L - Lr).abs().max().item())
    assert (O - Or).abs().max().item() <= 2e-3 * mag and (L - Lr).abs().max().item() <= 2e-3


@ring_only
@pytest.mark.parametrize("N,P", [(2048, 4), (1536, 3)])
def test_ring_random_ours_vs_reference_cuda_ring_kernel(N, P):
    rng = np.random.default_rng(N + P)
    Q, K, V = [rng.standard_normal((N, 64)).astype(np.float32) for _ in range(3)]
    q, k, v = dev(Q), dev(K), dev(V)
    sc = 0.125
    Or, Lr = oracle.ref_cuda_ring_forward(q, k, v, P, sc)
    O, L = _ours_ring_emulated(q, k, v, P, sc)
    Oo, Lo = oracle.forward(Q, K, V, sc)
    assert np.abs(host(Or) - Oo).max() <= 1e-4 and np.abs(host(Lr) - Lo).max() <= 1e-4      # their fp32 kernel vs the CPU oracle
    print(N, P, (O - Or).abs().max().item(), (L - Lr).abs().max().item())
    assert (O - Or).abs().max().item() <= 2e-3 and (L - Lr).abs().max().item() <= 2e-3
