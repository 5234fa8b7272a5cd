"""torchrun worker for test_gpu_ring.py::test_ring_nccl_multi_gpu: real ring fwd + bwd over NCCL."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle  # noqa: E402
import cuda_flashattention_b200 as fa  # noqa: E402
from cuda_flashattention_b200._lib import lib, check  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    comm = fa.RingComm()
    ok = True
    # every data plane / forward schedule the library has (include/b200attn.h: b200_ring_set_transport,
    # b200_ring_set_forward_mode): copy engines + persistent forward (the default), copy engines + one launch per step,
    # NCCL send/recv (the reference's ring_exchange_kv; always stepwise)
    for (transport, fwd_mode, bwd_mode, label) in ((0, 2, 0, "copy-engine / persistent forward / home backward"),
                                                   (0, 1, 1, "copy-engine / stepwise forward / travelling-accumulator backward"),
                                                   (1, 1, 0, "nccl / stepwise forward / travelling-accumulator backward")):
        check(lib.b200_ring_set_transport(transport), "set_transport")
        check(lib.b200_ring_set_forward_mode(fwd_mode), "set_forward_mode")
        check(lib.b200_ring_set_backward_mode(bwd_mode), "set_backward_mode")
        if rank == 0:
            print("=== ring configuration:", label)
        ok = battery(rank, world, comm) and ok
    check(lib.b200_ring_set_transport(-1), "set_transport")
    check(lib.b200_ring_set_forward_mode(0), "set_forward_mode")
    check(lib.b200_ring_set_backward_mode(0), "set_backward_mode")
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    comm.close()
    dist.destroy_process_group()
    if rank == 0:
        print("RING_OK" if flag.item() == 1 else "RING_FAIL")
    sys.exit(0 if flag.item() == 1 else 1)


def battery(rank, world, comm):
    ok = True
    for (dtype, H, N, d, tol) in ((torch.bfloat16, 2, 256 * world, 128, 2e-2), (torch.bfloat16, 3, 448 * world, 128, 2e-2),
                                  (torch.float32, 1, 5096 // world * world, 64, 2e-3)):
        rng = np.random.default_rng(7)
        Q, K, V, dO = [rng.standard_normal((1, H, N, d)).astype(np.float32) for _ in range(4)]
        n = N // world
        sl = slice(rank * n, (rank + 1) * n)
        to = lambda a: torch.from_numpy(a[:, :, sl].copy()).cuda().to(dtype).contiguous()
        q, k, v, do = to(Q), to(K), to(V), to(dO)
        k0, v0 = k.clone(), v.clone()
        for _rep in range(2):                          # twice back to back: the second call runs into the first one's scratch
            O, L = fa.ring_attention_forward(q, k, v, comm, 0.0)
            dQ, dK, dV = fa.ring_attention_backward(q, k, v, O, L, do, comm, 0.0)
        torch.cuda.synchronize()
        assert torch.equal(k, k0) and torch.equal(v, v0), "K_local/V_local must not be modified"
        # gather the (rounded) full inputs so the oracle sees exactly what the GPUs saw
        def gather(t):
            lst = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(lst, t)
            return torch.cat(lst, dim=2).float().cpu().numpy()
        Qf, Kf, Vf, dOf = gather(q), gather(k), gather(v), gather(do)
        for h in range(H):
            Oo, Lo = oracle.forward(Qf[0, h], Kf[0, h], Vf[0, h], 0.0)
            gq, gk, gv = oracle.backward(Qf[0, h], Kf[0, h], Vf[0, h], dOf[0, h], 0.0)
            errs = {"O": np.abs(O[0, h].float().cpu().numpy() - Oo[sl]).max(), "L": np.abs(L[0, h].cpu().numpy() - Lo[sl]).max(),
                    "dQ": np.abs(dQ[0, h].float().cpu().numpy() - gq[sl]).max(), "dK": np.abs(dK[0, h].float().cpu().numpy() - gk[sl]).max(),
                    "dV": np.abs(dV[0, h].float().cpu().numpy() - gv[sl]).max()}
            if rank == 0:
                print(dtype, "head", h, {k_: float(v_) for k_, v_ in errs.items()})
            ok = ok and all(e <= tol for e in errs.values())
    # ---- causal ring forward, zigzag sharding (extension): every rank holds chunk r and chunk 2P-1-r
    for (dtype, H, N, d, tol) in ((torch.bfloat16, 2, 256 * world, 128, 2e-2), (torch.float32, 1, 384 * world, 64, 2e-3)):
        rng = np.random.default_rng(11)
        Q, K, V, dO = [rng.standard_normal((1, H, N, d)).astype(np.float32) for _ in range(4)]
        rows = fa.zigzag_rows(rank, world, N)
        to = lambda a: torch.from_numpy(a[:, :, rows].copy()).cuda().to(dtype).contiguous()
        q, k, v, do = to(Q), to(K), to(V), to(dO)
        O, L = fa.ring_attention_forward(q, k, v, comm, 0.0, causal=True)
        grads = fa.ring_attention_backward(q, k, v, O, L, do, comm, 0.0, causal=True) if dtype == torch.bfloat16 else None
        torch.cuda.synchronize()
        rt = lambda a: torch.from_numpy(a).cuda().to(dtype).float().cpu().numpy()      # what the GPUs saw
        for h in range(H):
            Oo, Lo = oracle.forward(rt(Q[0, h]), rt(K[0, h]), rt(V[0, h]), 0.0, causal=True)
            scl = max(1.0, float(np.abs(Oo).max()))
            errs = {"O": np.abs(O[0, h].float().cpu().numpy() - Oo[rows]).max() / scl, "L": np.abs(L[0, h].cpu().numpy() - Lo[rows]).max()}
            if grads is not None:
                ref = oracle.backward(rt(Q[0, h]), rt(K[0, h]), rt(V[0, h]), rt(dO[0, h]), 0.0, causal=True)
                for name, got, want in zip(("dQ", "dK", "dV"), grads, ref):
                    errs[name] = np.abs(got[0, h].float().cpu().numpy() - want[rows]).max() / max(1.0, float(np.abs(want).max()))
            if rank == 0:
                print("causal zigzag", dtype, "head", h, {k_: float(v_) for k_, v_ in errs.items()})
            ok = ok and all(e <= tol for e in errs.values())
    return ok


if __name__ == "__main__":
    main()
