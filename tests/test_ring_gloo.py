"""CPU test of the N>1 ring path's host logic with world_size 2 over gloo: the block schedule comes
from the library (b200_ring_kv_block, the reference's (rank - step) mod P), K/V shards rotate
send(next)/recv(prev) exactly like ring_exchange_kv, the per-block attention and the LSE merge are
the oracle's.  Every rank must end with its rows of the full attention (forward) and, rotating
(K, V, dK, dV), with its rows of dQ/dK/dV (backward)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
import cuda_flashattention_b200 as fa


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _exchange(t_send, nxt, prv):
    recv = torch.empty_like(t_send)
    ops = [dist.P2POp(dist.isend, t_send, nxt), dist.P2POp(dist.irecv, recv, prv)]
    for r in dist.batch_isend_irecv(ops):
        r.wait()
    return recv


def _worker(rank, world, port, N, d, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)                       # same full tensors on every rank (like 04_ring_attention.cu:66-84)
    Q, K, V, dO = [rng.standard_normal((N, d)).astype(np.float32) for _ in range(4)]
    n = N // world
    sl = slice(rank * n, (rank + 1) * n)
    q, do = Q[sl], dO[sl]
    nxt, prv = (rank + 1) % world, (rank - 1) % world
    # ---- forward
    kc, vc = torch.from_numpy(K[sl].copy()), torch.from_numpy(V[sl].copy())
    O = L = None
    for step in range(world):
        blk = fa.ring_kv_block(rank, step, world)
        assert np.array_equal(kc.numpy(), K[blk * n:(blk + 1) * n])          # the shard we hold IS block (rank-step) mod P
        Ob, Lb = oracle.forward(q, kc.numpy(), vc.numpy(), 0.25)
        if O is None:
            O, L = Ob, Lb
        else:
            oracle.lse_merge(O, L, Ob, Lb)
        if step < world - 1:
            kc, vc = _exchange(kc, nxt, prv), _exchange(vc, nxt, prv)
    # ---- backward: (K, V, dK, dV) rotate, P hops bring dK/dV home
    kc, vc = torch.from_numpy(K[sl].copy()), torch.from_numpy(V[sl].copy())
    dQ = np.zeros((n, d), np.float32)
    dKa, dVa = torch.zeros(n, d), torch.zeros(n, d)
    for step in range(world):
        # partial grads of local queries w.r.t. the resident block, using the GLOBAL softmax statistics L
        S = (q @ kc.numpy().T) * 0.25
        Pm = np.exp(S - L[:, None])
        dP = do @ vc.numpy().T
        Dv = (do * O).sum(1, keepdims=True)
        dS = Pm * (dP - Dv) * 0.25
        dQ += dS @ kc.numpy()
        dKa += torch.from_numpy((dS.T @ q).astype(np.float32)); dVa += torch.from_numpy((Pm.T @ do).astype(np.float32))
        dKa, dVa = _exchange(dKa, nxt, prv), _exchange(dVa, nxt, prv)
        if step < world - 1:
            kc, vc = _exchange(kc, nxt, prv), _exchange(vc, nxt, prv)
    out[rank] = (O, L, dQ, dKa.numpy(), dVa.numpy())
    dist.destroy_process_group()


def test_ring_schedule_world2_gloo():
    world, N, d = 2, 64, 16
    mgr = mp.Manager(); out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), N, d, out), nprocs=world, join=True)
    rng = np.random.default_rng(0)
    Q, K, V, dO = [rng.standard_normal((N, d)).astype(np.float32) for _ in range(4)]
    Of, Lf = oracle.forward(Q, K, V, 0.25)
    gQ, gK, gV = oracle.backward(Q, K, V, dO, 0.25)
    n = N // world
    for r in range(world):
        O, L, dQ, dK, dV = out[r]
        sl = slice(r * n, (r + 1) * n)
        assert np.abs(O - Of[sl]).max() < 1e-5 and np.abs(L - Lf[sl]).max() < 1e-5
        assert np.abs(dQ - gQ[sl]).max() < 1e-4 and np.abs(dK - gK[sl]).max() < 1e-4 and np.abs(dV - gV[sl]).max() < 1e-4
