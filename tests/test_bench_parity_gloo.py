"""CPU test (world_size 2, gloo) of bench.py's parity gate -- the code that decides whether a contract line at N > 1 is
trustworthy: sharded "device outputs" that ARE the oracle's results (rounded to bf16 like a kernel's) must pass, outputs
with one K/V tile's gradient dropped or a shifted LSE on one rank must fail.  Also checks the reference arm's honest
config labelling."""
import importlib.util
import json
import os
import socket
import subprocess
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
import cuda_flashattention_b200 as fa
from conftest import ROOT


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    bench = _bench()
    B, H, Nt, d = 1, 2, 512, 32
    Nl = Nt // world
    rng = np.random.default_rng(3)
    bf = lambda a: torch.from_numpy(a.astype(np.float32)).bfloat16()
    Q, K, V = [bf(rng.standard_normal((B, H, Nt, d))) for _ in range(3)]
    dO = bf(0.2 * rng.standard_normal((B, H, Nt, d)))
    O = torch.empty(B, H, Nt, d); L = torch.empty(B, H, Nt)
    dQ, dK, dV = torch.empty_like(O), torch.empty_like(O), torch.empty_like(O)
    for h in range(H):
        f = lambda t: t[0, h].float().numpy()
        o, l = oracle.forward(f(Q), f(K), f(V), 0.0)
        gq, gk, gv = oracle.backward(f(Q), f(K), f(V), f(dO), 0.0)
        O[0, h], L[0, h] = torch.from_numpy(o), torch.from_numpy(l)
        dQ[0, h], dK[0, h], dV[0, h] = torch.from_numpy(gq), torch.from_numpy(gk), torch.from_numpy(gv)
    sl = slice(rank * Nl, (rank + 1) * Nl)
    shard = lambda t: t[:, :, sl].contiguous()
    good = [shard(Q), shard(K), shard(V), shard(dO), shard(O.bfloat16()), shard(L), shard(dQ.bfloat16()), shard(dK.bfloat16()),
            shard(dV.bfloat16())]
    res = {}
    res["good"] = bench.parity_check(torch, dist, fa, world, rank, False, (B, H, Nl, Nt, d), good)
    # (a) rank 1 lost the contribution of its first 64 query rows to one 128-row K/V tile of dK (a dropped tile update)
    bad = [t.clone() for t in good]
    if rank == 1:
        bad[7][:, :, :128] *= 0.5
    res["dK_tile"] = bench.parity_check(torch, dist, fa, world, rank, False, (B, H, Nl, Nt, d), bad)
    # (b) rank 0 reports an LSE that is off by 0.05 on every row
    bad = [t.clone() for t in good]
    if rank == 0:
        bad[5] += 0.05
    res["lse"] = bench.parity_check(torch, dist, fa, world, rank, False, (B, H, Nl, Nt, d), bad)
    if rank == 0:
        out["res"] = res
    dist.destroy_process_group()


def test_parity_gate_world2_gloo():
    mgr = mp.Manager(); out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    r = out["res"]
    assert r["good"]["ok"] is True, r["good"]
    assert r["good"]["stats"] == "oracle" and max(r["good"]["max_abs"].values()) <= 2e-2
    assert all(v <= 1.0 for v in r["good"]["identities_err_over_bound"].values()), r["good"]
    assert r["dK_tile"]["ok"] is False, r["dK_tile"]
    assert r["lse"]["ok"] is False, r["lse"]


def test_reference_arm_says_what_it_sampled():
    env = dict(os.environ, OMP_NUM_THREADS="2")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    j = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][0])
    assert j["config"]["reference_sample_differs_from_gpu_config"] is True
    assert j["config"]["sampled"]["N"] == 1024 and j["config"]["sampled"]["dtype"] == "f32"
