"""GPU tests of the ring path.  (1) one GPU: the ranks of a P-way ring are emulated one after the
other with the library's own per-block forward + LSE merge (what b200_ring_fwd enqueues per step);
(2) >= 2 GPUs: the real thing -- one process per GPU, NCCL K/V rotation -- against the oracle."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest
import torch

import oracle
import cuda_flashattention_b200 as fa
from conftest import ROOT, simple_test_data
from gpu_util import dev, host

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("P,dtype,N,d", [(2, torch.float32, 5096, 64), (4, torch.float32, 5096, 64), (8, torch.float32, 5096, 64),
                                          (4, torch.bfloat16, 2048, 128)])
def test_ring_emulated_on_one_gpu(P, dtype, N, d, golden):
    if dtype == torch.float32:
        Q, K, V = simple_test_data(N, d)                  # the reference ring fixture (04_ring_attention.cu:19-21)
        scale, want = 1.0, golden("G5")["O"]
    else:
        rng = np.random.default_rng(P)
        Q, K, V = [rng.standard_normal((N, d)).astype(np.float32) for _ in range(3)]
        scale = 0.0
    q, k, v = dev(Q, dtype), dev(K, dtype), dev(V, dtype)
    if dtype != torch.float32:
        want, _ = oracle.forward(host(q), host(k), host(v), scale)
    n = N // P
    outs = []
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
        outs.append(host(O))
    got = np.concatenate(outs)
    if dtype == torch.float32:
        assert oracle.compare_outputs(want, got, 5e-3, 1.0) == 0       # the reference's own rule (attention_helper.h:184)
        assert np.abs(got - want).max() / np.abs(want).max() < 2e-3
    else:
        assert np.abs(got - want).max() <= 2e-2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_ring_nccl_multi_gpu():
    n = min(torch.cuda.device_count(), 8)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "ring_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    print(r.stdout[-3000:], r.stderr[-3000:])
    assert r.returncode == 0 and "RING_OK" in r.stdout


def test_single_rank_communicator_bootstrap_exchange_and_ring_calls():
    """The NCCL side of the ring on the one GPU a test box has: b200_ring_unique_id -> b200_ring_comm_init (a real
    1-rank communicator, init_nccl_comm src/util/nccl_utils.h:62-103), b200_ring_exchange_kv to itself (the grouped
    send(next)/recv(prev) of ring_exchange_kv, nccl_utils.h:133-142, with next = prev = 0), the ring calls on that
    communicator, b200_ring_comm_destroy."""
    import ctypes as C
    lib = fa.lib
    idbuf = (C.c_ubyte * 128)()
    assert lib.b200_ring_unique_id(idbuf) == 0
    assert any(idbuf)
    comm = C.c_void_p(None)
    assert lib.b200_ring_comm_init(C.byref(comm), idbuf, 0, 1) == 0 and comm.value
    try:
        rng = np.random.default_rng(11)
        N, d = 640, 128
        q, k, v, do = [dev(rng.standard_normal((1, 2, N, d)).astype(np.float32), torch.bfloat16) for _ in range(4)]
        kr, vr = torch.zeros_like(k), torch.zeros_like(v)
        st = torch.cuda.current_stream().cuda_stream
        assert lib.b200_ring_exchange_kv(k.data_ptr(), kr.data_ptr(), v.data_ptr(), vr.data_ptr(), k.numel() * 2, 0, 0, comm, st) == 0
        torch.cuda.synchronize()
        assert torch.equal(kr, k) and torch.equal(vr, v)
        rc = fa.RingComm(rank=0, nranks=1)
        rc.handle = comm
        O, L = fa.ring_attention_forward(q, k, v, rc)
        dQ, dK, dV = fa.ring_attention_backward(q, k, v, O, L, do, rc)
        O1, L1 = fa.flash_attention_2_forward(q, k, v, 0.0)
        g1 = fa.flash_attention_2_backward(q, k, v, O1, L1, do, 0.0)
        torch.cuda.synchronize()
        assert torch.equal(O, O1) and torch.equal(L, L1)
        for a, b in zip((dQ, dK, dV), g1):
            assert np.abs(host(a) - host(b)).max() <= 2e-2
        want, _ = oracle.forward(host(q[0, 1]), host(k[0, 1]), host(v[0, 1]), 0.0)
        assert np.abs(host(O[0, 1]) - want).max() <= 2e-2
    finally:
        assert lib.b200_ring_comm_destroy(comm) == 0
