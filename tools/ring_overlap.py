"""Compute / transfer overlap probe for the ring path (the B200 counterpart of the reference's
src/03_flash_attention_v2_ring/02_overlap.cu, which simulates the pattern with a dummy kernel): times, per rank,
  (a) the P attention launches of a ring forward with no transfer at all,
  (b) the P-1 K/V block exchanges alone (b200_ring_exchange_kv, NCCL send/recv),
  (c) the real b200_ring_fwd (transport chosen by B200_RING_TRANSPORT: copy engines by default, nccl),
and prints the exposed (non-overlapped) time c - a and the overlap efficiency a / c.
  torchrun --nproc-per-node P tools/ring_overlap.py [B H N_local d]"""
import os
import sys

import torch
import torch.distributed as dist

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import cuda_flashattention_b200 as fa  # noqa: E402
from cuda_flashattention_b200._lib import lib, check  # noqa: E402


def timed(fn, n=5):
    for _ in range(2):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    dist.barrier(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    comm = fa.RingComm()
    B, H, Nl, d = [int(x) for x in sys.argv[1:5]] if len(sys.argv) >= 5 else (max(1, 8 // world), 16, 8192, 128)
    q, k, v = [torch.randn(B, H, Nl, d, device="cuda", dtype=torch.bfloat16) for _ in range(3)]
    kr, vr = torch.empty_like(k), torch.empty_like(v)
    st = torch.cuda.current_stream().cuda_stream

    def compute_only():
        for _ in range(world):
            fa.flash_attention_2_forward(q, k, v, 0.0)

    def transfer_only():
        for _ in range(world - 1):
            check(lib.b200_ring_exchange_kv(k.data_ptr(), kr.data_ptr(), v.data_ptr(), vr.data_ptr(), k.numel() * 2,
                                            (rank + 1) % world, (rank - 1 + world) % world, comm.handle, st), "exchange")

    ta, tb = timed(compute_only), timed(transfer_only)
    tc = timed(lambda: fa.ring_attention_forward(q, k, v, comm, 0.0))
    if rank == 0:
        mb = 2 * k.numel() * 2 / 1e6
        print(f"P={world} B={B} H={H} N_local={Nl} d={d}  transport={os.environ.get('B200_RING_TRANSPORT', 'copy-engine')}")
        print(f"  (a) {world} attention launches, no transfer : {ta:8.3f} ms")
        print(f"  (b) {world - 1} K/V exchanges alone ({mb:.0f} MB each)   : {tb:8.3f} ms  ({(world - 1) * mb / tb:.0f} GB/s per direction)")
        print(f"  (c) ring forward                          : {tc:8.3f} ms")
        print(f"  exposed transfer/merge time c - a = {tc - ta:.3f} ms;  overlap efficiency a / c = {100 * ta / tc:.1f} %")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
