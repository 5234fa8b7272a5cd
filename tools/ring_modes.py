"""Ring forward: stepwise (one launch per step + n-way merge) against persistent (one launch, in-kernel arrival waits)
for a list of per-rank shapes.  torchrun worker:
    python -m torch.distributed.run --nproc-per-node P tools/ring_modes.py [B,H,Nl ...]
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import cuda_flashattention_b200 as fa  # noqa: E402
from cuda_flashattention_b200._lib import lib, check  # noqa: E402


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    comm = fa.RingComm()
    shapes = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]] or [(1, 1, 16384), (1, 2, 16384), (1, 4, 8192), (1, 16, 8192)]
    for (B, H, Nl) in shapes:
        g = torch.Generator(device="cuda").manual_seed(rank)
        q, k, v = [torch.randn(B, H, Nl, 128, device="cuda", generator=g).bfloat16() for _ in range(3)]
        res = {}
        outs = {}
        for mode, name in ((1, "stepwise"), (2, "persistent")):
            check(lib.b200_ring_set_forward_mode(mode), "mode")
            for _ in range(3):
                O, L = fa.ring_attention_forward(q, k, v, comm, 0.0)
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 10
            e0.record()
            for _ in range(n):
                O, L = fa.ring_attention_forward(q, k, v, comm, 0.0)
            e1.record()
            dist.barrier(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / n], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[name] = float(t.item())
            outs[name] = (O.float(), L.clone())
        dO = (outs["stepwise"][0] - outs["persistent"][0]).abs().max().item()
        dL = (outs["stepwise"][1] - outs["persistent"][1]).abs().max().item()
        if rank == 0:
            ctas = B * H * ((Nl + 255) // 256)
            print(f"P={world} B={B} H={H} Nl={Nl}: {ctas} CTAs/rank  stepwise {res['stepwise']:.3f} ms  persistent {res['persistent']:.3f} ms  "
                  f"(x{res['stepwise'] / res['persistent']:.3f})  max|O diff| {dO:.2e}  max|L diff| {dL:.2e}", flush=True)
    check(lib.b200_ring_set_forward_mode(0), "mode")
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
