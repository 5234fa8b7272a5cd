"""Quick device-timed TFLOP/s probe (not the contract bench): python tools/quick_bench.py [fwd|bwd] B H N d dtype"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import cuda_flashattention_b200 as fa

mode = sys.argv[1] if len(sys.argv) > 1 else "fwd"
B, H, N, d = [int(x) for x in (sys.argv[2:6] if len(sys.argv) > 5 else (8, 16, 8192, 128))]
dt = torch.float32 if (len(sys.argv) > 6 and sys.argv[6] == "f32") else torch.bfloat16
torch.manual_seed(0)
q, k, v = [torch.randn(B, H, N, d, device="cuda", dtype=torch.float32).to(dt) for _ in range(3)]
O, L = fa.flash_attention_2_forward(q, k, v, 0.0)
if mode == "bwd":
    dO = torch.randn_like(O)
    fn = lambda: fa.flash_attention_2_backward(q, k, v, O, L, dO, 0.0)
    flops = 10.0 * B * H * N * N * d
else:
    fn = lambda: fa.flash_attention_2_forward(q, k, v, 0.0)
    flops = 4.0 * B * H * N * N * d
for _ in range(3):
    fn()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
iters = 10
e0.record()
for _ in range(iters):
    fn()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"{mode} B={B} H={H} N={N} d={d} {dt}: {ms:.3f} ms  {flops / ms / 1e9:.1f} TFLOP/s")
