"""Denominator of the ring-efficiency definition in BASELINE.md section 3: one GPU, the per-rank work of a P-way ring
with no communication -- Nq = N/P local queries against all Nkv = N keys (B=1 H=16 d=128 bf16, C4's shape)."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import cuda_flashattention_b200 as fa  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 8
B, H, N, d = 1, 16, 65536, 128
q = torch.randn(B, H, N // P, d, device="cuda", dtype=torch.bfloat16)
k, v = [torch.randn(B, H, N, d, device="cuda", dtype=torch.bfloat16) for _ in range(2)]
do = torch.randn_like(q) * 0.2
ts = []
for it in range(8):
    e0, e1, e2 = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e0.record()
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0)
    e1.record()
    g = fa.flash_attention_2_backward(q, k, v, O, L, do, 0.0)
    e2.record()
    torch.cuda.synchronize()
    if it >= 3:
        ts.append((e0.elapsed_time(e1), e1.elapsed_time(e2)))
f = sorted(t[0] for t in ts)[len(ts) // 2]
b = sorted(t[1] for t in ts)[len(ts) // 2]
print(f"P={P}: one GPU, Nq={N // P} x Nkv={N}, B={B} H={H}: fwd {f:.3f} ms  bwd {b:.3f} ms  fwd+bwd {f + b:.3f} ms")
