"""Forward kernel time against the number of CTAs (Nq = 256 * n rows of one head, Nkv = 8192):
shows the per-CTA duration and how the grid fills the 148 SMs."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import cuda_flashattention_b200 as fa  # noqa: E402

Nkv, d = 8192, 128
k = torch.randn(1, 1, Nkv, d, device="cuda", dtype=torch.bfloat16)
v = torch.randn_like(k)
for n in [1, 2, 74, 140, 148, 149, 160, 222, 296, 297, 444, 445, 512, 592, 593, 1024]:
    q = torch.randn(1, 1, 256 * n, d, device="cuda", dtype=torch.bfloat16)
    for _ in range(3):
        fa.flash_attention_2_forward(q, k, v, 0.0)
    torch.cuda.synchronize()
    ts = []
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(7):
        e0.record(); fa.flash_attention_2_forward(q, k, v, 0.0); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    print("CTAs %4d  (%.2f waves)  med %.4f ms  best %.4f ms" % (n, n / 148, ts[3], ts[0]))
