"""Tiny ncu target: the C2-shaped fp32 (tf32 tensor-core) forward and backward."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import cuda_flashattention_b200 as fa
B, H, N, d = 1, 8, 4096, 64
torch.manual_seed(0)
q, k, v, do = [torch.randn(B, H, N, d, device="cuda") for _ in range(4)]
for _ in range(3):
    O, L = fa.flash_attention_2_forward(q, k, v, 0.0)
    g = fa.flash_attention_2_backward(q, k, v, O, L, do, 0.0)
torch.cuda.synchronize()
print("ok")
