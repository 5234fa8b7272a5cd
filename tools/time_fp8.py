"""C3-shaped forward in fp8 (e4m3 in, bf16 out) beside the bf16 forward: CUDA-event timing + error against the bf16 result."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import cuda_flashattention_b200 as fa

B, H, N, d = 8, 16, 8192, 128
g = torch.Generator(device="cuda").manual_seed(0)
q, k, v = [torch.randn(B, H, N, d, device="cuda", generator=g).to(torch.float8_e4m3fn) for _ in range(3)]
qb, kb, vb = q.to(torch.bfloat16), k.to(torch.bfloat16), v.to(torch.bfloat16)      # the same values, exactly representable


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best, tot = 1e9, 0.0
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        t = e0.elapsed_time(e1); best = min(best, t); tot += t
    return tot / n, best


fl = 4.0 * B * H * N * N * d
for name, fn in (("bf16", lambda: fa.flash_attention_2_forward(qb, kb, vb, 0.0)), ("fp8 ", lambda: fa.flash_attention_2_forward(q, k, v, 0.0))):
    avg, best = timed(fn)
    print(f"{name} forward B={B} H={H} N={N} d={d}: avg {avg:.3f} ms ({fl / avg / 1e9:.0f} TFLOP/s)  best {best:.3f} ms ({fl / best / 1e9:.0f})")
O8, L8 = fa.flash_attention_2_forward(q, k, v, 0.0)
Ob, Lb = fa.flash_attention_2_forward(qb, kb, vb, 0.0)
print("fp8 vs bf16 kernel on the same values: max|dO| %.3e  rel-L2 %.3e  max|dL| %.3e" %
      ((O8.float() - Ob.float()).abs().max().item(), ((O8.float() - Ob.float()).norm() / Ob.float().norm()).item(), (L8 - Lb).abs().max().item()))
