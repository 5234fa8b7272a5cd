"""Packed variable-length batch: ONE b200_fa2_*_varlen call (sequences spread over the library's side streams) against the
same sequences as single-sequence calls back to back on one stream (what the call did before r2h).  bf16, H = 16, d = 128."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import cuda_flashattention_b200 as fa

def run(n_seqs, n_tok, H=16, d=128, iters=20):
    torch.manual_seed(0)
    total = n_seqs * n_tok
    q, k, v, do = [torch.randn(total, H, d, device="cuda").to(torch.bfloat16) for _ in range(4)]
    cu = [i * n_tok for i in range(n_seqs + 1)]
    def packed():
        O, L = fa.attention_varlen_forward(q, k, v, cu, cu, causal=True)
        return fa.attention_varlen_backward(q, k, v, O, L, do, cu, cu, causal=True)
    def serial():
        outs = []
        for i in range(n_seqs):
            s = slice(cu[i], cu[i + 1])
            O, L = fa.attention_varlen_forward(q[s], k[s], v[s], [0, n_tok], [0, n_tok], causal=True)
            outs.append(fa.attention_varlen_backward(q[s], k[s], v[s], O, L, do[s], [0, n_tok], [0, n_tok], causal=True))
        return outs
    a = packed(); b = serial()
    torch.cuda.synchronize()
    err = max((a[t][cu[i]:cu[i + 1]].float() - b[i][t].float()).abs().max().item() for i in range(n_seqs) for t in range(3))
    res = {}
    for name, fn in (("packed", packed), ("serial", serial)):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters): fn()
        e1.record(); torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / iters
    print(f"{n_seqs} x {n_tok} tokens, H={H} d={d} causal fwd+bwd: one packed call {res['packed']:.3f} ms, "
          f"per-sequence calls {res['serial']:.3f} ms, max |packed - serial| {err:.2e}")

for n_seqs, n_tok in ((32, 512), (16, 1024), (8, 4096), (64, 256)):
    run(n_seqs, n_tok)
