"""Single-GPU timing of what one rank of a P-way ring forward enqueues (P attention launches on two
alternating streams into P partial buffers + one n-way merge), with no transport at all, beside one
launch over the same total work.  Separates kernel-granularity cost from transport/skew cost."""
import os
import sys
import torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import cuda_flashattention_b200 as fa  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 4
BH_total, N, d = 128, 8192, 128
B = BH_total // 16 // P
dev = "cuda"
q = torch.randn(B, 16, N, d, device=dev, dtype=torch.bfloat16)
ks = [torch.randn(B, 16, N, d, device=dev, dtype=torch.bfloat16) for _ in range(P)]
vs = [torch.randn(B, 16, N, d, device=dev, dtype=torch.bfloat16) for _ in range(P)]
streams = [torch.cuda.Stream(), torch.cuda.Stream()]


def ring_like(two_streams):
    main = torch.cuda.current_stream()
    outs = []
    for s in range(P):
        st = streams[s & 1] if two_streams else main
        st.wait_stream(main) if s < 2 and two_streams else None
        with torch.cuda.stream(st):
            outs.append(fa.flash_attention_2_forward(q, ks[s], vs[s], 0.0))
    if two_streams:
        main.wait_stream(streams[0]); main.wait_stream(streams[1])
    O, L = outs[0]
    for Ob, Lb in outs[1:]:
        fa.lse_merge_(O, L, Ob, Lb)
    return O


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


qq = torch.randn(8, 16, N, d, device=dev, dtype=torch.bfloat16)
kk, vv = torch.randn_like(qq), torch.randn_like(qq)
print("one launch B=8 H=16 N=8192          : med %.3f best %.3f ms" % timeit(lambda: fa.flash_attention_2_forward(qq, kk, vv, 0.0)))
print("P=%d launches B=%d, one stream + merges: med %.3f best %.3f ms" % ((P, B) + timeit(lambda: ring_like(False))))
print("P=%d launches B=%d, two streams + merges: med %.3f best %.3f ms" % ((P, B) + timeit(lambda: ring_like(True))))
