"""Generate tests/golden/*.npz by running the UNMODIFIED reference CPU code (oracle/_ref,
compiled by oracle/Makefile from /root/reference) on the reference's own test fixtures.

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
The .npz files are committed; nothing at test time reads /root/reference.

Fixtures:
  G0  2x2            src/00_naive_attention/main.cpp:41-61 (expected values hard-coded there)
  G1  fwd 4x4        src/02_flash_attention_v2_forward/main.cu:134-155
  G2  bwd 4x4        src/02_flash_attention_v2_backward/main.cu:78-107
  G3  fwd N=512 d=64 src/02_flash_attention_v2_forward/main.cu:14-33   (glibc srand(42))
  G4  bwd N=128 d=64 src/02_flash_attention_v2_backward/main.cu:200-227 (glibc srand(42))
  G5  ring N=5096 d=64 scale 1  src/util/attention_helper.h:151-173, src/03_flash_attention_v2_ring/04_ring_attention.cu:19-21
  FA1 the ten cases of src/01_flash_attention_v1/main.cu:195-347
  C1  BASELINE config 1: N=1024 d=64 (G3-style inputs)
"""
import os
import sys
import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
import oracle  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
f32 = np.float32


def simple_test_data(N, d):
    """create_simple_test_data, src/util/attention_helper.h:151-173."""
    Q = np.zeros((N, d), f32)
    for i in range(min(N, d)):
        Q[i, i] = 1.0
    K = Q.copy()
    i = np.arange(N, dtype=f32)[:, None]
    j = np.arange(d, dtype=f32)[None, :]
    V = (i * f32(4.0) + j + f32(1.0)).astype(f32)
    return Q, K, V


def main():
    oracle.build(force=True)
    assert oracle.ref_available(), "needs /root/reference"
    g = {}

    # G0
    Q = np.eye(2, dtype=f32); V = np.array([[1, 2], [3, 4]], f32)
    g["G0"] = dict(Q=Q, K=Q, V=V, O=oracle.ref_attention_rowwise(Q, Q, V),
                   O_literal=np.array([[1.6604769, 2.6604770], [2.3395231, 3.3395231]], f32))

    # G1
    Q = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0], [1, 1, 0, 0]], f32)
    V = np.arange(1, 17, dtype=f32).reshape(4, 4)
    O, L = oracle.ref_forward_pass(Q, Q, V, 1.0)
    g["G1"] = dict(Q=Q, K=Q, V=V, O=O, L=L, scale=f32(1.0))

    # G2
    Q = np.eye(4, dtype=f32)
    V = np.repeat(np.arange(1, 5, dtype=f32)[:, None], 4, axis=1)
    dO = np.eye(4, dtype=f32)
    O, L = oracle.ref_forward_pass(Q, Q, V, 1.0)
    dQ, dK, dV = oracle.ref_backward(Q, Q, V, dO, 1.0)
    g["G2"] = dict(Q=Q, K=Q, V=V, dO=dO, O=O, L=L, dQ=dQ, dK=dK, dV=dV, scale=f32(1.0))

    # G3
    N, d = 512, 64
    Q, K, V = [a.reshape(N, d) for a in oracle.ref_fixture("fwd_random", N * d)]
    O, L = oracle.ref_forward_pass(Q, K, V, 1.0 / np.sqrt(f32(d)))
    g["G3"] = dict(Q=Q, K=K, V=V, O=O, L=L, scale=f32(1.0 / np.sqrt(f32(d))))

    # G4
    N, d = 128, 64
    Q, K, V, dO = [a.reshape(N, d) for a in oracle.ref_fixture("bwd_random", N * d)]
    sc = f32(1.0 / np.sqrt(f32(d)))
    O, L = oracle.ref_forward_pass(Q, K, V, sc)
    dQ, dK, dV = oracle.ref_backward(Q, K, V, dO, sc)
    g["G4"] = dict(Q=Q, K=K, V=V, dO=dO, O=O, L=L, dQ=dQ, dK=dK, dV=dV, scale=sc)

    # G5 (inputs are formulaic; store outputs as fp32, compressed)
    N, d = 5096, 64
    Q, K, V = simple_test_data(N, d)
    O, L = oracle.ref_forward_pass(Q, K, V, 1.0)
    g["G5"] = dict(O=O, L=L, N=np.int32(N), d=np.int32(d), scale=f32(1.0))

    # FA1 cases (run_test compares O with naive_attention, src/01_flash_attention_v1/main.cu:110-165)
    fa1 = {}
    def add(name, Q, K, V):
        Q = np.asarray(Q, f32); K = np.asarray(K, f32); V = np.asarray(V, f32)
        fa1[name + "_Q"] = Q; fa1[name + "_K"] = K; fa1[name + "_V"] = V
        fa1[name + "_O"] = oracle.ref_attention_rowwise(Q, K, V)
    add("t1", [[1, 0, 1, 0], [0, 1, 0, 1]], [[1, 0, 1, 0], [0, 1, 0, 1]], [[10, 20, 30, 40], [50, 60, 70, 80]])
    add("t2", np.eye(4), np.eye(4), np.diag([1, 2, 3, 4]))
    add("t3", np.ones((3, 2)), np.ones((3, 2)), [[1, 2], [3, 4], [5, 6]])
    add("t4", [[1, 0], [0, 1]], [[0, 1], [-1, 0]], [[10, 20], [30, 40]])
    add("t5", [[1]], [[1]], [[42]])
    Qd = np.zeros((8, 4), f32); Vd = np.zeros((8, 4), f32)
    for i in range(8):
        for j in range(4):
            Qd[i, j] = 1.0 if i == j else 0.0
            Vd[i, j] = i * 10 + j
    add("t6", Qd, Qd, Vd)
    Qr, Kr, Vr = [a.reshape(64, 32) for a in oracle.ref_fixture("fa1_random", 64 * 32)]
    add("t7", Qr, Kr, Vr)
    add("t8", np.eye(4), np.eye(4), np.arange(1, 17, dtype=f32).reshape(4, 4))
    g["FA1"] = fa1

    # C1: BASELINE config 1 (N=1024, d=64), G3-style inputs, forward + backward via the reference
    N, d = 1024, 64
    Q, K, V = [a.reshape(N, d) for a in oracle.ref_fixture("fwd_random", N * d)]
    sc = f32(1.0 / np.sqrt(f32(d)))
    O, L = oracle.ref_forward_pass(Q, K, V, sc)
    g["C1"] = dict(Q=Q, K=K, V=V, O=O, L=L, scale=sc)

    for name, arrs in g.items():
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrs)
        print(name, {k: getattr(v, "shape", ()) for k, v in arrs.items()})


if __name__ == "__main__":
    main()
