"""ctypes bindings for oracle/libnaive_oracle.so (our C restatement of the reference CPU math)
and, when present, oracle/_ref/libref_naive.so (the unmodified reference CPU code compiled by
oracle/Makefile).  TEST INFRASTRUCTURE ONLY -- see naive_oracle.c header.

All functions take/return contiguous float32 numpy arrays in the reference layout:
Q,K,V,O,dO,dQ,dK,dV are row-major [N, d]; L/l/m are [N]
(reference: src/util/naive_attention.h:7-8, :84-87).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "libnaive_oracle.so")
_REF = os.path.join(_HERE, "_ref", "libref_naive.so")
_REF_CUDA_FWD = os.path.join(_HERE, "_ref", "libref_cuda_fwd.so")
_REF_CUDA_BWD = os.path.join(_HERE, "_ref", "libref_cuda_bwd.so")
_REF_CUDA_RING = os.path.join(_HERE, "_ref", "libref_cuda_ring.so")

_fp = C.POINTER(C.c_float)
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def build(force=False):
    """Compile the checker (gcc only).  Called by __graft_entry__.build() and lazily by tests."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(os.path.join(_HERE, "naive_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "libnaive_oracle.so"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/src") and (force or not os.path.exists(_REF)):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    if os.path.isdir("/root/reference/src") and (force or not os.path.exists(_REF_CUDA_FWD) or not os.path.exists(_REF_CUDA_BWD)
                                               or not os.path.exists(_REF_CUDA_RING)):
        import shutil
        if shutil.which("nvcc"):
            subprocess.check_call(["make", "-C", _HERE, "refcuda"], stdout=subprocess.DEVNULL)


_lib = None
_ref = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
    return _lib


def ref_available():
    return os.path.exists(_REF)


def ref():
    """The unmodified reference CPU implementation (oracle/_ref); None if it was never built."""
    global _ref
    if _ref is None and os.path.exists(_REF):
        _ref = C.CDLL(_REF)
    return _ref


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_fp)


def num_threads():
    return int(lib().oracle_num_threads())


def forward(Q, K, V, scale=0.0, want_L=True, causal=False):
    """oracle_forward == naive_forward_pass (src/util/naive_attention.h:7-61); Nq may differ from Nkv.
    causal=True is an extension (the reference is non-causal): row i sees keys 0 .. i + (Nkv - Nq)."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    Nq, d = Q.shape; Nkv = K.shape[0]
    O = np.empty((Nq, d), np.float32); L = np.empty((Nq,), np.float32)
    fn = lib().oracle_forward_causal if causal else lib().oracle_forward
    fn(q, k, v, O.ctypes.data_as(_fp), L.ctypes.data_as(_fp) if want_L else None,
                         C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return (O, L) if want_L else O


def forward_rows(Q, K, V, rows, scale=0.0):
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    d = Q.shape[1]; Nkv = K.shape[0]
    O = np.empty((len(rows), d), np.float32); L = np.empty((len(rows),), np.float32)
    lib().oracle_forward_rows(q, k, v, O.ctypes.data_as(_fp), L.ctypes.data_as(_fp),
                              rows.ctypes.data_as(_ip), C.c_int(len(rows)), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return O, L


def forward_batched(Q, K, V, scale=0.0, nthreads=1):
    """Q [BH,Nq,d], K,V [BH,Nkv,d]: one reference call per (b,h) slab, optionally on several cores."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    BH, Nq, d = Q.shape; Nkv = K.shape[1]
    O = np.empty_like(Q); L = np.empty((BH, Nq), np.float32)
    lib().oracle_forward_batched(q, k, v, O.ctypes.data_as(_fp), L.ctypes.data_as(_fp), C.c_int(BH),
                                 C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale), C.c_int(nthreads))
    return O, L


def forward_f64(Q, K, V, scale=0.0):
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    Nq, d = Q.shape; Nkv = K.shape[0]
    O = np.empty((Nq, d), np.float64); L = np.empty((Nq,), np.float64)
    lib().oracle_forward_f64(q, k, v, O.ctypes.data_as(_dp), L.ctypes.data_as(_dp),
                             C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return O, L


def attention_rowwise(Q, K, V):
    """oracle_attention_rowwise == naive_attention (src/00_naive_attention/main.cpp:8-38)."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    N, d = Q.shape
    O = np.empty((N, d), np.float32)
    lib().oracle_attention_rowwise(q, k, v, O.ctypes.data_as(_fp), C.c_int(N), C.c_int(d))
    return O


def fa1(Q, K, V):
    """FlashAttention-1 outputs (O, l, m), src/01_flash_attention_v1/flash_attention_kernel.cu:113-128,156-157."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    N, d = Q.shape
    O = np.empty((N, d), np.float32); l = np.empty((N,), np.float32); m = np.empty((N,), np.float32)
    lib().oracle_fa1(q, k, v, O.ctypes.data_as(_fp), l.ctypes.data_as(_fp), m.ctypes.data_as(_fp), C.c_int(N), C.c_int(d))
    return O, l, m


def backward_jacobian(Q, K, V, dO, scale):
    """oracle_backward_jacobian == naive_attention_backward (src/util/naive_attention.h:84-161), O(N^3)."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V); dO, g = _f(dO)
    N, d = Q.shape
    dQ = np.empty((N, d), np.float32); dK = np.empty((N, d), np.float32); dV = np.empty((N, d), np.float32)
    lib().oracle_backward_jacobian(q, k, v, None, None, g, dQ.ctypes.data_as(_fp), dK.ctypes.data_as(_fp),
                                   dV.ctypes.data_as(_fp), C.c_int(N), C.c_int(d), C.c_float(scale))
    return dQ, dK, dV


def backward(Q, K, V, dO, scale=0.0, causal=False):
    """O(N^2 d) restatement (dS = P o (dP - D)), double accumulation; Nq may differ from Nkv.  causal: see forward."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V); dO, g = _f(dO)
    Nq, d = Q.shape; Nkv = K.shape[0]
    dQ = np.empty((Nq, d), np.float32); dK = np.empty((Nkv, d), np.float32); dV = np.empty((Nkv, d), np.float32)
    fn = lib().oracle_backward_causal if causal else lib().oracle_backward
    fn(q, k, v, g, dQ.ctypes.data_as(_fp), dK.ctypes.data_as(_fp), dV.ctypes.data_as(_fp),
                          C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return dQ, dK, dV


def row_stats(Q, K, V, dO, scale=0.0):
    """Per query row: (log-sum-exp of the scaled scores, delta = sum_j P_ij dP_ij), fp64, all host threads."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V); dO, g = _f(dO)
    Nq, d = Q.shape; Nkv = K.shape[0]
    lse = np.empty((Nq,), np.float64); delta = np.empty((Nq,), np.float64)
    lib().oracle_row_stats(q, k, v, g, lse.ctypes.data_as(_dp), delta.ctypes.data_as(_dp),
                           C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return lse, delta


def backward_rows(Q, K, V, dO, qrows, kvrows, scale=0.0, stats=None):
    """Row-sampled backward: dQ for the query rows `qrows`, (dK, dV) for the key rows `kvrows` -- the arithmetic of
    oracle.backward (naive_attention_backward, src/util/naive_attention.h:84-161) on those rows only.  `stats` =
    (lse, delta) for every query row; None computes them here (row_stats: O(Nq Nkv d))."""
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V); dO, g = _f(dO)
    Nq, d = Q.shape; Nkv = K.shape[0]
    lse, delta = row_stats(Q, K, V, dO, scale) if stats is None else stats
    lse = np.ascontiguousarray(lse, dtype=np.float64); delta = np.ascontiguousarray(delta, dtype=np.float64)
    assert lse.shape == (Nq,) and delta.shape == (Nq,)
    qrows = np.ascontiguousarray(qrows, dtype=np.int32); kvrows = np.ascontiguousarray(kvrows, dtype=np.int32)
    dQ = np.empty((len(qrows), d), np.float32); dK = np.empty((len(kvrows), d), np.float32); dV = np.empty((len(kvrows), d), np.float32)
    lib().oracle_backward_rows(q, k, v, g, lse.ctypes.data_as(_dp), delta.ctypes.data_as(_dp),
                               qrows.ctypes.data_as(_ip), C.c_int(len(qrows)), kvrows.ctypes.data_as(_ip), C.c_int(len(kvrows)),
                               dQ.ctypes.data_as(_fp), dK.ctypes.data_as(_fp), dV.ctypes.data_as(_fp),
                               C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale))
    return dQ, dK, dV


def ring_step(Q, Kblk, Vblk, O, Lbuf, M, scale, last_step):
    """In-place carry step, src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:67-139."""
    Q, q = _f(Q); Kblk, k = _f(Kblk); Vblk, v = _f(Vblk)
    assert O.dtype == np.float32 and Lbuf.dtype == np.float32 and M.dtype == np.float32
    Nq, d = Q.shape; Nkv = Kblk.shape[0]
    lib().oracle_ring_step(q, k, v, O.ctypes.data_as(_fp), Lbuf.ctypes.data_as(_fp), M.ctypes.data_as(_fp),
                           C.c_int(Nq), C.c_int(Nkv), C.c_int(d), C.c_float(scale), C.c_int(1 if last_step else 0))


def lse_merge(Oa, La, Ob, Lb):
    """In place on (Oa, La): merge two normalised partials by log-sum-exp."""
    Ob, ob = _f(Ob); Lb, lb = _f(Lb)
    N, d = Oa.shape
    lib().oracle_lse_merge(Oa.ctypes.data_as(_fp), La.ctypes.data_as(_fp), ob, lb, C.c_int(N), C.c_int(d))


def compare_outputs(ref_arr, test_arr, rtol=1e-3, atol=1.0):
    """Failing-element count under the ring test's rule (src/util/attention_helper.h:174-208)."""
    a, pa = _f(np.ravel(ref_arr)); b, pb = _f(np.ravel(test_arr))
    return int(lib().oracle_compare_outputs(pa, pb, C.c_int(a.size), C.c_float(rtol), C.c_float(atol)))


# ---- the unmodified reference (oracle/_ref) -------------------------------------------------
def ref_forward_pass(Q, K, V, scale):
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    N, d = Q.shape
    O = np.empty((N, d), np.float32); L = np.empty((N,), np.float32)
    ref().ref_naive_forward_pass(q, k, v, O.ctypes.data_as(_fp), L.ctypes.data_as(_fp), C.c_int(N), C.c_int(d), C.c_float(scale))
    return O, L


def ref_backward(Q, K, V, dO, scale):
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V); dO, g = _f(dO)
    N, d = Q.shape
    dQ = np.empty((N, d), np.float32); dK = np.empty((N, d), np.float32); dV = np.empty((N, d), np.float32)
    ref().ref_naive_attention_backward(q, k, v, None, None, g, dQ.ctypes.data_as(_fp), dK.ctypes.data_as(_fp),
                                       dV.ctypes.data_as(_fp), C.c_int(N), C.c_int(d), C.c_float(scale))
    return dQ, dK, dV


def ref_attention_rowwise(Q, K, V):
    Q, q = _f(Q); K, k = _f(K); V, v = _f(V)
    N, d = Q.shape
    O = np.empty((N, d), np.float32)
    ref().ref_naive_attention(q, k, v, O.ctypes.data_as(_fp), C.c_int(N), C.c_int(d))
    return O


def ref_fixture(name, n_elems):
    """Reference test inputs that depend on glibc rand() (see ref_wrapper.cpp)."""
    arrs = [np.empty((n_elems,), np.float32) for _ in range(4 if name == "bwd_random" else 3)]
    fn = getattr(ref(), "ref_fixture_" + name)
    fn(*[a.ctypes.data_as(_fp) for a in arrs], C.c_int(n_elems))
    return arrs


# ---- the unmodified reference CUDA kernels (oracle/_ref/libref_cuda_*.so, built by `make refcuda`) ----------------
# Device-pointer calls on torch CUDA float32 tensors [N, d]; they launch on the legacy default stream like the
# reference wrappers do and return cudaGetLastError().  d <= 64 only (reference assert).
_ref_cuda = {}


def ref_cuda_available():
    return os.path.exists(_REF_CUDA_FWD) and os.path.exists(_REF_CUDA_BWD)


def _ref_cuda_lib(which):
    if which not in _ref_cuda:
        _ref_cuda[which] = C.CDLL(_REF_CUDA_FWD if which == "fwd" else _REF_CUDA_BWD)
    return _ref_cuda[which]


def ref_cuda_forward(Q, K, V, O, L, scale):
    """flash_attention_2_forward, reference src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu:133-172."""
    N, d = Q.shape
    vp = C.c_void_p
    return int(_ref_cuda_lib("fwd").ref_cuda_fa2_forward(vp(Q.data_ptr()), vp(K.data_ptr()), vp(V.data_ptr()), vp(O.data_ptr()),
                                                         vp(L.data_ptr()), C.c_int(N), C.c_int(d), C.c_float(scale)))


def ref_cuda_backward(Q, K, V, O, L, dO, dQ, dK, dV, scale):
    """flash_attention_2_backward, reference src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:249-299."""
    N, d = Q.shape
    vp = C.c_void_p
    return int(_ref_cuda_lib("bwd").ref_cuda_fa2_backward(*[vp(t.data_ptr()) for t in (Q, K, V, O, L, dO, dQ, dK, dV)],
                                                          C.c_int(N), C.c_int(d), C.c_float(scale)))


def ref_cuda_ring_available():
    return os.path.exists(_REF_CUDA_RING)


def ref_cuda_ring_forward(Q, K, V, P, scale):
    """The reference's ring_attention_forward_kernel (src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:13-140)
    run for the P ranks of a ring one after the other on one GPU, with the schedule and launch configuration of its host
    loop (:159-214); torch CUDA float32 tensors [N, d], N % P == 0, d <= 64.  -> (O [N, d], L [N])."""
    import torch
    N, d = Q.shape
    if "ring" not in _ref_cuda:
        _ref_cuda["ring"] = C.CDLL(_REF_CUDA_RING)
    O = torch.empty_like(Q)
    L = torch.empty(N + 64, dtype=torch.float32, device=Q.device)
    M = torch.empty(N + 64, dtype=torch.float32, device=Q.device)
    torch.cuda.synchronize()
    vp = C.c_void_p
    rc = int(_ref_cuda["ring"].ref_cuda_ring_forward_emulated(vp(Q.data_ptr()), vp(K.data_ptr()), vp(V.data_ptr()), vp(O.data_ptr()),
                                                              vp(L.data_ptr()), vp(M.data_ptr()), C.c_int(N), C.c_int(d), C.c_int(P),
                                                              C.c_float(scale)))
    if rc != 0:
        raise RuntimeError(f"ref_cuda_ring_forward_emulated: {rc}")
    return O, L[:N]
