"""Host-side mirror of the reference's operator interface, on torch CUDA tensors.

Same names, argument meaning and layouts as the reference's C++ wrappers (citations are paths in
the reference repository); each call goes straight through the C-ABI of libb200attn.so.
torch is used only for device memory and streams.

    flash_attention_2_forward(Q, K, V, softmax_scale)            -> (O, L)
        src/02_flash_attention_v2_forward/flash_attention_kernel.cu:300
    flash_attention_2_backward(Q, K, V, O, L, dO, softmax_scale) -> (dQ, dK, dV)
        src/02_flash_attention_v2_backward/flash_attention_backward_kernel.cu:249
    flash_attention(Q, K, V)                                     -> (O, l, m)
        src/01_flash_attention_v1/main.cu:7

Tensors are [N, d] (one head, the reference's shape) or [B, H, N, d]; fp32 or bf16; contiguous.
The forward also takes torch.float8_e4m3fn Q/K/V (d = 128; O comes back in bf16).
"""
import torch

from ._lib import lib, check, B200_DTYPE_BF16, B200_DTYPE_F32, B200_DTYPE_FP8_E4M3, AttnDesc, Layout

FLAG_CAUSAL = 1  # B200_FLAG_CAUSAL (include/b200attn.h)


def _dtype_code(t):
    if t.dtype == torch.float32:
        return B200_DTYPE_F32
    if t.dtype == torch.bfloat16:
        return B200_DTYPE_BF16
    if t.dtype == torch.float8_e4m3fn:
        return B200_DTYPE_FP8_E4M3           # forward only, d = 128: fp8 Q/K/V in, bf16 O out
    raise TypeError(f"unsupported dtype {t.dtype}: fp32, bf16 (and fp8 e4m3 for the forward) only")


def _check_inputs(*ts):
    dev, dt = ts[0].device, ts[0].dtype
    for t in ts:
        if not t.is_cuda:
            raise ValueError("cuda_flashattention_b200 has no CPU path: tensors must live on a CUDA device")
        if t.device != dev or t.dtype != dt:
            raise ValueError("all tensors must share device and dtype")
        if not t.is_contiguous():
            raise ValueError("tensors must be contiguous row-major [.., N, d] (the reference layout)")


def _bhnd(t):
    if t.dim() == 2:
        return 1, 1, t.shape[0], t.shape[1]
    if t.dim() == 4:
        return tuple(t.shape)
    raise ValueError("expected [N, d] or [B, H, N, d]")


def _stream():
    return torch.cuda.current_stream().cuda_stream


def flash_attention_2_forward(Q, K, V, softmax_scale=0.0, return_lse=True, causal=False):
    """O, L = FA2 forward.  causal=True (an extension, the reference is non-causal) hides key j from query i
    when j > i + (Nkv - Nq)."""
    _check_inputs(Q, K, V)
    B, H, Nq, d = _bhnd(Q)
    Bk, Hk, Nkv, dk = _bhnd(K)
    if (Bk, Hk, dk) != (B, H, d) or tuple(V.shape) != tuple(K.shape):
        raise ValueError("Q/K/V shape mismatch")
    O = torch.empty(Q.shape, dtype=torch.bfloat16 if Q.dtype == torch.float8_e4m3fn else Q.dtype, device=Q.device)
    L = torch.empty(Q.shape[:-1], dtype=torch.float32, device=Q.device)
    with torch.cuda.device(Q.device):
        check(lib.b200_fa2_fwd_ex(Q.data_ptr(), K.data_ptr(), V.data_ptr(), O.data_ptr(), L.data_ptr(),
                                  B, H, Nq, Nkv, d, float(softmax_scale), _dtype_code(Q), FLAG_CAUSAL if causal else 0,
                                  _stream()), "b200_fa2_fwd_ex")
    return (O, L) if return_lse else O


def flash_attention_2_backward(Q, K, V, O, L, dO, softmax_scale=0.0, causal=False):
    _check_inputs(Q, K, V, O, dO)
    B, H, Nq, d = _bhnd(Q)
    Bk, Hk, Nkv, dk = _bhnd(K)
    if (Bk, Hk, dk) != (B, H, d) or tuple(V.shape) != tuple(K.shape) or tuple(O.shape) != tuple(Q.shape) or \
            tuple(dO.shape) != tuple(Q.shape):
        raise ValueError("Q/K/V/O/dO shape mismatch")
    if L.dtype != torch.float32 or not L.is_contiguous() or L.device != Q.device or L.numel() != B * H * Nq:
        raise ValueError("L must be contiguous fp32 [.., N] on the tensors' device")
    dQ, dK, dV = torch.empty_like(Q), torch.empty_like(K), torch.empty_like(V)
    with torch.cuda.device(Q.device):
        check(lib.b200_fa2_bwd_ex(Q.data_ptr(), K.data_ptr(), V.data_ptr(), O.data_ptr(), L.data_ptr(), dO.data_ptr(),
                                  dQ.data_ptr(), dK.data_ptr(), dV.data_ptr(), B, H, Nq, Nkv, d, float(softmax_scale),
                                  _dtype_code(Q), FLAG_CAUSAL if causal else 0, None, 0, _stream()), "b200_fa2_bwd_ex")
    return dQ, dK, dV


def flash_attention(Q, K, V):
    """FlashAttention-1 API: returns (O, l, m) with l = sum exp(s - m), m = rowmax; scale 1/sqrt(d); synchronous."""
    _check_inputs(Q, K, V)
    if Q.dim() != 2 or Q.dtype != torch.float32:
        raise ValueError("flash_attention (FA1 API) takes one fp32 head [N, d]")
    N, d = Q.shape
    O = torch.empty_like(Q)
    l = torch.empty(N, dtype=torch.float32, device=Q.device)
    m = torch.empty(N, dtype=torch.float32, device=Q.device)
    with torch.cuda.device(Q.device):
        check(lib.b200_fa1_fwd_f32(Q.data_ptr(), K.data_ptr(), V.data_ptr(), O.data_ptr(), l.data_ptr(), m.data_ptr(),
                                   N, d, _stream()), "b200_fa1_fwd_f32")
    return O, l, m


def lse_merge_(Oa, La, Ob, Lb):
    """In place on (Oa, La): merge two normalised partials by log-sum-exp."""
    _check_inputs(Oa, Ob)
    B, H, N, d = _bhnd(Oa)
    if tuple(Ob.shape) != tuple(Oa.shape) or any(
            t.dtype != torch.float32 or not t.is_contiguous() or t.device != Oa.device or t.numel() != B * H * N for t in (La, Lb)):
        raise ValueError("lse_merge_: Ob must match Oa; La / Lb must be contiguous fp32 [.., N] on the same device")
    with torch.cuda.device(Oa.device):
        check(lib.b200_lse_merge(Oa.data_ptr(), La.data_ptr(), Ob.data_ptr(), Lb.data_ptr(), B, H, N, d,
                                 _dtype_code(Oa), _stream()), "b200_lse_merge")
    return Oa, La


def uses_tensor_cores(dtype, d, Nq, Nkv, backward=False):
    code = B200_DTYPE_F32 if dtype == torch.float32 else B200_DTYPE_BF16
    fn = lib.b200_bwd_uses_tensor_cores if backward else lib.b200_fwd_uses_tensor_cores
    return bool(fn(code, d, Nq, Nkv))


def _check_host(*ts):
    dt = ts[0].dtype
    for t in ts:
        if t.is_cuda:
            raise ValueError("host entry point: tensors must be CPU tensors (pin them for full PCIe speed)")
        if t.dtype != dt or not t.is_contiguous():
            raise ValueError("host tensors must share dtype and be contiguous row-major [.., N, d]")


def flash_attention_2_forward_host(Q, K, V, softmax_scale=0.0, out=None, device=None):
    """Host tensors in, host tensors out: (O, L).  What the reference's drivers do around every call
    (cudaMemcpy H2D, kernel, cudaMemcpy D2H -- src/02_flash_attention_v2_forward/main.cu:42-64), but
    chunked over the (b,h) slabs with copies and attention overlapped (b200_fa2_fwd_host).
    Stream-ordered: synchronise the current stream before reading the results."""
    _check_host(Q, K, V)
    B, H, N, d = _bhnd(Q)
    O, L = out if out is not None else (torch.empty_like(Q).pin_memory(),
                                        torch.empty(Q.shape[:-1], dtype=torch.float32).pin_memory())
    with torch.cuda.device(device if device is not None else torch.cuda.current_device()):
        check(lib.b200_fa2_fwd_host(Q.data_ptr(), K.data_ptr(), V.data_ptr(), O.data_ptr(), L.data_ptr(),
                                    B, H, N, d, float(softmax_scale), _dtype_code(Q), _stream()), "b200_fa2_fwd_host")
    return O, L


def flash_attention_2_forward_backward_host(Q, K, V, dO, softmax_scale=0.0, out=None, device=None):
    """Host tensors in, host tensors out: (O, L, dQ, dK, dV) -- forward and backward of every chunk of
    (b,h) slabs, pipelined with the PCIe copies in both directions (b200_fa2_fwd_bwd_host; the
    reference's host-side pattern is src/02_flash_attention_v2_backward/main.cu:261-289)."""
    _check_host(Q, K, V, dO)
    B, H, N, d = _bhnd(Q)
    if out is None:
        out = tuple(torch.empty_like(Q).pin_memory() for _ in range(1)) + \
              (torch.empty(Q.shape[:-1], dtype=torch.float32).pin_memory(),) + \
              tuple(torch.empty_like(Q).pin_memory() for _ in range(3))
    O, L, dQ, dK, dV = out
    with torch.cuda.device(device if device is not None else torch.cuda.current_device()):
        check(lib.b200_fa2_fwd_bwd_host(Q.data_ptr(), K.data_ptr(), V.data_ptr(), dO.data_ptr(), O.data_ptr(), L.data_ptr(),
                                        dQ.data_ptr(), dK.data_ptr(), dV.data_ptr(), B, H, N, d, float(softmax_scale),
                                        _dtype_code(Q), _stream()), "b200_fa2_fwd_bwd_host")
    return O, L, dQ, dK, dV


# ---------------------------------------------------------------------------------------------------
# The callers' side of the boundary: strided [B,H,N,d] / token-major [B,N,H,d] tensors,
# grouped K/V heads (GQA / MQA) and an autograd function, all through b200_fa2_fwd_desc / _bwd_desc.
# The reference stops at contiguous single-head slabs (__info__/README.md:175).
# ---------------------------------------------------------------------------------------------------
def _dense(t):
    """A view the descriptors can express: last dimension contiguous and no zero stride on a dimension of extent > 1
    (autograd hands out expanded gradients, e.g. O.sum().backward(); the all-zero layout means "contiguous")."""
    if t.stride(-1) != 1 or any(st == 0 and sz > 1 for st, sz in zip(t.stride(), t.shape)):
        return t.contiguous()
    return t


def _dims(t, layout):
    """(B, H, N, d, Layout) of a 4-D tensor whose last dimension is contiguous."""
    if t.dim() != 4 or t.stride(-1) != 1:
        raise ValueError("expected a 4-D tensor with a contiguous last dimension")
    if layout == "bhnd":
        B, H, N, d = t.shape
        sb, sh, sn = t.stride(0), t.stride(1), t.stride(2)
    elif layout == "bnhd":
        B, N, H, d = t.shape
        sb, sh, sn = t.stride(0), t.stride(2), t.stride(1)
    else:
        raise ValueError("layout must be 'bhnd' or 'bnhd'")
    # a dimension of extent 1 is never stepped over: give it a harmless non-zero stride
    sn = sn if N > 1 else d
    sh = sh if H > 1 else N * sn
    sb = sb if B > 1 else max(H * sh, N * sn)
    return B, H, N, d, Layout(sb, sh, sn)


def _make_desc(q, k, v, layout, causal, softmax_scale):
    for t in (q, k, v):
        if not t.is_cuda:
            raise ValueError("cuda_flashattention_b200 has no CPU path: tensors must live on a CUDA device")
    if not (q.dtype == k.dtype == v.dtype) or not (q.device == k.device == v.device):
        raise ValueError("q, k, v must share device and dtype")
    B, H, Nq, d, lq = _dims(q, layout)
    Bk, Hkv, Nkv, dk, lk = _dims(k, layout)
    Bv, Hv, Nv, dv, lv = _dims(v, layout)
    if (Bk, dk) != (B, d) or (Bv, Hv, Nv, dv) != (Bk, Hkv, Nkv, dk) or Hkv < 1 or H % Hkv != 0:
        raise ValueError("q/k/v shape mismatch (K/V heads must divide the query heads)")
    desc = AttnDesc(B=B, H=H, Hkv=Hkv, Nq=Nq, Nkv=Nkv, d=d, dtype=_dtype_code(q), flags=FLAG_CAUSAL if causal else 0,
                    scale=float(softmax_scale or 0.0), q=lq, k=lk, v=lv)
    return desc, (B, H, Nq)


def attention_forward(q, k, v, causal=False, softmax_scale=None, layout="bhnd"):
    """O (same dimension order as q, contiguous) and L [B,H,Nq] for strided / grouped-head inputs."""
    q, k, v = _dense(q), _dense(k), _dense(v)
    desc, (B, H, Nq) = _make_desc(q, k, v, layout, causal, softmax_scale)
    O = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    L = torch.empty(B, H, Nq, dtype=torch.float32, device=q.device)
    desc.o = _dims(O, layout)[4]
    with torch.cuda.device(q.device):
        check(lib.b200_fa2_fwd_desc(desc, q.data_ptr(), k.data_ptr(), v.data_ptr(), O.data_ptr(), L.data_ptr(), _stream()),
              "b200_fa2_fwd_desc")
    return O, L


def attention_backward(q, k, v, O, L, dO, causal=False, softmax_scale=None, layout="bhnd"):
    q, k, v = _dense(q), _dense(k), _dense(v)
    desc, _ = _make_desc(q, k, v, layout, causal, softmax_scale)
    dO, O = _dense(dO), _dense(O)
    dQ = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    dK = torch.empty(k.shape, dtype=k.dtype, device=k.device)
    dV = torch.empty(v.shape, dtype=v.dtype, device=v.device)
    desc.o, desc.d_o = _dims(O, layout)[4], _dims(dO, layout)[4]
    desc.dq, desc.dk, desc.dv = _dims(dQ, layout)[4], _dims(dK, layout)[4], _dims(dV, layout)[4]
    with torch.cuda.device(q.device):
        check(lib.b200_fa2_bwd_desc(desc, q.data_ptr(), k.data_ptr(), v.data_ptr(), O.data_ptr(), L.data_ptr(), dO.data_ptr(),
                                    dQ.data_ptr(), dK.data_ptr(), dV.data_ptr(), None, 0, _stream()), "b200_fa2_bwd_desc")
    return dQ, dK, dV


# Registered operators ("a torch.library custom-op binding so the kernels are usable from a model"): opaque to
# torch.compile / AOT autograd (fake implementations give shapes and strides, the backward is its own op), so a model
# that calls flash_attn_func traces as ONE node per direction instead of graph-breaking on ctypes.
@torch.library.custom_op("b200attn::fwd", mutates_args=(), device_types="cuda")
def _op_fwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, causal: bool, softmax_scale: float, layout: str) -> tuple[torch.Tensor, torch.Tensor]:
    return attention_forward(q, k, v, causal, softmax_scale, layout)


@_op_fwd.register_fake
def _(q, k, v, causal, softmax_scale, layout):
    B, H, Nq = (q.shape[0], q.shape[1], q.shape[2]) if layout == "bhnd" else (q.shape[0], q.shape[2], q.shape[1])
    return torch.empty_like(q, memory_format=torch.contiguous_format), q.new_empty((B, H, Nq), dtype=torch.float32)


@torch.library.custom_op("b200attn::bwd", mutates_args=(), device_types="cuda")
def _op_bwd(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, O: torch.Tensor, L: torch.Tensor, dO: torch.Tensor, causal: bool,
            softmax_scale: float, layout: str) -> tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
    return attention_backward(q, k, v, O, L, dO, causal, softmax_scale, layout)


@_op_bwd.register_fake
def _(q, k, v, O, L, dO, causal, softmax_scale, layout):
    return (torch.empty_like(q, memory_format=torch.contiguous_format), torch.empty_like(k, memory_format=torch.contiguous_format),
            torch.empty_like(v, memory_format=torch.contiguous_format))


def _op_setup(ctx, inputs, output):
    q, k, v, causal, softmax_scale, layout = inputs
    O, L = output
    ctx.save_for_backward(q, k, v, O, L)
    ctx.cfg = (causal, softmax_scale, layout)


def _op_backward(ctx, dO, dL):
    q, k, v, O, L = ctx.saved_tensors
    causal, softmax_scale, layout = ctx.cfg
    dQ, dK, dV = _op_bwd(q, k, v, O, L, dO.contiguous(), causal, softmax_scale, layout)
    return dQ, dK, dV, None, None, None


_op_fwd.register_autograd(_op_backward, setup_context=_op_setup)


def flash_attn_func(q, k, v, causal=False, softmax_scale=None, layout="bnhd"):
    """Differentiable exact attention on the tcgen05 kernels.  q [B,Nq,H,d], k/v [B,Nkv,Hkv,d] (layout='bnhd', the
    token-major order models use) or [B,H,N,d] (layout='bhnd'); any batch/head/row strides with a contiguous last
    dimension (e.g. slices of a packed qkv projection); Hkv may divide H (GQA / MQA)."""
    return _op_fwd(q, k, v, bool(causal), float(softmax_scale or 0.0), layout)[0]


# ---- packed variable-length batches (b200_fa2_fwd_varlen / b200_fa2_bwd_varlen) ----------------------
def _cu_array(cu):
    import ctypes as C
    vals = [int(x) for x in (cu.tolist() if hasattr(cu, "tolist") else cu)]
    if len(vals) < 1 or vals[0] != 0 or any(b < a for a, b in zip(vals, vals[1:])):
        raise ValueError("cu_seqlens must start at 0 and be non-decreasing")
    return (C.c_int * len(vals))(*vals), vals


def _packed_desc(q, k, v, causal, softmax_scale):
    for t in (q, k, v):
        if t.dim() != 3 or t.stride(-1) != 1 or not t.is_cuda:
            raise ValueError("expected packed CUDA tensors [total, heads, d] with a contiguous last dimension")
    H, d = q.shape[1], q.shape[2]
    Hkv = k.shape[1]
    if k.shape[2] != d or tuple(v.shape) != tuple(k.shape) or H % Hkv != 0:
        raise ValueError("q/k/v shape mismatch")
    lay = lambda t: Layout(0, t.stride(1), t.stride(0))
    return AttnDesc(B=1, H=H, Hkv=Hkv, Nq=0, Nkv=0, d=d, dtype=_dtype_code(q), flags=FLAG_CAUSAL if causal else 0,
                    scale=float(softmax_scale or 0.0), q=lay(q), k=lay(k), v=lay(v)), lay


def attention_varlen_forward(q, k, v, cu_seqlens_q, cu_seqlens_k, causal=False, softmax_scale=None):
    """q [total_q, H, d], k/v [total_k, Hkv, d]; sequence i = rows cu_seqlens[i] .. cu_seqlens[i+1]-1.
    -> O [total_q, H, d], L [H, total_q]."""
    desc, lay = _packed_desc(q, k, v, causal, softmax_scale)
    cq, vq = _cu_array(cu_seqlens_q)
    ck, vk = _cu_array(cu_seqlens_k)
    if len(vq) != len(vk) or vq[-1] != q.shape[0] or vk[-1] != k.shape[0]:
        raise ValueError("cu_seqlens do not match the packed tensors")
    O = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    L = torch.empty(q.shape[1], q.shape[0], dtype=torch.float32, device=q.device)
    desc.o = lay(O)
    with torch.cuda.device(q.device):
        check(lib.b200_fa2_fwd_varlen(desc, len(vq) - 1, cq, ck, q.data_ptr(), k.data_ptr(), v.data_ptr(), O.data_ptr(),
                                      L.data_ptr(), _stream()), "b200_fa2_fwd_varlen")
    return O, L


def attention_varlen_backward(q, k, v, O, L, dO, cu_seqlens_q, cu_seqlens_k, causal=False, softmax_scale=None):
    desc, lay = _packed_desc(q, k, v, causal, softmax_scale)
    cq, vq = _cu_array(cu_seqlens_q)
    ck, vk = _cu_array(cu_seqlens_k)
    dO, O = _dense(dO), _dense(O)
    dQ = torch.empty(q.shape, dtype=q.dtype, device=q.device)
    dK = torch.empty(k.shape, dtype=k.dtype, device=k.device)
    dV = torch.empty(v.shape, dtype=v.dtype, device=v.device)
    desc.o, desc.d_o, desc.dq, desc.dk, desc.dv = lay(O), lay(dO), lay(dQ), lay(dK), lay(dV)
    with torch.cuda.device(q.device):
        check(lib.b200_fa2_bwd_varlen(desc, len(vq) - 1, cq, ck, q.data_ptr(), k.data_ptr(), v.data_ptr(), O.data_ptr(),
                                      L.data_ptr(), dO.data_ptr(), dQ.data_ptr(), dK.data_ptr(), dV.data_ptr(), _stream()),
              "b200_fa2_bwd_varlen")
    return dQ, dK, dV


class _FlashAttnVarlenFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, cu_q, cu_k, causal, softmax_scale):
        O, L = attention_varlen_forward(q, k, v, cu_q, cu_k, causal, softmax_scale)
        ctx.save_for_backward(q, k, v, O, L)
        ctx.cfg = (list(cu_q.tolist() if hasattr(cu_q, "tolist") else cu_q), list(cu_k.tolist() if hasattr(cu_k, "tolist") else cu_k),
                   causal, softmax_scale)
        return O

    @staticmethod
    def backward(ctx, dO):
        q, k, v, O, L = ctx.saved_tensors
        cu_q, cu_k, causal, softmax_scale = ctx.cfg
        dQ, dK, dV = attention_varlen_backward(q, k, v, O, L, dO, cu_q, cu_k, causal, softmax_scale)
        return dQ, dK, dV, None, None, None, None


def flash_attn_varlen_func(q, k, v, cu_seqlens_q, cu_seqlens_k, causal=False, softmax_scale=None):
    """Differentiable attention over a packed batch of variable-length sequences (cu_seqlens on the host or any tensor)."""
    return _FlashAttnVarlenFn.apply(q, k, v, cu_seqlens_q, cu_seqlens_k, causal, softmax_scale)
