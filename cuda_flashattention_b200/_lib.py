"""ctypes binding of libb200attn.so (the C-ABI in include/b200attn.h).

The library is built in-tree by ``make -C cuda_flashattention_b200`` (or ``__graft_entry__.build()``).
There is no fallback of any kind: if the shared object is missing, importing this module raises.
"""
import ctypes as C
import os

# torch first: it loads its bundled libnccl.so.2 (2.28.x); libb200attn.so then binds to that copy
# instead of pulling the older system libnccl in ahead of torch (same SONAME).
import torch  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200ATTN_LIB: development override (tools/kbench reads the same variable) to load a side build -- libb200attn_exp.so,
# the product plus the experimental kernel variants, or libb200attn_trace.so -- instead of the product library.
LIB_PATH = os.environ.get("B200ATTN_LIB") or os.path.join(_HERE, "libb200attn.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (nvcc, sm_100a). "
        "cuda_flashattention_b200 has no CPU or PyTorch fallback.")

lib = C.CDLL(LIB_PATH)

_vp, _fp, _i, _f, _sz = C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_size_t



class Layout(C.Structure):
    """b200_layout: batch / head / row strides in elements (all zero = contiguous [B,H,N,d])."""
    _fields_ = [("batch_stride", C.c_longlong), ("head_stride", C.c_longlong), ("row_stride", C.c_longlong)]


class AttnDesc(C.Structure):
    """b200_attn_desc (include/b200attn.h)."""
    _fields_ = [("B", _i), ("H", _i), ("Hkv", _i), ("Nq", _i), ("Nkv", _i), ("d", _i), ("dtype", _i), ("flags", _i),
                ("scale", _f),
                ("q", Layout), ("k", Layout), ("v", Layout), ("o", Layout), ("d_o", Layout), ("dq", Layout), ("dk", Layout),
                ("dv", Layout)]


_dp = C.POINTER(AttnDesc)

_SIGS = {
    "b200_strerror": (C.c_char_p, [_i]),
    "b200_last_cuda_error": (_i, []),
    "b200_version": (_i, []),
    "b200_debug_force_generic": (_i, [_i]),
    "b200_debug_launch_count": (C.c_ulonglong, []),
    "b200_fwd_uses_tensor_cores": (_i, [_i, _i, _i, _i]),
    "b200_bwd_uses_tensor_cores": (_i, [_i, _i, _i, _i]),
    "b200_fa2_fwd_route": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _f, _i, _i]),
    "b200_fa2_bwd_route": (_i, [_vp] * 7 + [_i, _i, _i, _f, _i, _i]),
    "b200_fa2_fwd_f32": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _f, _vp]),
    "b200_fa2_bwd_f32": (_i, [_vp] * 9 + [_i, _i, _f, _vp]),
    "b200_fa1_fwd_f32": (_i, [_vp] * 6 + [_i, _i, _vp]),
    "b200_fa2_fwd": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_fwd_ex": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _i, _f, _i, _i, _vp]),
    "b200_fa2_bwd_ex": (_i, [_vp] * 9 + [_i, _i, _i, _i, _i, _f, _i, _i, _vp, _sz, _vp]),
    "b200_fa2_fwd_desc": (_i, [_dp, _vp, _vp, _vp, _vp, _fp, _vp]),
    "b200_fa2_bwd_desc": (_i, [_dp] + [_vp] * 9 + [_vp, _sz, _vp]),
    "b200_fa2_fwd_varlen": (_i, [_dp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _fp, _vp]),
    "b200_fa2_bwd_varlen": (_i, [_dp, _i, _vp, _vp] + [_vp] * 9 + [_vp]),
    "b200_fa2_fwd_host": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_fwd_bwd_host": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b200_fa2_bwd": (_i, [_vp] * 9 + [_i, _i, _i, _i, _i, _f, _i, _vp, _sz, _vp]),
    "b200_fa2_bwd_static": (_i, [_vp] * 9 + [_i, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_rowdot": (_i, [_vp, _vp, _fp, _i, _i, _i, _i, _i, _vp]),
    "b200_lse_merge": (_i, [_vp, _fp, _vp, _fp, _i, _i, _i, _i, _i, _vp]),
    "b200_ring_kv_block": (_i, [_i, _i, _i]),
    "b200_ring_fwd": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _f, _i, _vp, _i, _i, _vp]),
    "b200_ring_fwd_ex": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _f, _i, _i, _vp, _i, _i, _vp]),
    "b200_ring_bwd_ex": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _i, _vp, _i, _i, _vp]),
    "b200_ring_zigzag_chunk": (_i, [_i, _i, _i]),
    "b200_ring_bwd": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _vp, _i, _i, _vp]),
    "b200_ring_set_transport": (_i, [_i]),
    "b200_ring_set_forward_mode": (_i, [_i]),
    "b200_ring_set_backward_mode": (_i, [_i]),
    "b200_ring_fwd_bwd_host": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _i, _vp, _i, _i, _vp]),
    "b200_ring_unique_id": (_i, [_vp]),
    "b200_ring_comm_init": (_i, [C.POINTER(C.c_void_p), _vp, _i, _i]),
    "b200_ring_comm_destroy": (_i, [_vp]),
    "b200_ring_exchange_kv": (_i, [_vp, _vp, _vp, _vp, _sz, _i, _i, _vp, _vp]),
    "b200_release_workspace": (_i, []),
}
for _name, (_res, _args) in _SIGS.items():
    _fn = getattr(lib, _name)          # AttributeError here == header/library mismatch
    _fn.restype = _res
    _fn.argtypes = _args

EXPORTS = tuple(_SIGS)

B200_DTYPE_BF16 = 0
B200_DTYPE_F32 = 1
B200_DTYPE_FP8_E4M3 = 2
B200_STATIC_SCRATCH_BYTES = 32 << 20     # include/b200attn.h


class B200Error(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = lib.b200_strerror(status).decode()
        if status == -3:
            msg += f" [cudaError {lib.b200_last_cuda_error()}]"
        super().__init__(f"{where}: {msg} (status {status})")


def check(status, where):
    if status != 0:
        raise B200Error(status, where)
