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
LIB_PATH = os.path.join(_HERE, "libb200attn.so")

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (nvcc, sm_100a). "
        "cuda_flashattention_b200 has no CPU or PyTorch fallback.")

lib = C.CDLL(LIB_PATH)

_vp, _fp, _i, _f, _sz = C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_size_t

_SIGS = {
    "b200_strerror": (C.c_char_p, [_i]),
    "b200_last_cuda_error": (_i, []),
    "b200_version": (_i, []),
    "b200_debug_force_generic": (_i, [_i]),
    "b200_debug_launch_count": (C.c_ulonglong, []),
    "b200_fwd_uses_tensor_cores": (_i, [_i, _i, _i, _i]),
    "b200_bwd_uses_tensor_cores": (_i, [_i, _i, _i, _i]),
    "b200_fa2_fwd_f32": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _f, _vp]),
    "b200_fa2_bwd_f32": (_i, [_vp] * 9 + [_i, _i, _f, _vp]),
    "b200_fa1_fwd_f32": (_i, [_vp] * 6 + [_i, _i, _vp]),
    "b200_fa2_fwd": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_fwd_ex": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _i, _f, _i, _i, _vp]),
    "b200_fa2_bwd_ex": (_i, [_vp] * 9 + [_i, _i, _i, _i, _i, _f, _i, _i, _vp, _sz, _vp]),
    "b200_fa2_fwd_host": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_fwd_bwd_host": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _vp]),
    "b200_fa2_bwd_workspace_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b200_fa2_bwd": (_i, [_vp] * 9 + [_i, _i, _i, _i, _i, _f, _i, _vp, _sz, _vp]),
    "b200_rowdot": (_i, [_vp, _vp, _fp, _i, _i, _i, _i, _i, _vp]),
    "b200_lse_merge": (_i, [_vp, _fp, _vp, _fp, _i, _i, _i, _i, _i, _vp]),
    "b200_ring_kv_block": (_i, [_i, _i, _i]),
    "b200_ring_fwd": (_i, [_vp, _vp, _vp, _vp, _fp, _i, _i, _i, _i, _f, _i, _vp, _i, _i, _vp]),
    "b200_ring_bwd": (_i, [_vp] * 9 + [_i, _i, _i, _i, _f, _i, _vp, _i, _i, _vp]),
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
