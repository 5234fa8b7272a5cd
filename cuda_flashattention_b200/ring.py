"""Ring attention host mirror (reference: src/03_flash_attention_v2_ring/common/ring_attention_kernel.cu:143,
bootstrap src/util/nccl_utils.h:29-103).  One process per GPU; the NCCL communicator is the library's
own (ncclCommInitRank), bootstrapped by shipping the 128-byte unique id through torch.distributed
instead of MPI_Bcast."""
import ctypes as C

import torch
import torch.distributed as dist

from ._lib import lib, check
from .attention import _check_inputs, _bhnd, _dtype_code, _stream


class RingComm:
    """init_nccl_comm / cleanup_mpi_nccl equivalent."""

    def __init__(self, rank=None, nranks=None, group=None):
        self.rank = dist.get_rank(group) if rank is None else rank
        self.nranks = dist.get_world_size(group) if nranks is None else nranks
        self.handle = C.c_void_p(None)
        if self.nranks > 1:
            buf = (C.c_ubyte * 128)()
            if self.rank == 0:
                check(lib.b200_ring_unique_id(buf), "b200_ring_unique_id")
            obj = [bytes(buf)]
            dist.broadcast_object_list(obj, src=0, group=group)
            idbuf = (C.c_ubyte * 128).from_buffer_copy(obj[0])
            check(lib.b200_ring_comm_init(C.byref(self.handle), idbuf, self.rank, self.nranks), "b200_ring_comm_init")

    def close(self):
        if self.handle:
            lib.b200_ring_comm_destroy(self.handle)
            self.handle = C.c_void_p(None)


def ring_kv_block(rank, step, nranks):
    return lib.b200_ring_kv_block(rank, step, nranks)


def zigzag_rows(rank, nranks, n_total):
    """Global row indices that rank `rank` holds under the zigzag sharding of the causal ring: chunk `rank` followed by
    chunk `2P-1-rank` of the 2P chunks of n_total/(2P) rows."""
    c = n_total // (2 * nranks)
    lo, hi = lib.b200_ring_zigzag_chunk(rank, 0, nranks), lib.b200_ring_zigzag_chunk(rank, 1, nranks)
    return list(range(lo * c, (lo + 1) * c)) + list(range(hi * c, (hi + 1) * c))


def ring_attention_forward(Q_local, K_local, V_local, comm, softmax_scale=0.0, causal=False):
    """-> (O_local, L_local).  K_local/V_local are left untouched (the reference overwrites them).
    causal=True expects the zigzag sharding (zigzag_rows)."""
    _check_inputs(Q_local, K_local, V_local)
    B, H, Nl, d = _bhnd(Q_local)
    O = torch.empty_like(Q_local)
    L = torch.empty(Q_local.shape[:-1], dtype=torch.float32, device=Q_local.device)
    with torch.cuda.device(Q_local.device):
        check(lib.b200_ring_fwd_ex(Q_local.data_ptr(), K_local.data_ptr(), V_local.data_ptr(), O.data_ptr(), L.data_ptr(),
                                   B, H, Nl, d, float(softmax_scale), _dtype_code(Q_local), 1 if causal else 0,
                                   comm.handle, comm.rank, comm.nranks, _stream()), "b200_ring_fwd_ex")
    return O, L


def ring_attention_backward(Q_local, K_local, V_local, O_local, L_local, dO_local, comm, softmax_scale=0.0, causal=False):
    _check_inputs(Q_local, K_local, V_local, O_local, dO_local)
    B, H, Nl, d = _bhnd(Q_local)
    dQ, dK, dV = torch.empty_like(Q_local), torch.empty_like(K_local), torch.empty_like(V_local)
    with torch.cuda.device(Q_local.device):
        check(lib.b200_ring_bwd_ex(Q_local.data_ptr(), K_local.data_ptr(), V_local.data_ptr(), O_local.data_ptr(),
                                   L_local.data_ptr(), dO_local.data_ptr(), dQ.data_ptr(), dK.data_ptr(), dV.data_ptr(),
                                   B, H, Nl, d, float(softmax_scale), _dtype_code(Q_local), 1 if causal else 0,
                                   comm.handle, comm.rank, comm.nranks, _stream()), "b200_ring_bwd_ex")
    return dQ, dK, dV
