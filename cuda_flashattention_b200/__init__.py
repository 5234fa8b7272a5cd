"""cuda_flashattention_b200 -- B200-native (sm_100a) exact attention: FlashAttention-2 forward /
backward and ring attention behind the reference's host-callable interface.

Importing this package loads libb200attn.so; it raises if the library was not built.
"""
from ._lib import lib, LIB_PATH, EXPORTS, B200Error, B200_DTYPE_BF16, B200_DTYPE_F32  # noqa: F401
from .attention import (flash_attention_2_forward, flash_attention_2_backward, flash_attention,  # noqa: F401
                        lse_merge_, uses_tensor_cores, flash_attention_2_forward_host,
                        flash_attention_2_forward_backward_host, attention_forward, attention_backward,
                        flash_attn_func, attention_varlen_forward, attention_varlen_backward, flash_attn_varlen_func)
from .ring import RingComm, ring_attention_forward, ring_attention_backward, ring_kv_block, zigzag_rows  # noqa: F401
