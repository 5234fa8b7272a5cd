// Source-compat shim for src/03_flash_attention_v2_ring/common/flash_attention_kernel.cu (see compat/b200_compat.h).
#pragma once
#include "../../util/cuda_helper.h"
#include "../../util/attention_helper.h"
#include "../../b200_compat.h"
