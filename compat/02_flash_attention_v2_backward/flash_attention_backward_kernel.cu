// Source-compat shim: same path as the reference kernel file its driver #includes textually;
// the kernels themselves live in libb200attn.so.  See compat/b200_compat.h.
#pragma once
#include "../util/cuda_helper.h"
#include "../b200_compat.h"
