// tmap.h -- host-side construction of TMA tensor maps (cuTensorMapEncodeTiled) without linking
// libcuda: the driver entry point is fetched through the runtime.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            return nullptr;
        fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// 3-D map over a [slabs, rows, cols] row-major tensor (cols contiguous), 128-byte swizzle.
// box = {box_cols (bytes must be 128), box_rows, 1}.  Out-of-bounds rows/cols are zero-filled on
// load and clipped on store, per slab -- which is what lets ragged N share the full-tile code.
inline int make_tmap_3d(CUtensorMap* out, const void* base, CUtensorMapDataType dt, int elem_bytes,
                        uint64_t slabs, uint64_t rows, uint64_t cols, uint32_t box_rows, uint32_t box_cols,
                        CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) return -1;
    cuuint64_t gdim[3] = {cols, rows, slabs};
    cuuint64_t gstr[2] = {cols * (uint64_t)elem_bytes, rows * cols * (uint64_t)elem_bytes};
    cuuint32_t box[3] = {box_cols, box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(out, dt, 3, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

// 4-D map over a [B, H, rows, cols] tensor with arbitrary (16-byte aligned) batch / head / row strides, cols contiguous:
// one description for the reference's contiguous [B,H,N,d], token-major [B,N,H,d], sub-blocks of rows (ring steps) and
// grouped K/V heads.  Strides in ELEMENTS.  box = {box_cols, box_rows, 1, 1}; OOB rows/cols zero-fill / clip per (b,h).
inline int make_tmap_4d(CUtensorMap* out, const void* base, CUtensorMapDataType dt, int elem_bytes,
                        uint64_t B, uint64_t H, uint64_t rows, uint64_t cols, uint64_t sb, uint64_t sh, uint64_t sn,
                        uint32_t box_rows, uint32_t box_cols, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
    PFN_encodeTiled enc = get_encode_tiled();
    if (!enc) return -1;
    const uint64_t eb = (uint64_t)elem_bytes;
    cuuint64_t gdim[4] = {cols, rows, H, B};
    // a dimension of extent 1 is never stepped over: give it any legal stride
    cuuint64_t gstr[3] = {sn * eb, (H > 1 ? sh : rows * sn) * eb, (B > 1 ? sb : H * rows * sn) * eb};
    for (int i = 0; i < 3; ++i) if (gstr[i] == 0 || (gstr[i] & 15)) return -2;
    cuuint32_t box[4] = {box_cols, box_rows, 1, 1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = enc(out, dt, 4, const_cast<void*>(base), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

}  // namespace b200
