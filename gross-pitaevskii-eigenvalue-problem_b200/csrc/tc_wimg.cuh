// Weight images of the tensor-core path: layout, tf32 split and the packing loop -- shared by the tensor-path kernels
// (jet_tc_kernels.cuh) and the one pack launch of the C ABI (capi.cu).
#pragma once
#include "jet_host.h"

namespace gpe {
namespace tcp {

constexpr int kImgFloats = 64 * 128;               // one half-matrix image (hi or lo)

// byte offset of element (r, c) of a [rows x 128] matrix image: 8-row x 16-byte core matrices, cores of one
// row group contiguous along c (LBO = 128 B), row groups 4096 B apart (SBO)
__host__ __device__ __forceinline__ uint32_t img_off128(int r, int c) {
    return (uint32_t)((r >> 3) * 4096 + (c >> 2) * 128 + (r & 7) * 16 + (c & 3) * 4);
}

// RN (ties away) to tf32: what `cvt.rna.tf32.f32` computes -- ptxas emulates that instruction on sm_100a with four
// instructions (add, inf test, select, mask; measured in the round-1 SASS: 20 of the 82 instructions per (point, feature)
// of the forward epilogue); the add + mask below is the same rounding for every finite value below 2^127.
__device__ __forceinline__ uint32_t tf32_hi(float v) { return (__float_as_uint(v) + 0x1000u) & 0xffffe000u; }

// hidden weight images for the tensor path.  transpose = 0: B[n = out][k = in] (forward);
// transpose = 1: B[n = in][k = out] (reverse sweep).  Zero padded to 128 x 128.
__device__ __forceinline__ void pack_wimg_elems(const PtrArray& W, int L, int Wd, int transpose, float* __restrict__ wimg, long long first,
                                                long long stride) {
    const long long total = (long long)(L - 1) * 128 * 128;
    for (long long idx = first; idx < total; idx += stride) {
        const int l = (int)(idx / (128 * 128)), rem = (int)(idx % (128 * 128));
        const int nn = rem / 128, k = rem % 128;
        const int o = transpose ? k : nn, i = transpose ? nn : k;
        const float v = (o < Wd && i < Wd) ? W.p[l + 1][(size_t)o * Wd + i] : 0.f;
        const float hi = __uint_as_float(tf32_hi(v));
        float* base = wimg + ((size_t)l * 2 + (nn >> 6)) * 2 * kImgFloats;
        const uint32_t off = img_off128(nn & 63, k) >> 2;
        base[off] = hi;
        base[kImgFloats + off] = v - hi;
    }
}

}  // namespace tcp
}  // namespace gpe
