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

// Correction products of the fp32-accurate split.  x = hi + lo with hi = tf32(x):  a w ~= a_hi w_hi + (a_lo w_hi + a_hi w_lo).
// bf16 corrections: the bracket, 2^-11 of the result, runs as kind::f16 MMAs on bf16 operands (K = 16 per MMA, i.e. half
// the MMAs of a tf32 product; 8 mantissa bits are plenty for a term that small): 8 MMAs per 32-feature chunk instead of 12.
// Otherwise all three products run in tf32 (the round-1 scheme).
// Per kernel family (A/B builds: csrc/build_variant.sh tag -DGPE_TC_BF16_FWD=0 ...).  Measured on B200 against the fp64 oracle
// (scripts/grad_err.py / scripts/debias_sweep.py: gradient of the GPE2D golden, bar 1e-5) and at 4M points (scripts/power_probe.py):
//   forward sweeps   bf16: stash forward 9.1 -> 7.9 ms.  Gradient error 7.8e-6 (tf32) / 1.19e-5 (bf16) with the raw accumulators,
//                    1.0e-6 / 1.4e-6 with the de-biased epilogue (kTcDebiasFwd below): what the residual of the loss amplifies
//                    is the tensor core's truncation bias, not the width of the correction products                -> default 1
//   reverse sweeps   bf16: 13.6 -> 12.4 ms, gradient error unchanged to 2 %                                        -> default 1
//   dW contraction   bf16: 10.4 -> 10.2 ms (it is bound by its operand streams, not by the MMAs)                   -> default 0
#ifndef GPE_TC_BF16_FWD
#define GPE_TC_BF16_FWD 1
#endif
#ifndef GPE_TC_BF16_REV
#define GPE_TC_BF16_REV 1
#endif
#ifndef GPE_TC_BF16_DW
#define GPE_TC_BF16_DW 0
#endif
constexpr bool kBfFwd = GPE_TC_BF16_FWD != 0, kBfRev = GPE_TC_BF16_REV != 0;
// de-bias factor of the forward epilogue in units of 2^-24 (see tc_debias, jet_launch_tc.cu): measured optimum per scheme
constexpr float kTcDebiasFwd = kBfFwd ? 8.f : 4.f;

// bf16x2 {upper = hi16, lower = lo16}, round to nearest
__device__ __forceinline__ uint32_t pack_bf16x2(float lower, float upper) {
    uint32_t d;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(upper), "f"(lower));
    return d;
}

// byte offset of (row r, k) in the correction image of one half matrix: [64 rows x 256 k] bf16, 8-row x 16-byte core
// matrices (8 k each), cores contiguous along k (LBO = 128 B), row groups 4096 B apart (SBO).  The k axis is ordered the way
// the epilogue warps write the A operand: for every group of 8 features (chunk c, warp column group s) 16 consecutive k:
//   k = 16 (4 c + s) + j      j = 0..7  : bf16(w_lo) of feature 32 c + 8 s + j      (meets bf16(a)    in TMEM columns 0..3)
//                             j = 8..15 : bf16(w)    of feature 32 c + 8 s + j - 8  (meets bf16(a_lo) in TMEM columns 4..7)
__host__ __device__ __forceinline__ uint32_t corr_off(int r, int kk) {
    return (uint32_t)((r >> 3) * 4096 + (kk >> 3) * 128 + (r & 7) * 16 + (kk & 7) * 2);
}

// hidden weight images for the tensor path.  transpose = 0: B[n = out][k = in] (forward);
// transpose = 1: B[n = in][k = out] (reverse sweep).  Zero padded to 128 x 128.  Per (layer, 64-row half): the tf32 hi image
// (kImgFloats floats) followed by the correction image of the same size (bf16 pairs, or the fp32 lo image).
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
        if (transpose ? kBfRev : kBfFwd) {
            unsigned short* corr = reinterpret_cast<unsigned short*>(base + kImgFloats);
            const int kk = 16 * (k >> 3) + (k & 7);
            corr[corr_off(nn & 63, kk) >> 1] = (unsigned short)(pack_bf16x2(v - hi, 0.f) & 0xffffu);
            corr[corr_off(nn & 63, kk + 8) >> 1] = (unsigned short)(pack_bf16x2(v, 0.f) & 0xffffu);
        } else {
            base[kImgFloats + off] = v - hi;
        }
    }
}

}  // namespace tcp
}  // namespace gpe
