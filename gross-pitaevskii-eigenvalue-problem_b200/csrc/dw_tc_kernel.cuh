// Weight gradients of the tensor-core path (included by jet_tc_kernels.cuh).
//
//   hidden matrix l = 1..L-1 :  dW_l[o][i] = sum_rows zbar_l[row][o] * a_{l-1}[row][i]
//                               db_l[o]    = sum_{value rows} zbar_l[row][o]
//   input layer  l = 0       :  dW_0[o][i] = sum_points zbar_0,val[o] x_i + zbar_0,d_i[o],  db_0[o] = sum zbar_0,val[o]
//
// i.e. a contraction over rows (= points x streams) -- what loss.backward() does with the saved activations in the
// reference (NBroot cell 10:100, S/gross_pitaevskii_2D.py:376,422).  Here: 3xTF32 tensor-core contractions with
// K = 32 rows (one block of 8 points) per pipeline stage, fed by TMA.
//
// Operand blocks in global memory.  The stash forward and the reverse sweep write st_l (pre-activation jets) and
// zbar_l in "image blocks": 16 KB per (layer, block of 8 points), laid out exactly like the SWIZZLE_NONE K-major
// shared-memory operand image [128 features x 32 rows] (blk_off below), where one 16-byte core-matrix row is "the 4
// streams of one point for one feature".  One cp.async.bulk per operand and block lands the raw fp32 image in shared
// memory; no thread ever computes an address per element.
//   A operand (M = 128 features o, K): zbar_l^T.  The 16 conversion warps (4 per TMEM lane quadrant, thread = feature,
//              2 points each) read the raw image, split it hi/lo and write it to TENSOR MEMORY (tcgen05.st), so the
//              MMA reads it from TMEM and shared-memory bandwidth is spent on the B operand only.
//   B operand (N = 128 features i, K): a_{l-1}^T.  The same warps turn the raw stash image into post-activation jets
//              in place (hi) and write the lo image next to it.  For l = 0: a 16-row image of x_i / ones, built from x.
//
// Accumulation accuracy.  tcgen05 rounds the fp32 accumulator toward zero after every MMA (measured: scripts/
// tc_probe, scripts/tc_bign.py), a bias that grows with the length of the accumulate chain.  So every CTA owns ONE
// layer (CTAs are split into L classes), accumulates only kDwGroup blocks (48 MMAs) into one of two ping-pong TMEM
// accumulators, and the 16 conversion warps add each finished short accumulator into fp32 REGISTERS (32 per thread,
// round-to-nearest adds) while the next group is contracted.  One vector RED per 4 elements and CTA at the end.
// Bias gradients: compensated per-thread sums of the value-stream elements seen during the zbar conversion.
#pragma once

namespace gpe {
namespace tcp {

constexpr int kDwConvWarps = 16;                              // 4 per TMEM lane quadrant
constexpr int kDwThreads = (kDwConvWarps + 2) * 32;           // + MMA warp + TMA producer warp
constexpr int kDwPts = 8;                                     // points per block (K = 32 rows)
constexpr int kDwStages = 4;
constexpr int kDwGroup = 4;                                   // blocks per short accumulate chain
constexpr int kDwPf = 0;                                      // L2 prefetch distance, blocks (0: the 4-stage TMA ring alone is best, measured)
constexpr uint32_t kDwImgBytes = 128 * 32 * 4;                // one operand image block
constexpr int kBlkFloats = 128 * 32;
constexpr uint32_t kDwStageBytes = 3 * kDwImgBytes;           // raw zbar | a hi (the raw stash lands here) | a lo
constexpr uint32_t kDwColAcc = 0, kDwColZ = 256;              // TMEM: 2 x 128 accumulator columns, 4 x (32 hi + 32 lo)
__host__ __device__ constexpr size_t dw_smem_bytes() { return (size_t)kDwStages * kDwStageBytes + 256; }

// byte offset of (feature f, point-in-block p) inside an image block; the 4 streams follow contiguously.
// cores contiguous along k (LBO = 128 B), 8-feature groups 1024 B apart (SBO)
__host__ __device__ __forceinline__ uint32_t blk_off(int f, int p) { return (uint32_t)((f >> 3) * 1024 + p * 128 + (f & 7) * 16); }

__device__ __forceinline__ void split4(const float4& v, uint4& h, uint4& l) {
    h.x = tf32_hi(v.x), h.y = tf32_hi(v.y), h.z = tf32_hi(v.z), h.w = tf32_hi(v.w);
    l.x = __float_as_uint(v.x - __uint_as_float(h.x)), l.y = __float_as_uint(v.y - __uint_as_float(h.y));
    l.z = __float_as_uint(v.z - __uint_as_float(h.z)), l.w = __float_as_uint(v.w - __uint_as_float(h.w));
}
__device__ __forceinline__ void red_add_v4(float* g, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(g), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// which layer a CTA works on, and its rank / count inside that class
struct DwClass { int layer, rank, count; };
__host__ __device__ inline DwClass dw_class(int cta, int n_cta, int L, int n0) {
    // the host launches at least L CTAs; n0 = CTAs of the input-layer class (host: bwd_tc)
    if (n0 < 1) n0 = 1;
    if (n0 > n_cta - (L - 1)) n0 = n_cta - (L - 1);
    if (cta < n0) return DwClass{0, cta, n0};
    const int rest = n_cta - n0, per = rest / (L - 1), extra = rest % (L - 1);
    int c = cta - n0, l = 1;
    for (; l < L; ++l) {
        const int cnt = per + (l <= extra ? 1 : 0);
        if (c < cnt) return DwClass{l, c, cnt};
        c -= cnt;
    }
    return DwClass{L - 1, 0, 1};
}

// stash / zb: [layer][block][kBlkFloats], lstride = floats per layer.  stash[l] = st_l, zb[l] = zbar_l.
template <int D>
__global__ void __launch_bounds__(kDwThreads, 1)
dw_tc_kernel(int L, int act, const float* __restrict__ x, long long n, long long lstride, const float* __restrict__ stash,
             const float* __restrict__ zb, float* __restrict__ dWacc, float* __restrict__ small_acc, int num_blocks, int n0_cta, int pf_dist) {
    using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before; using tc::tc_fence_after; using tc::tmem_wait_ld;
    using tc::tmem_wait_st; using tc::smem_desc; using tc::instr_desc; using tc::mma_ts; using tc::mma_commit;
    using tc::fence_async_smem; using tc::tmem_ld_32x32b_x32; using tc::tmem_ld_32x32b_x16; using tc::tmem_st_32x32b_x8;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* raw_full = reinterpret_cast<uint64_t*>(smem + (size_t)kDwStages * kDwStageBytes);   // TMA bytes landed
    uint64_t* full = raw_full + kDwStages;      // operands converted (one arrive per conversion warp)
    uint64_t* empty = full + kDwStages;         // stage consumed (tcgen05.commit)
    uint64_t* acc_full = empty + kDwStages;     // [2] short accumulator complete (tcgen05.commit)
    uint64_t* acc_free = acc_full + 2;          // [2] short accumulator drained (one arrive per conversion warp)
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(acc_free + 2);
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const DwClass cls = dw_class((int)blockIdx.x, (int)gridDim.x, L, n0_cta);
    const int l = cls.layer;

    if (tid == 0) {
        for (int s = 0; s < kDwStages; ++s) {
            mbar_init(&raw_full[s], 1);
            mbar_init(&full[s], kDwConvWarps);
            mbar_init(&empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(&acc_full[b], 1);
            mbar_init(&acc_free[b], kDwConvWarps);
        }
        mbar_fence_init();
    }
    if (w == kDwConvWarps) tmem_alloc<1>(tbase_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const int items = num_blocks > cls.rank ? (num_blocks - cls.rank + cls.count - 1) / cls.count : 0;   // blocks of this CTA
    const int groups = (items + kDwGroup - 1) / kDwGroup;
    const float* zsrc = zb + (size_t)l * lstride;
    const float* ssrc = stash + (size_t)(l > 0 ? l - 1 : 0) * lstride;

    if (w < kDwConvWarps) {
        const int q = w & 3, grp = w >> 2;
        float acc[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) acc[c] = 0.f;
        float bsum = 0.f, bcomp = 0.f;          // bias gradient of feature 32 q + lane over this warp's points, compensated
        // add the finished short accumulator of group g into the register accumulator (round to nearest)
        auto drain = [&](int g) {
            const int buf = g & 1;
            mbar_wait(&acc_full[buf], (uint32_t)((g >> 1) & 1));
            tc_fence_after();
            const uint32_t tl = tb + ((uint32_t)(32 * q) << 16) + kDwColAcc + (uint32_t)buf * 128;
            if (l != 0) {
                uint32_t r[32];
                tmem_ld_32x32b_x32(tl + 32 * grp, r);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 32; ++c) acc[c] += __uint_as_float(r[c]);
            } else if (grp == 0) {
                uint32_t r[16];
                tmem_ld_32x32b_x16(tl, r);
                tmem_wait_ld();
#pragma unroll
                for (int c = 0; c < 16; ++c) acc[c] += __uint_as_float(r[c]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_free[buf]);
        };
        // input-layer class: coordinate (tid >> 3) of point (tid & 7) of block i, for the threads that own an x row
        auto load_x = [&](int i) -> float {
            const long long p = ((long long)cls.rank + (long long)i * cls.count) * kDwPts + (tid & 7);
            return (i < items && p < n) ? x[p * D + (tid >> 3)] : 0.f;
        };
        float xcur = (l == 0 && (tid >> 3) < D) ? load_x(0) : 0.f;
        for (int it = 0; it < items; ++it) {
            const int stg = it % kDwStages;
            const uint32_t par = (uint32_t)((it / kDwStages) & 1);
            unsigned char* ZR = smem + (size_t)stg * kDwStageBytes;
            unsigned char* AH = ZR + kDwImgBytes;
            unsigned char* AL = AH + kDwImgBytes;
            mbar_wait(&raw_full[stg], par);         // every warp waits: the stage (and its barrier phase) is free only now
            {
                // ---- zbar^T of feature f, points 2 grp and 2 grp + 1 -> TMEM lane f, columns 8 grp .. 8 grp + 7, hi | lo ----
                const int f = 32 * q + lane;
                const uint32_t tz = tb + ((uint32_t)(32 * q) << 16) + kDwColZ + (uint32_t)stg * 64 + 8 * grp;
                uint32_t hi[8], lo[8];
                float s2 = 0.f;
#pragma unroll
                for (int pp = 0; pp < 2; ++pp) {
                    const float4 v = *reinterpret_cast<const float4*>(ZR + blk_off(f, 2 * grp + pp));
                    uint4 h, lw;
                    split4(v, h, lw);
                    hi[4 * pp] = h.x, hi[4 * pp + 1] = h.y, hi[4 * pp + 2] = h.z, hi[4 * pp + 3] = h.w;
                    lo[4 * pp] = lw.x, lo[4 * pp + 1] = lw.y, lo[4 * pp + 2] = lw.z, lo[4 * pp + 3] = lw.w;
                    s2 += v.x;
                }
                tmem_st_32x32b_x8(tz, hi);
                tmem_st_32x32b_x8(tz + 32, lo);
                const float y = s2 - bcomp, t = bsum + y;
                bcomp = (t - bsum) - y;
                bsum = t;
            }
            if (l != 0) {
                // ---- stash image -> post-activation jets, hi in place, lo next to it ----
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int r = tid + kDwConvWarps * 32 * j;
                    const float4 sv = *reinterpret_cast<const float4*>(AH + 16 * r);
                    const float st[4] = {sv.x, sv.y, sv.z, sv.w};
                    float a[4] = {0.f, 0.f, 0.f, 0.f};
                    act_jet_from_stash<D, JET_LAPLACIAN>(st, a, act);
                    uint4 h, lw;
                    split4(make_float4(a[0], a[1], a[2], a[3]), h, lw);
                    *reinterpret_cast<uint4*>(AH + 16 * r) = h;
                    *reinterpret_cast<uint4*>(AL + 16 * r) = lw;
                }
            } else if (tid < 16 * kDwPts) {
                // ---- input layer: 16-row image, row f < D = (x_f | 1 on derivative stream f), other rows zero ----
                const int f = tid >> 3, pp = tid & 7;
                float v[4] = {0.f, 0.f, 0.f, 0.f};
                if (f < D) {
                    v[0] = xcur;
#pragma unroll
                    for (int k = 0; k < D; ++k) v[1 + k] = f == k ? 1.f : 0.f;
                    xcur = load_x(it + 1);          // one block ahead: the (L2-prefetched) load completes under this block
                }
                uint4 h, lw;
                split4(make_float4(v[0], v[1], v[2], v[3]), h, lw);
                *reinterpret_cast<uint4*>(AH + blk_off(f, pp)) = h;
                *reinterpret_cast<uint4*>(AL + blk_off(f, pp)) = lw;
            }
            fence_async_smem();
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[stg]);
            // group g = it / kDwGroup has just been fed: the previous group's MMAs are long done -- drain it
            if ((it + 1) % kDwGroup == 0 || it + 1 == items) {
                const int g = it / kDwGroup;
                if (g >= 1) drain(g - 1);
            }
        }
        if (groups >= 1) drain(groups - 1);
        // ---- register accumulators -> global (vector RED, one per 4 elements and CTA) ----
        if (items > 0) {
            const int o = 32 * q + lane;
            if (l != 0) {
                float* dst = dWacc + ((size_t)(l - 1) * 128 + o) * 128 + 32 * grp;
#pragma unroll
                for (int c = 0; c < 32; c += 4) red_add_v4(dst + c, acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
            } else if (grp == 0) {
#pragma unroll
                for (int c = 0; c < D; ++c) atomicAdd(&small_acc[c * 128 + o], acc[c]);
            }
            atomicAdd(&small_acc[(D + l) * 128 + o], bsum);
        }
    } else if (w == kDwConvWarps) {
        if (lane == 0) {
            // ---- MMA issue thread ----
            const uint32_t idesc = l == 0 ? instr_desc(2, 128, 16, 0, 0) : instr_desc(2, 128, 128, 0, 0);
            for (int it = 0; it < items; ++it) {
                const int stg = it % kDwStages, g = it / kDwGroup, buf = g & 1;
                const bool first = it % kDwGroup == 0;
                mbar_wait(&full[stg], (uint32_t)((it / kDwStages) & 1));
                if (first && g >= 2) mbar_wait(&acc_free[buf], (uint32_t)(((g >> 1) - 1) & 1));
                tc_fence_after();
                const uint32_t ah = smem_u32(smem + (size_t)stg * kDwStageBytes + kDwImgBytes), al = ah + kDwImgBytes;
                const uint32_t zh = tb + kDwColZ + (uint32_t)stg * 64, zl = zh + 32;
                const uint32_t dcol = tb + kDwColAcc + (uint32_t)buf * 128;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint64_t bah = smem_desc(ah + j * 256, 128, 1024), bal = smem_desc(al + j * 256, 128, 1024);
                    mma_ts<0, 1>(dcol, zl + 8 * j, bah, idesc, (first && j == 0) ? 0u : 1u);
                    mma_ts<0, 1>(dcol, zh + 8 * j, bal, idesc, 1u);
                    mma_ts<0, 1>(dcol, zh + 8 * j, bah, idesc, 1u);
                }
                mma_commit(&empty[stg]);
                if ((it + 1) % kDwGroup == 0 || it + 1 == items) mma_commit(&acc_full[buf]);
            }
        }
    } else if (lane == 0) {
        // ---- TMA producer: one bulk copy per operand image, L2 bulk prefetch kDwPf blocks ahead ----
        auto prefetch = [&](int i) {
            const size_t off = ((size_t)cls.rank + (size_t)i * cls.count) * kBlkFloats;
            tc::prefetch_l2_bulk(zsrc + off, kDwImgBytes);
            if (l != 0) tc::prefetch_l2_bulk(ssrc + off, kDwImgBytes);
            else asm volatile("prefetch.global.L2 [%0];" ::"l"(x + ((size_t)cls.rank + (size_t)i * cls.count) * kDwPts * D));
        };
        for (int i = 0; i < pf_dist && i < items; ++i) prefetch(i);
        for (int it = 0; it < items; ++it) {
            const int stg = it % kDwStages;
            if (it + pf_dist < items) prefetch(it + pf_dist);
            if (it >= kDwStages) mbar_wait(&empty[stg], (uint32_t)(((it / kDwStages) - 1) & 1));
            unsigned char* ZR = smem + (size_t)stg * kDwStageBytes;
            const size_t off = ((size_t)cls.rank + (size_t)it * cls.count) * kBlkFloats;
            mbar_arrive_expect_tx(&raw_full[stg], l != 0 ? 2 * kDwImgBytes : kDwImgBytes);
            tma_bulk_g2s(ZR, zsrc + off, kDwImgBytes, &raw_full[stg]);
            if (l != 0) tma_bulk_g2s(ZR + kDwImgBytes, ssrc + off, kDwImgBytes, &raw_full[stg]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (w == kDwConvWarps) tmem_dealloc<1>(tb, 512);
}

}  // namespace tcp
}  // namespace gpe
