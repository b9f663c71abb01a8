// Weight gradients of the tensor-core path (included by jet_tc_kernels.cuh).
//
//   hidden matrix l = 1..L-1 :  dW_l[o][i] = sum_rows zbar_l[row][o] * a_{l-1}[row][i]
//                               db_l[o]    = sum_{value rows} zbar_l[row][o]
//   (the input layer's 3 x 128 numbers are summed in registers by the reverse sweep, jet_rev_tc_kernel)
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
//              in place (hi) and write the lo image next to it.  For l = 1 there is no stash: a_0 = act(W_0 x + b_0) is
//              recomputed from the 8 points of the block (64 bytes of x, also TMA-fed).
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
// stages | barriers (256 B) | x blocks (kDwStages x 64 B) | W0t, b0 ((D + 1) x 128 floats)
__host__ __device__ constexpr size_t dw_smem_bytes() { return (size_t)kDwStages * kDwStageBytes + 256 + kDwStages * 64 + 3 * 128 * 4; }

// byte offset of (feature f, point-in-block p) inside an image block; the 4 streams follow contiguously.
// cores contiguous along k (LBO = 128 B), 8-feature groups 1024 B apart (SBO)
__host__ __device__ __forceinline__ uint32_t blk_off(int f, int p) { return (uint32_t)((f >> 3) * 1024 + p * 128 + (f & 7) * 16); }

__device__ __forceinline__ void split4(const float4& v, uint4& h, uint4& l) {
    h.x = tf32_hi(v.x), h.y = tf32_hi(v.y), h.z = tf32_hi(v.z), h.w = tf32_hi(v.w);
    l.x = __float_as_uint(v.x - __uint_as_float(h.x)), l.y = __float_as_uint(v.y - __uint_as_float(h.y));
    l.z = __float_as_uint(v.z - __uint_as_float(h.z)), l.w = __float_as_uint(v.w - __uint_as_float(h.w));
}
// One 16-byte row of the B operand (a_{l-1}: the 4 streams of one point for one feature) at byte offset `off` = blk_off(f, p)
// of the stage's images: tf32 hi values into AH; into AC the correction operand -- the fp32 lo values at the same offset, or
// (GPE_TC_BF16_DW) bf16 quadruples in the k order of the A operand's correction columns: per pair of points (2 g, 2 g + 1)
// one 16-byte core row of bf16(a_lo) (meets bf16(zbar)) followed, 128 B further, by one of bf16(a) (meets bf16(zbar_lo)).
__device__ __forceinline__ void store_b_row(unsigned char* AH, unsigned char* AC, uint32_t off, const float4& v) {
    uint4 h, lw;
    split4(v, h, lw);
    *reinterpret_cast<uint4*>(AH + off) = h;
#if GPE_TC_BF16_DW
    const uint32_t odd = (off >> 7) & 1u;                       // p & 1: blk_off has p * 128 and nothing else in bits 7..9
    unsigned char* c = AC + off - odd * 128u + odd * 8u;
    *reinterpret_cast<uint2*>(c) = make_uint2(pack_bf16x2(__uint_as_float(lw.x), __uint_as_float(lw.y)), pack_bf16x2(__uint_as_float(lw.z), __uint_as_float(lw.w)));
    *reinterpret_cast<uint2*>(c + 128) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
#else
    *reinterpret_cast<uint4*>(AC + off) = lw;
#endif
}
__device__ __forceinline__ void red_add_v4(float* g, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(g), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// which hidden matrix l = 1..L-1 a CTA works on, and its rank / count inside that class
struct DwClass { int layer, rank, count; };
__host__ __device__ inline DwClass dw_class(int cta, int n_cta, int L, int n1) {
    // the host launches at least L - 1 CTAs; n1 = CTAs of class l = 1 (it reads one operand stream, not two)
    if (L == 2) return DwClass{1, cta, n_cta};
    if (n1 < 1) n1 = 1;
    if (n1 > n_cta - (L - 2)) n1 = n_cta - (L - 2);
    if (cta < n1) return DwClass{1, cta, n1};
    const int rest = n_cta - n1, per = rest / (L - 2), extra = rest % (L - 2);
    int c = cta - n1, l = 2;
    for (; l < L; ++l) {
        const int cnt = per + (l - 1 <= extra ? 1 : 0);
        if (c < cnt) return DwClass{l, c, cnt};
        c -= cnt;
    }
    return DwClass{L - 1, 0, 1};
}

// stash / zb: [layer][block][kBlkFloats]; stash_ls / lstride = floats per layer of the stash / the zbar chunk buffer.  stash[l-1] = st_l, zb[l-1] = zbar_l (l = 1..L-1).
// small: the packed small parameters (W0t[D][128] | b[L][128] | ...), small_acc: their gradient, same layout.
// RING = false: the blocks of one chunk, zbar from the chunk buffer (dw_tc_kernel).
// RING = true : all blocks of the launch; zbar of half tile h = 2 T + c comes from ring slot T % ns once the reverse-sweep
//               CTA that owns it has signalled, and the slot is handed back when its images have landed (bwd_fused_tc_kernel).
// cta_idx / n_cta: this CTA's index among the dW CTAs of the launch.
constexpr int kDwFlushGroups = 512;   // RING: register accumulators go to the fp64 accumulator every 2048 blocks (16384 points)
template <int D, int ACT, bool RING>
__device__ __forceinline__ void dw_role(int L, int act_rt, const float* __restrict__ small, const float* __restrict__ x, long long n,
                                        long long stash_ls, long long lstride, const float* __restrict__ stash,
                                        const float* __restrict__ zb, double* __restrict__ dWacc, float* __restrict__ small_acc,
                                        int num_blocks, int n1_cta, int pf_dist, int cta_idx, int n_cta, const RingCtl& ring) {
    using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before; using tc::tc_fence_after; using tc::tmem_wait_ld;
    using tc::tmem_wait_st; using tc::smem_desc; using tc::instr_desc; using tc::mma_ts; using tc::mma_commit;
    using tc::fence_async_smem; using tc::tmem_ld_32x32b_x32; using tc::tmem_st_32x32b_x8;
    extern __shared__ __align__(1024) unsigned char smem[];
    uint64_t* raw_full = reinterpret_cast<uint64_t*>(smem + (size_t)kDwStages * kDwStageBytes);   // TMA bytes landed
    uint64_t* full = raw_full + kDwStages;      // operands converted (one arrive per conversion warp)
    uint64_t* empty = full + kDwStages;         // stage consumed (tcgen05.commit)
    uint64_t* acc_full = empty + kDwStages;     // [2] short accumulator complete (tcgen05.commit)
    uint64_t* acc_free = acc_full + 2;          // [2] short accumulator drained (one arrive per conversion warp)
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(acc_free + 2);
    float* xbuf = reinterpret_cast<float*>(smem + (size_t)kDwStages * kDwStageBytes + 256);   // [stage][8 points x D]
    float* w0s = xbuf + kDwStages * 16;                                                         // W0t[D][128] | b0[128]
    const int act = ACT >= 0 ? ACT : act_rt;
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const DwClass cls = dw_class(cta_idx, n_cta, L, n1_cta);
    const int l = cls.layer;
    if (l == 1)
        for (int i = tid; i < (D + 1) * 128; i += (int)blockDim.x) w0s[i] = small[i];

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
    // items of this CTA.  Chunk mode: blocks rank, rank + count, ...  Ring mode: half tiles (4 blocks) rank, rank + count, ...;
    // only the last half tile of the point set can be partial, and its missing blocks are the trailing ones.
    int items;
    if (RING) {
        const int halves = (num_blocks + 3) / 4;
        const int mine = halves > cls.rank ? (halves - cls.rank + cls.count - 1) / cls.count : 0;
        items = 4 * mine;
        if (mine > 0) {
            const int last_h = cls.rank + (mine - 1) * cls.count;
            const int rem = num_blocks - 4 * last_h;
            items -= 4 - (rem < 4 ? rem : 4);
        }
    } else {
        items = num_blocks > cls.rank ? (num_blocks - cls.rank + cls.count - 1) / cls.count : 0;
    }
    const int groups = (items + kDwGroup - 1) / kDwGroup;
    const float* zsrc = zb + (size_t)(l - 1) * lstride;
    const float* ssrc = stash + (size_t)(l > 1 ? l - 2 : 0) * stash_ls;
    auto item_half = [&](int i) -> int { return cls.rank + (i >> 2) * cls.count; };                       // RING: half tile of item i
    auto item_block = [&](int i) -> long long {                                                            // global block (8 points)
        return RING ? 4ll * item_half(i) + (i & 3) : (long long)cls.rank + (long long)i * cls.count;
    };
    auto zsrc_block = [&](int i) -> long long {                                                            // block inside zb
        if (!RING) return item_block(i);
        const int h = item_half(i);
        return (long long)((h >> 1) % ring.ns) * 8 + (h & 1) * 4 + (i & 3);
    };
    // block `i` of this CTA: all 8 points exist (then its 8 x D coordinates come by TMA; the one ragged block is read directly)
    auto blk_first = [&](int i) -> long long { return item_block(i) * kDwPts; };
    const bool x_aligned = (reinterpret_cast<uintptr_t>(x) & 15) == 0;   // cp.async.bulk needs a 16-byte aligned source

    if (w < kDwConvWarps) {
        const int q = w & 3, grp = w >> 2;
        float acc[32];
#pragma unroll
        for (int c = 0; c < 32; ++c) acc[c] = 0.f;
        float bsum = 0.f, bcomp = 0.f;          // bias gradient of feature 32 q + lane over this warp's points, compensated
        // add the finished short accumulator of group g into the register accumulator (round to nearest)
        // register accumulators -> the global fp64 accumulator (one RED per element): the sum over CTAs, launches and
        // flushes is exact to fp64, only the partial sums between two flushes are fp32
        auto flush = [&]() {
            double* dst = dWacc + ((size_t)(l - 1) * 128 + 32 * q + lane) * 128 + 32 * grp;
#pragma unroll
            for (int c = 0; c < 32; ++c) {
                atomicAdd(dst + c, (double)acc[c]);
                acc[c] = 0.f;
            }
        };
        auto drain = [&](int g) {
            const int buf = g & 1;
            mbar_wait(&acc_full[buf], (uint32_t)((g >> 1) & 1));
            tc_fence_after();
            const uint32_t tl = tb + ((uint32_t)(32 * q) << 16) + kDwColAcc + (uint32_t)buf * 128;
            uint32_t r[32];
            tmem_ld_32x32b_x32(tl + 32 * grp, r);
            tmem_wait_ld();
#pragma unroll
            for (int c = 0; c < 32; ++c) acc[c] += __uint_as_float(r[c]);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_free[buf]);
            if (RING && (g + 1) % kDwFlushGroups == 0) flush();
        };
        for (int it = 0; it < items; ++it) {
            const int stg = it % kDwStages;
            const uint32_t par = (uint32_t)((it / kDwStages) & 1);
            unsigned char* ZR = smem + (size_t)stg * kDwStageBytes;
            unsigned char* AH = ZR + kDwImgBytes;
            unsigned char* AL = AH + kDwImgBytes;
            mbar_wait(&raw_full[stg], par);         // every warp waits: the stage (and its barrier phase) is free only now
            if (RING && w == 0 && lane == 0 && ((it & 3) == 3 || it + 1 == items)) {
                // the last image of this half tile has landed in shared memory: hand the ring slot back
                const int h = item_half(it);
                red_release_gpu_add(ring.cons + 2 * ((h >> 1) % ring.ns) + (h & 1), 1u);
            }
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
#if GPE_TC_BF16_DW
                    // correction columns of this warp: 0..3 = bf16 pairs of the 8 values, 4..7 = bf16 pairs of their lo parts
                    lo[2 * pp] = pack_bf16x2(v.x, v.y), lo[2 * pp + 1] = pack_bf16x2(v.z, v.w);
                    lo[4 + 2 * pp] = pack_bf16x2(__uint_as_float(lw.x), __uint_as_float(lw.y));
                    lo[4 + 2 * pp + 1] = pack_bf16x2(__uint_as_float(lw.z), __uint_as_float(lw.w));
#else
                    lo[4 * pp] = lw.x, lo[4 * pp + 1] = lw.y, lo[4 * pp + 2] = lw.z, lo[4 * pp + 3] = lw.w;
#endif
                    s2 += v.x;
                }
                tmem_st_32x32b_x8(tz, hi);
                tmem_st_32x32b_x8(tz + 32, lo);
                const float y = s2 - bcomp, t = bsum + y;
                bcomp = (t - bsum) - y;
                bsum = t;
            }
            if (l != 1) {
                // ---- stash image -> post-activation jets, hi in place, lo next to it ----
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int r = tid + kDwConvWarps * 32 * j;
                    const float4 sv = *reinterpret_cast<const float4*>(AH + 16 * r);
                    const float sv4[4] = {sv.x, sv.y, sv.z, sv.w};
                    float st[D + 2], a[D + 2];
#pragma unroll
                    for (int s = 0; s < D + 2; ++s) st[s] = sv4[s];
                    tc_act_from_stash<D, ACT>(st, a, act);
                    store_b_row(AH, AL, 16u * (uint32_t)r, make_float4(a[0], a[1], a[2], D == 2 ? a[D + 1] : 0.f));
                }
            } else {
                // ---- l = 1: a_0 = act(W_0 x + b_0) of point pp, features f and f + 64, recomputed from x ----
                const int pp = (tid >> 3) & 7, f = (tid >> 6) * 8 + (tid & 7);
                const long long p0 = blk_first(it);
                float xv[D];
#pragma unroll
                for (int k = 0; k < D; ++k)
                    xv[k] = (x_aligned && p0 + kDwPts <= n) ? xbuf[stg * 16 + pp * D + k] : (p0 + pp < n ? x[(p0 + pp) * D + k] : 0.f);
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int fj = f + 64 * j;
                    float z[D + 2];
                    z[0] = w0s[D * 128 + fj];
                    z[D + 1] = 0.f;
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        const float wk = w0s[k * 128 + fj];
                        z[0] = fmaf(wk, xv[k], z[0]);
                        z[1 + k] = wk;
                    }
                    tc_act_fwd<D, ACT>(z, act);
                    store_b_row(AH, AL, blk_off(fj, pp), make_float4(z[0], z[1], z[2], D == 2 ? z[D + 1] : 0.f));
                }
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
        // ---- register accumulators -> the global fp64 accumulator (one RED per element and CTA): the sum over the ~50 CTAs
        // of a layer and the 64 launches of a 16M-point step is exact to fp64, only the per-CTA partial sums are fp32 ----
        if (items > 0) {
            flush();
            atomicAdd(&small_acc[(D + l) * 128 + 32 * q + lane], bsum);
        }
    } else if (w == kDwConvWarps) {
        if (lane == 0) {
            // ---- MMA issue thread ----
            // descriptors of the stage's two operand images are built once (the MMA thread's per-MMA descriptor arithmetic
            // was the issue bottleneck, scripts/tc_trace.py); +16 per K = 8 step, + one stage = kDwStageBytes / 16
            const uint32_t idesc = instr_desc(2, 128, 128, 0, 0);
            const uint64_t bah0 = smem_desc(smem_u32(smem + kDwImgBytes), 128, 1024), bal0 = smem_desc(smem_u32(smem + 2 * kDwImgBytes), 128, 1024);
            static_assert(kDwStages == 4 && kDwGroup == 4, "the issue loop below is unrolled over the four stages of a group");
            for (int it0 = 0; it0 < items; it0 += kDwStages) {
                const int g = it0 / kDwGroup, buf = g & 1;
                const uint32_t dcol = tb + kDwColAcc + (uint32_t)buf * 128;
                const uint32_t par = (uint32_t)((it0 / kDwStages) & 1);
#pragma unroll
                for (int stg = 0; stg < kDwStages; ++stg) {
                    const int it = it0 + stg;
                    if (it < items) {
                        mbar_wait(&full[stg], par);
                        if (stg == 0 && g >= 2) mbar_wait(&acc_free[buf], (uint32_t)(((g >> 1) - 1) & 1));
                        tc_fence_after();
                        const uint32_t zh = tb + kDwColZ + (uint32_t)stg * 64, zl = zh + 32;
                        const uint64_t bah = bah0 + (uint64_t)(stg * (kDwStageBytes / 16)), bal = bal0 + (uint64_t)(stg * (kDwStageBytes / 16));
#if GPE_TC_BF16_DW
                        constexpr uint32_t idesc16 = instr_desc(1, 128, 128, 0, 0);
#pragma unroll
                        for (int j = 0; j < 4; ++j) mma_ts<1, 1>(dcol, zl + 8 * j, bal + (uint64_t)(16 * j), idesc16, (stg == 0 && j == 0) ? 0u : 1u);
#pragma unroll
                        for (int j = 0; j < 4; ++j) mma_ts<0, 1>(dcol, zh + 8 * j, bah + (uint64_t)(16 * j), idesc, 1u);
#else
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            mma_ts<0, 1>(dcol, zl + 8 * j, bah + (uint64_t)(16 * j), idesc, (stg == 0 && j == 0) ? 0u : 1u);
                            mma_ts<0, 1>(dcol, zh + 8 * j, bal + (uint64_t)(16 * j), idesc, 1u);
                            mma_ts<0, 1>(dcol, zh + 8 * j, bah + (uint64_t)(16 * j), idesc, 1u);
                        }
#endif
                        mma_commit(&empty[stg]);
                        if (stg == kDwStages - 1 || it + 1 == items) mma_commit(&acc_full[buf]);
                    }
                }
            }
        }
    } else if (w == kDwConvWarps + 1 && lane == 0) {
        // ---- TMA producer: one bulk copy per operand image (optionally an L2 bulk prefetch pf_dist blocks ahead) ----
        auto prefetch = [&](int i) {
            if (!RING) tc::prefetch_l2_bulk(zsrc + (size_t)zsrc_block(i) * kBlkFloats, kDwImgBytes);
            if (l != 1) tc::prefetch_l2_bulk(ssrc + (size_t)item_block(i) * kBlkFloats, kDwImgBytes);
        };
        for (int i = 0; i < pf_dist && i < items; ++i) prefetch(i);
        for (int it = 0; it < items; ++it) {
            const int stg = it % kDwStages;
            if (pf_dist > 0 && it + pf_dist < items) prefetch(it + pf_dist);
            if (it >= kDwStages) mbar_wait(&empty[stg], (uint32_t)(((it / kDwStages) - 1) & 1));
            if (RING && (it & 3) == 0) {
                // first block of a half tile: wait until all kEpiWarps warps of the reverse CTA that owns it have released
                // zbar_l of that tile, then order the async-proxy reads below after this generic-proxy acquire
                const int h = item_half(it), T = h >> 1;
                ring_wait(ring.prod + (size_t)(l - 1) * 2 * ring.nrc + 2 * (T % ring.nrc) + (h & 1),
                          (unsigned int)(kEpiWarps * (T / ring.nrc + 1)), ring.abort);
                asm volatile("fence.proxy.async.global;" ::: "memory");
            }
            unsigned char* ZR = smem + (size_t)stg * kDwStageBytes;
            const long long p0 = blk_first(it);
            const bool xtma = l == 1 && x_aligned && p0 + kDwPts <= n;
            mbar_arrive_expect_tx(&raw_full[stg], (l != 1 ? 2 * kDwImgBytes : kDwImgBytes) + (xtma ? kDwPts * D * 4u : 0u));
            tma_bulk_g2s(ZR, zsrc + (size_t)zsrc_block(it) * kBlkFloats, kDwImgBytes, &raw_full[stg]);
            if (l != 1) tma_bulk_g2s(ZR + kDwImgBytes, ssrc + (size_t)item_block(it) * kBlkFloats, kDwImgBytes, &raw_full[stg]);
            else if (xtma) tma_bulk_g2s(xbuf + stg * 16, x + p0 * D, kDwPts * D * 4u, &raw_full[stg]);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (w == kDwConvWarps) tmem_dealloc<1>(tb, 512);
}

template <int D, int ACT = -1>
__global__ void __launch_bounds__(kDwThreads, 1)
dw_tc_kernel(int L, int act_rt, const float* __restrict__ small, const float* __restrict__ x, long long n, long long stash_ls, long long lstride,
             const float* __restrict__ stash, const float* __restrict__ zb, double* __restrict__ dWacc, float* __restrict__ small_acc,
             int num_blocks, int n1_cta, int pf_dist) {
    dw_role<D, ACT, false>(L, act_rt, small, x, n, stash_ls, lstride, stash, zb, dWacc, small_acc, num_blocks, n1_cta, pf_dist, (int)blockIdx.x,
                           (int)gridDim.x, RingCtl{});
}

// ------------------------------------------------------------------------------------------
// Fused backward: one persistent launch over ALL points.  Clusters [0, nrc) are reverse-sweep CTA pairs (cta_group::2),
// every CTA of the remaining clusters runs the dW contraction of one hidden matrix on its own (cta_group::1; one kernel
// holding both group sizes was checked with scripts/tc_probe, test 5).  The grid never exceeds one CTA per SM, so all
// CTAs are co-resident and the counters of RingCtl cannot deadlock: the lowest unconsumed tile is always producible.
// Needs the resident stash of all points (the forward wrote it).  HBM traffic: the stash is read once by the sweep and
// (layers >= 2) once by dW; zbar stays in L2.
// ------------------------------------------------------------------------------------------
constexpr int kFusedThreads = kDwThreads > kTcThreads ? kDwThreads : kTcThreads;
template <int D, int ACT = -1, bool TWO = true>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kFusedThreads, 1)
bwd_fused_tc_kernel(TcParams Pr, const float* __restrict__ x, long long n, const float* __restrict__ u_bar,
                    const float* __restrict__ g_bar, const float* __restrict__ h_bar, const float* __restrict__ stash,
                    float* __restrict__ zb, double* __restrict__ dWacc, float* __restrict__ small_acc, long long stash_ls, int nblk8,
                    int num_pairs, int pf_dist, int n1_cta, RingCtl ring) {
    const int cluster_id = (int)(blockIdx.x >> 1);
    if (cluster_id < ring.nrc) {
        if (TWO) rev2_role<D, ACT, true>(Pr, x, n, u_bar, g_bar, h_bar, stash, zb, small_acc, stash_ls, nblk8, num_pairs, cluster_id, ring.nrc, ring);
        else rev_role<D, ACT, true>(Pr, x, n, u_bar, g_bar, h_bar, stash, zb, small_acc, stash_ls, nblk8, num_pairs, pf_dist, cluster_id, ring.nrc, ring);
    } else {
        dw_role<D, ACT, true>(Pr.n_hidden, Pr.act, Pr.small, x, n, stash_ls, (long long)ring.ns * 8 * kBlkFloats, stash, zb, dWacc, small_acc, nblk8,
                              n1_cta, 0, (int)blockIdx.x - 2 * ring.nrc, (int)gridDim.x - 2 * ring.nrc, ring);
    }
}

}  // namespace tcp
}  // namespace gpe
