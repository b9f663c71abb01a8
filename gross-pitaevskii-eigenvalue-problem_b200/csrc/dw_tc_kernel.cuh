// Weight gradients of the tensor-core path (included by jet_tc_kernels.cuh).
//
//   hidden matrix l = 1..L-1 :  dW_l[o][i] = sum_rows zbar_l[row][o] * a_{l-1}[row][i]
//                               db_l[o]    = sum_{value rows} zbar_l[row][o]
//   input layer  l = 0       :  dW_0[o][i] = sum_points zbar_0,val[o] x_i + zbar_0,d_i[o],  db_0[o] = sum zbar_0,val[o]
//
// i.e. a contraction over rows (= points x streams) -- what loss.backward() does with the saved activations in the
// reference (NBroot cell 10:100, S/gross_pitaevskii_2D.py:376,422).  Here: 3xTF32 tensor-core contractions with
// K = 32 rows (8 points) per pipeline stage.
//   A operand: image of zbar_l^T  [M = 128 features x K rows]
//   B operand: image of a_{l-1}^T [N x K rows], N = 128 features + a ones column (bias gradient), padded to 144;
//              for l = 0: N = 16 holding x_i / ones on the right streams.
// One 16-byte core-matrix row of either image is "the 4 streams of one point for one feature", so the conversion
// threads (feature = tid & 127) build a = act(z) thread-locally and store hi / lo with one STS.128 each.
//
// Accumulation accuracy.  tcgen05 rounds the fp32 accumulator toward zero after every MMA (measured: scripts/
// tc_probe, scripts/tc_bign.py), a bias that grows with the length of the accumulate chain: one TMEM accumulator
// over a whole launch was 4-6e-5 off at 1M points.  So every CTA owns ONE layer (CTAs are split into L classes),
// accumulates only kDwGroup blocks (48 MMAs) into one of two ping-pong TMEM accumulators, and the conversion
// warps add each finished short accumulator into an fp32 shared-memory accumulator with ordinary round-to-nearest
// adds while the next group is contracted.  One RED per element and CTA at the end of the launch.
#pragma once

namespace gpe {
namespace tcp {

constexpr int kDwConvWarps = 16;
constexpr int kDwThreads = (kDwConvWarps + 1) * 32;
constexpr int kDwPts = 8;                                     // points per stage (K = 32 rows)
constexpr int kDwPpt = kDwPts * 128 / (kDwConvWarps * 32);    // points per conversion thread and item
constexpr int kDwStages = 2;
constexpr int kDwGroup = 4;                                   // blocks per short accumulate chain
constexpr int kDwN = 144;
constexpr int kDwPf = 8;                                      // L2 prefetch distance, items
constexpr uint32_t kDwZBytes = 128 * 32 * 4, kDwABytes = kDwN * 32 * 4;
constexpr uint32_t kDwStageBytes = 2 * kDwZBytes + 2 * kDwABytes;
constexpr uint32_t kDwLongBytes = 128 * kDwN * 4;             // long accumulator, [column][feature m]
__host__ __device__ constexpr size_t dw_smem_bytes(int D, int L) {
    return (size_t)kDwStages * kDwStageBytes + kDwLongBytes + ((size_t)(D + L + 1) * 128 + 4) * 4 + 128;
}
// [rows x 32] image: cores contiguous along k (LBO = 128 B), 8-row groups 1024 B apart (SBO)
__device__ __forceinline__ uint32_t dw_off(int m, int kgroup) { return (uint32_t)((m >> 3) * 1024 + kgroup * 128 + (m & 7) * 16); }

__device__ __forceinline__ void sts_split4(unsigned char* hi_img, unsigned char* lo_img, uint32_t off, const float (&v)[4]) {
    uint4 h, l;
    h.x = tf32_hi(v[0]), h.y = tf32_hi(v[1]), h.z = tf32_hi(v[2]), h.w = tf32_hi(v[3]);
    l.x = __float_as_uint(v[0] - __uint_as_float(h.x)), l.y = __float_as_uint(v[1] - __uint_as_float(h.y));
    l.z = __float_as_uint(v[2] - __uint_as_float(h.z)), l.w = __float_as_uint(v[3] - __uint_as_float(h.w));
    *reinterpret_cast<uint4*>(hi_img + off) = h;
    *reinterpret_cast<uint4*>(lo_img + off) = l;
}

// which layer a CTA works on, and its rank / count inside that class.  The input layer converts only one
// operand per block, so it gets about half the CTAs of a hidden layer.
struct DwClass { int layer, rank, count; };
__host__ __device__ inline DwClass dw_class(int cta, int n_cta, int L, int n0) {
    // the host launches at least L CTAs; n0 = CTAs of the input-layer class (host: dw_n0)
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

template <int D>
__global__ void __launch_bounds__(kDwThreads, 1)
dw_tc_kernel(TcParams P, const float* __restrict__ x, long long n, const float* __restrict__ stash,
             const float* __restrict__ zb, float* __restrict__ dWacc, float* __restrict__ small_acc, int num_blocks, int n0_cta) {
    using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before; using tc::tc_fence_after; using tc::tmem_wait_ld;
    using tc::smem_desc; using tc::instr_desc; using tc::mma_ss; using tc::mma_commit; using tc::fence_async_smem;
    using tc::tmem_ld_32x32b_x32; using tc::tmem_ld_32x32b_x4;
    constexpr int R = D + 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int L = P.n_hidden;
    const int n_small = (D + L + 1) * 128 + 4;
    float* longacc = reinterpret_cast<float*>(smem + (size_t)kDwStages * kDwStageBytes);
    float* small = longacc + 128 * kDwN;
    uint64_t* full = reinterpret_cast<uint64_t*>(small + n_small);
    uint64_t* empty = full + kDwStages;
    uint64_t* acc_full = empty + kDwStages;     // [2] short accumulator complete (tcgen05.commit)
    uint64_t* acc_free = acc_full + 2;          // [2] short accumulator drained (one arrive per conversion warp)
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(acc_free + 2);
    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const DwClass cls = dw_class((int)blockIdx.x, (int)gridDim.x, L, n0_cta);
    const int l = cls.layer;
    const int NC = l == 0 ? 16 : kDwN;          // accumulator columns of this CTA's layer

    if (tid == 0) {
        for (int s = 0; s < kDwStages; ++s) {
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
    for (int i = tid; i < n_small; i += kDwThreads) small[i] = P.small[i];
    for (int i = tid; i < 128 * kDwN; i += kDwThreads) longacc[i] = 0.f;
    // constant rows 128..143 of every B image: row 128 = indicator of the value stream, the rest zero
    for (int i = tid; i < kDwStages * 2 * 16 * 8; i += kDwThreads) {
        const int st = i / 256, rem = i % 256, lo = rem / 128, r = (rem % 128) / 8, kg = rem % 8;
        unsigned char* img = smem + (size_t)st * kDwStageBytes + 2 * kDwZBytes + (lo ? kDwABytes : 0);
        const float one = (!lo && r == 0) ? 1.f : 0.f;
        *reinterpret_cast<float4*>(img + dw_off(128 + r, kg)) = make_float4(one, 0.f, 0.f, 0.f);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* W0t = small;
    const float* b0 = small + D * 128;
    const int act = P.act;
    const int items = num_blocks > cls.rank ? (num_blocks - cls.rank + cls.count - 1) / cls.count : 0;   // blocks of this CTA
    const int groups = (items + kDwGroup - 1) / kDwGroup;

    if (w < kDwConvWarps) {
        const int f = tid & 127, ph = tid >> 7;
        const int q = w & 3, grp = w >> 2, m = 32 * q + lane;
        // add the finished short accumulator of group g into the shared-memory accumulator (round to nearest)
        auto drain = [&](int g) {
            const int buf = g & 1;
            mbar_wait(&acc_full[buf], (uint32_t)((g >> 1) & 1));
            tc_fence_after();
            const uint32_t tl = tb + ((uint32_t)(32 * q) << 16) + (uint32_t)buf * kDwN;
            if (l != 0) {
                uint32_t r[32], r4[4];
                tmem_ld_32x32b_x32(tl + 36 * grp, r);
                tmem_ld_32x32b_x4(tl + 36 * grp + 32, r4);
                tmem_wait_ld();
                float* dst = longacc + (36 * grp) * 128 + m;
#pragma unroll
                for (int c = 0; c < 32; ++c) dst[c * 128] += __uint_as_float(r[c]);
#pragma unroll
                for (int c = 0; c < 4; ++c) dst[(32 + c) * 128] += __uint_as_float(r4[c]);
            } else {
                uint32_t r4[4];
                tmem_ld_32x32b_x4(tl + 4 * grp, r4);
                tmem_wait_ld();
                float* dst = longacc + (4 * grp) * 128 + m;
#pragma unroll
                for (int c = 0; c < 4; ++c) dst[c * 128] += __uint_as_float(r4[c]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_free[buf]);
        };
        // global loads of block `i` of this CTA, issued one item ahead of their use
        auto load_item = [&](int i, float (&zv)[kDwPpt][R], float (&sv)[kDwPpt][R], float (&xv)[kDwPpt][D]) {
            const long long p0 = ((long long)cls.rank + (long long)i * cls.count) * kDwPts + ph * kDwPpt;
            const float* zl = zb + (size_t)l * n * 512 + f;
            const float* sl = stash + (size_t)(l >= 2 ? l - 2 : 0) * n * 512 + f;
#pragma unroll
            for (int j = 0; j < kDwPpt; ++j) {
                const bool valid = p0 + j < n;
                const long long pc = valid ? p0 + j : 0;       // clamp: loads stay in bounds, the mask zeroes zbar
                const float vm = valid ? 1.f : 0.f;
                const float* zp = zl + pc * 512;
#pragma unroll
                for (int s = 0; s < R; ++s) zv[j][s] = zp[s * 128] * vm;
                if (l >= 2) {
                    const float* sp = sl + pc * 512;
#pragma unroll
                    for (int s = 0; s < R; ++s) sv[j][s] = sp[s * 128];
                } else {
#pragma unroll
                    for (int k = 0; k < D; ++k) xv[j][k] = x[pc * D + k];
                }
            }
        };
        // L2 bulk prefetch of the operand blocks kDwPf items ahead (one thread): the register prefetch above then
        // waits on L2, not on HBM
        auto prefetch_item = [&](int i) {
            const long long p0 = ((long long)cls.rank + (long long)i * cls.count) * kDwPts;
            const long long left = n - p0;
            const uint32_t bytes = (uint32_t)(left < kDwPts ? left : kDwPts) * 512u * 4u;
            tc::prefetch_l2_bulk(zb + ((size_t)l * n + p0) * 512, bytes);
            if (l >= 2) tc::prefetch_l2_bulk(stash + ((size_t)(l - 2) * n + p0) * 512, bytes);
        };
        if (tid == 0)
            for (int i = 0; i < kDwPf && i < items; ++i) prefetch_item(i);
        float zv[kDwPpt][R], sv[kDwPpt][R], xv[kDwPpt][D];
        if (items > 0) load_item(0, zv, sv, xv);
        int stg = 0;
        uint32_t round = 0;   // how many times the stage ring has wrapped
        for (int it = 0; it < items; ++it) {
            unsigned char* ZH = smem + (size_t)stg * kDwStageBytes;
            unsigned char* ZL = ZH + kDwZBytes;
            unsigned char* AH = ZL + kDwZBytes;
            unsigned char* AL = AH + kDwABytes;
            float nzv[kDwPpt][R], nsv[kDwPpt][R], nxv[kDwPpt][D];
            if (tid == 0 && it + kDwPf < items) prefetch_item(it + kDwPf);
            if (it + 1 < items) load_item(it + 1, nzv, nsv, nxv);
            if (round > 0) mbar_wait(&empty[stg], (round - 1) & 1u);
#pragma unroll
            for (int j = 0; j < kDwPpt; ++j) {
                const int kg = ph * kDwPpt + j;
                float v[4];
#pragma unroll
                for (int s = 0; s < 4; ++s) v[s] = s < R ? zv[j][s] : 0.f;
                sts_split4(ZH, ZL, dw_off(f, kg), v);
                if (l >= 1) {
                    float a[R];
                    if (l >= 2) {
                        act_jet_from_stash<D, JET_LAPLACIAN>(sv[j], a, act);
                    } else {
                        float z0 = b0[f];
#pragma unroll
                        for (int k = 0; k < D; ++k) {
                            const float wk = W0t[k * 128 + f];
                            z0 = fmaf(wk, xv[j][k], z0);
                            a[1 + k] = wk;
                        }
                        a[0] = z0;
                        a[1 + D] = 0.f;
                        float st0;
                        act_jet_fwd<D, JET_LAPLACIAN>(a, act, st0);
                    }
#pragma unroll
                    for (int s = 0; s < 4; ++s) v[s] = s < R ? a[s] : 0.f;
                    sts_split4(AH, AL, dw_off(f, kg), v);
                } else if (f < 16) {
                    // input layer: B row f holds x_f on the value stream and 1 on derivative stream f (f < D); row D = ones
#pragma unroll
                    for (int s = 0; s < 4; ++s) v[s] = 0.f;
                    if (f < D) v[0] = xv[j][f], v[1 + f] = 1.f;
                    else if (f == D) v[0] = 1.f;
                    sts_split4(AH, AL, dw_off(f, kg), v);
                }
            }
            fence_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(&full[stg]);
#pragma unroll
            for (int j = 0; j < kDwPpt; ++j) {
#pragma unroll
                for (int s = 0; s < R; ++s) zv[j][s] = nzv[j][s], sv[j][s] = nsv[j][s];
#pragma unroll
                for (int k = 0; k < D; ++k) xv[j][k] = nxv[j][k];
            }
            if (++stg == kDwStages) stg = 0, ++round;
            // group g = it / kDwGroup has just been fed: the previous group's MMAs are long done -- drain it
            if ((it + 1) % kDwGroup == 0 || it + 1 == items) {
                const int g = it / kDwGroup;
                if (g >= 1) drain(g - 1);
            }
        }
        if (groups >= 1) drain(groups - 1);
        asm volatile("bar.sync 2, %0;" ::"n"(kDwConvWarps * 32) : "memory");
        // ---- shared-memory accumulator -> global (one atomic per element per CTA) ----
        if (items > 0) {
            for (int idx = tid; idx < NC * 128; idx += kDwConvWarps * 32) {
                const int c = idx >> 7, mm = idx & 127;
                const float v = longacc[idx];
                if (l != 0) {
                    if (c < 128) atomicAdd(&dWacc[((size_t)(l - 1) * 128 + mm) * 128 + c], v);
                    else if (c == 128) atomicAdd(&small_acc[(D + l) * 128 + mm], v);
                } else if (c <= D) {
                    atomicAdd(&small_acc[c * 128 + mm], v);
                }
            }
        }
    } else if (lane == 0) {
        const uint32_t idesc = l == 0 ? instr_desc(2, 128, 16, 0, 0) : instr_desc(2, 128, kDwN, 0, 0);
        int stg = 0;
        uint32_t round = 0;
        for (int it = 0; it < items; ++it) {
            const int g = it / kDwGroup, buf = g & 1;
            const bool first = it % kDwGroup == 0;
            mbar_wait(&full[stg], round & 1u);
            if (first && g >= 2) mbar_wait(&acc_free[buf], (uint32_t)(((g >> 1) - 1) & 1));
            tc_fence_after();
            const uint32_t zh = smem_u32(smem + (size_t)stg * kDwStageBytes), zl = zh + kDwZBytes, ah = zl + kDwZBytes, al = ah + kDwABytes;
            const uint32_t dcol = tb + (uint32_t)buf * kDwN;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint64_t azh = smem_desc(zh + j * 256, 128, 1024), azl = smem_desc(zl + j * 256, 128, 1024);
                const uint64_t bah = smem_desc(ah + j * 256, 128, 1024), bal = smem_desc(al + j * 256, 128, 1024);
                mma_ss<0, 1>(dcol, azl, bah, idesc, (first && j == 0) ? 0u : 1u);
                mma_ss<0, 1>(dcol, azh, bal, idesc, 1u);
                mma_ss<0, 1>(dcol, azh, bah, idesc, 1u);
            }
            mma_commit(&empty[stg]);
            if ((it + 1) % kDwGroup == 0 || it + 1 == items) mma_commit(&acc_full[buf]);
            if (++stg == kDwStages) stg = 0, ++round;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (w == kDwConvWarps) tmem_dealloc<1>(tb, 512);
}

}  // namespace tcp
}  // namespace gpe
