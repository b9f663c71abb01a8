// Tensor-core (tcgen05 / TMEM) path of the Taylor-mode MLP -- sm_100a, hidden width padded to 128.
//
// Same mathematics as jet_fwd_kernel (jet_kernels.cuh): the nn.Sequential forward plus the nested
// autograd.grad derivatives of the reference (S/gross_pitaevskii_2D.py:59-81,183-188; NBroot cell 6:30-54,
// 90-104), but the per-layer contraction  [rows x 128] . W^T  runs on the 5th-gen tensor cores.
//
// Precision: every fp32 operand is split  v = hi + lo  (hi = RN to tf32, lo = v - hi) and the product is
// formed as hi*hi + lo*hi + hi*lo with kind::tf32 MMAs accumulating in fp32 (3xTF32).  Measured on B200
// (scripts/tc_probe): max error 1.6e-5 on |D| <= 15 over K = 128, i.e. fp32-GEMM class.
//
// Work decomposition: a CTA *pair* (cluster of 2, tcgen05 cta_group::2) owns 2 x 32 collocation points.
//   rows of the M = 2 x 128 tile: TMEM lane 32q + 8s + p  <->  point 8q + p of the CTA, stream s
//                                 (s = 0 value, 1..D first derivatives, D+1 trace of second derivatives)
//   A operand (activations)     : TMEM, columns [0,128) hi and [128,256) lo, written by the epilogue warps
//   B operand (weights)         : shared memory, resident for the whole kernel; each CTA holds the 64 output
//                                 features of its half, hi and lo images in the SWIZZLE_NONE K-major
//                                 canonical layout (8 x 16 B core matrices)
//   D accumulators              : TMEM columns [256,384) and [384,512), alternating per layer
// Epilogue threads use the 16x256b TMEM fragment (thread t: rows t/4 and t/4+8, two adjacent columns), so
// with the lane map above one thread holds all streams of one point for two features: the activation jets
// are thread-local, with no shuffles and no shared-memory staging.
// The epilogue of layer l hands its output to the MMA of layer l+1 in 32-column chunks (mbarriers
// a_ready[0..3]), so tensor work trails the activation work by one chunk.
#pragma once
#include "jet_common.cuh"
#include "tc_tmem.cuh"
#include "tc_wimg.cuh"

namespace gpe {
namespace tcp {

constexpr int kEpiWarps = 16;                      // 4 per TMEM lane quadrant
constexpr int kWpq = kEpiWarps / 4;                // warps per quadrant
constexpr int kCw = 32 / kWpq;                     // columns per warp in every 32-column chunk
constexpr int kNI = kCw / 8;                       // 16x256b repetitions per warp and chunk
static_assert(kCw == 8 && kNI == 1, "the image-block slot addressing assumes one 8-column group per warp and chunk");
constexpr int kTcThreads = (kEpiWarps + 1) * 32;   // + MMA / control warp
constexpr int kTcPts = 32;                         // points per CTA tile
constexpr int kTcMaxMat = 3;                       // resident hidden matrices (64 KB each per CTA)
constexpr uint32_t kColAHi = 0, kColALo = 128, kColD0 = 256;

struct TcParams {
    int n_hidden;          // L
    int act;
    const float* wimg;     // [L-1][cta 2][hi,lo][kImgFloats]
    const float* small;    // W0t[D][128], b[L][128], wout[128], bout[4]   (same pack as the FFMA path)
    float xa[2], xb[2];    // forward sweeps only: x_eff = xa * x + xb per coordinate (mirror points of symmetry_loss; value-only calls)
    float debias;          // forward sweeps: multiplier on every accumulator read, 1 + beta 2^-24 (tc_debias, jet_launch_tc.cu)
    int interleave;        // order of the 16 hi*hi MMAs of a layer relative to its 32 correction MMAs (the accumulator is rounded
                           // toward zero after every MMA, so what lands late lands on a large accumulator):
                           //   0: all corrections first, hi*hi last -- 24 MMAs trail the last chunk, fewest roundings at size;
                           //   1: hi*hi of a chunk right after its corrections -- 12 MMAs trail and the tensor work is spread
                           //      evenly under the epilogue of the next chunk; most roundings at size.
                           // (Tried: corrections of chunks 0..2, then hi*hi of chunks 0..2 under the epilogue of chunk 3, then chunk
                           // 3 -- also 12 trailing MMAs and fewer late roundings, but 20 MMAs do not fit under one chunk's epilogue:
                           // forward 11.0 ms per 4M points against 11.7 (order 0) and 9.7 (order 1).)
};

constexpr int kDw0Stride = 12;   // floats per (row, feature) of the input-layer accumulator: 8 point slots, conflict-free banks
__host__ __device__ constexpr size_t tc_smem_base_bytes(int D, int L) {
    return (size_t)(L - 1) * 2 * kImgFloats * 4 + ((size_t)(D + L + 1) * 128 + 4) * 4 + kEpiWarps * 32 * 4 * 4 + 128;
}
// + the reverse sweep's shared-memory accumulator of the input-layer gradient, [D + 1][128][kDw0Stride]
__host__ __device__ constexpr size_t tc_smem_bytes(int D, int L) { return tc_smem_base_bytes(D, L) + (size_t)(D + 1) * 128 * kDw0Stride * 4; }

// tanh for the tensor-path epilogues: 1 - 2 / (1 + e^{2x}), one MUFU.EX2 + one MUFU.RCP, branch free, exact limits at
// +-inf.  Absolute error <= 2e-7 (two units in the last place of 1), i.e. the rounding class of the fp32 sums the
// activations feed; tanhf costs ~17 instructions with two predicated paths.
__device__ __forceinline__ float tanh_tc(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 2.8853900817779268f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.f));
    return fmaf(-2.f, r, 1.f);
}

// Activation jets of the tensor-path kernels (JET_LAPLACIAN streams: value, D first derivatives, trace).  ACT fixed at
// compile time for tanh / ShiftedTanh (no runtime branches); ACT = -1 takes the generic code of jet_common.cuh.
// z: pre-activation jets in, post-activation jets out; returns the value the reverse sweep needs (tanh(z0)).
template <int D, int ACT>
__device__ __forceinline__ float tc_act_fwd(float (&z)[D + 2], int act) {
    if (ACT == ACT_TANH || ACT == ACT_SHIFTED_TANH) {
        const float t = tanh_tc(z[0]);
        const float s1 = fmaf(-t, t, 1.f);
        const float s2 = (-2.f * t) * s1;
        float q = z[1] * z[1];
        if (D == 2) q = fmaf(z[2], z[2], q);
        z[1 + D] = fmaf(s1, z[1 + D], s2 * q);
#pragma unroll
        for (int i = 0; i < D; ++i) z[1 + i] *= s1;
        z[0] = ACT == ACT_SHIFTED_TANH ? t + 1.f : t;
        return t;
    } else {
        float st0;
        act_jet_fwd<D, JET_LAPLACIAN>(z, act, st0);
        return st0;
    }
}
// stash {t, z1.., ztrace} -> post-activation jets
template <int D, int ACT>
__device__ __forceinline__ void tc_act_from_stash(const float (&st)[D + 2], float (&a)[D + 2], int act) {
    if (ACT == ACT_TANH || ACT == ACT_SHIFTED_TANH) {
        const float t = st[0];
        const float s1 = fmaf(-t, t, 1.f);
        const float s2 = (-2.f * t) * s1;
        float q = st[1] * st[1];
        if (D == 2) q = fmaf(st[2], st[2], q);
        a[0] = ACT == ACT_SHIFTED_TANH ? t + 1.f : t;
#pragma unroll
        for (int i = 0; i < D; ++i) a[1 + i] = s1 * st[1 + i];
        a[1 + D] = fmaf(s1, st[1 + D], s2 * q);
    } else {
        act_jet_from_stash<D, JET_LAPLACIAN>(st, a, act);
    }
}
// reverse of the activation jets: ab (adjoint of the post-activation jets) -> adjoint of the pre-activation jets, in place
template <int D, int ACT>
__device__ __forceinline__ void tc_act_bwd(const float (&st)[D + 2], float (&ab)[D + 2], int act) {
    if (ACT == ACT_TANH || ACT == ACT_SHIFTED_TANH) {
        const float t = st[0];
        const float s1 = fmaf(-t, t, 1.f);
        const float s2 = (-2.f * t) * s1;
        const float s3 = (-2.f * s1) * fmaf(-3.f * t, t, 1.f);
        const float a2b = ab[1 + D];
        const float c2 = (2.f * s2) * a2b;
        float z0b = s1 * ab[0];
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const float z1 = st[1 + i];
            q = fmaf(z1, z1, q);
            z0b = fmaf(s2 * z1, ab[1 + i], z0b);
            ab[1 + i] = fmaf(s1, ab[1 + i], c2 * z1);
        }
        ab[0] = fmaf(fmaf(s3, q, s2 * st[1 + D]), a2b, z0b);
        ab[1 + D] = s1 * a2b;
    } else {
        act_jet_bwd<D, JET_LAPLACIAN>(st, ab, act);
    }
}

__global__ void pack_wimg_kernel(PtrArray W, int L, int Wd, int transpose, float* __restrict__ wimg) {
    pack_wimg_elems(W, L, Wd, transpose, wimg, (long long)blockIdx.x * blockDim.x + threadIdx.x, (long long)gridDim.x * blockDim.x);
}


// ------------------------------------------------------------------------------------------
// MMA issue (one thread).  Measured with clock64 stamps on B200 (scripts/tc_trace.py): with the descriptor built per
// MMA (shift, mask, or, R2UR -- a dependent chain of uniform-datapath instructions) one thread needs ~1000 cycles to
// issue the 12 MMAs of a 32-column chunk, more than the 774 cycles they execute, and every cluster-scope barrier wait
// adds an L1 invalidate (CCTL.IVALL).  So: the descriptors of a matrix are built once and advanced by a constant
// (+16 in the 16-byte address field = the 256 bytes of one K = 8 step), everything is unrolled, and the waits are the
// plain CTA-scope ones -- what travels between the two CTAs here is tensor memory, ordered by the tcgen05 fences.
// ------------------------------------------------------------------------------------------
// the three products of one K = 32 chunk: corrections first, then hi*hi (all read the same A columns)
// The MMAs of one 32-feature chunk of A (TMEM: ahi = 32 tf32 columns, ac = 32 correction columns) against the matching K
// slice of a weight image (bhi / bc: descriptors of the tf32 hi image / the correction image, already advanced to the chunk).
// Correction products (see tc_wimg.cuh): four kind::f16 MMAs on bf16 pairs, one per 8-feature group -- TMEM columns
// 8 s .. 8 s + 3 hold bf16(a), 8 s + 4 .. 8 s + 7 hold bf16(a_lo), the image holds bf16(w_lo) | bf16(w) in the same k order.
template <bool BF>
__device__ __forceinline__ void mma_chunk_corr(uint32_t dcol, uint32_t ahi, uint32_t ac, uint64_t bhi, uint64_t bc, uint32_t idesc,
                                               uint32_t accum_first) {
    if (BF) {
        constexpr uint32_t idesc16 = tc::instr_desc(1, 256, 128, 0, 0);
#pragma unroll
        for (int sgrp = 0; sgrp < 4; ++sgrp)
            tc::mma_ts<1, 2>(dcol, ac + 8 * sgrp, bc + (uint64_t)(16 * sgrp), idesc16, sgrp == 0 ? accum_first : 1u);
    } else {
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
            tc::mma_ts<0, 2>(dcol, ac + 8 * kk, bhi + (uint64_t)(16 * kk), idesc, kk == 0 ? accum_first : 1u);
            tc::mma_ts<0, 2>(dcol, ahi + 8 * kk, bc + (uint64_t)(16 * kk), idesc, 1u);
        }
    }
}
__device__ __forceinline__ void mma_chunk_hh(uint32_t dcol, uint32_t ahi, uint64_t bhi, uint32_t idesc, uint32_t accum_first) {
#pragma unroll
    for (int kk = 0; kk < 4; ++kk) tc::mma_ts<0, 2>(dcol, ahi + 8 * kk, bhi + (uint64_t)(16 * kk), idesc, kk == 0 ? accum_first : 1u);
}
// the small products first, then hi*hi: the accumulator is rounded toward zero after every MMA
template <bool BF>
__device__ __forceinline__ void mma_chunk_3x(uint32_t dcol, uint32_t ahi, uint32_t ac, uint64_t bhi, uint64_t bc, uint32_t idesc,
                                             uint32_t accum_first) {
#ifndef GPE_TC_PROBE_ONE_PRODUCT   // (energy probe builds issue hi*hi alone: wrong numerics)
    mma_chunk_corr<BF>(dcol, ahi, ac, bhi, bc, idesc, accum_first);
    mma_chunk_hh(dcol, ahi, bhi, idesc, 1u);
#else
    mma_chunk_hh(dcol, ahi, bhi, idesc, accum_first);
#endif
}

// ------------------------------------------------------------------------------------------
// forward: u, grad u, Laplacian per point (JET_LAPLACIAN semantics; for D = 1 the trace is u_xx)
// ------------------------------------------------------------------------------------------

template <int NI> struct FragIO;
template <> struct FragIO<1> {
    static __device__ __forceinline__ void ld(uint32_t a, uint32_t (&r)[4]) { tc::tmem_ld_16x256b_x1(a, r); }
    static __device__ __forceinline__ void st(uint32_t a, const uint32_t (&r)[4]) { tc::tmem_st_16x256b_x1(a, r); }
};
template <> struct FragIO<2> {
    static __device__ __forceinline__ void ld(uint32_t a, uint32_t (&r)[8]) { tc::tmem_ld_16x256b_x2(a, r); }
    static __device__ __forceinline__ void st(uint32_t a, const uint32_t (&r)[8]) { tc::tmem_st_16x256b_x2(a, r); }
};
// A-operand fragments of one (point, feature pair): the 4 stream slots are 4 TMEM lanes (lane half lo: slots 0, 1; lane
// half hi: slots 2, 3), the two features are adjacent columns.  Hi: tf32 values (16x256b store, 8 columns per warp);
// C: the correction columns of the same warp -- bf16 pairs {a | a_lo} (16x128b.x2 store: columns 0..3 | 4..7), or the
// fp32 lo values when BF = false.
struct AFrag { uint32_t LoHi[4], HiHi[4], LoC[4], HiC[4]; };
template <int R, bool BF>
__device__ __forceinline__ void make_afrag(const float (&a)[2][R], AFrag& f) {
#pragma unroll
    for (int s = 0; s < 4; ++s) {
        const float v0 = s < R ? a[0][s] : 0.f, v1 = s < R ? a[1][s] : 0.f;
        const uint32_t h0 = tf32_hi(v0), h1 = tf32_hi(v1);
        const float l0 = v0 - __uint_as_float(h0), l1 = v1 - __uint_as_float(h1);
        uint32_t* H = s < 2 ? f.LoHi : f.HiHi;
        uint32_t* C = s < 2 ? f.LoC : f.HiC;
        H[2 * (s & 1)] = h0, H[2 * (s & 1) + 1] = h1;
        if (BF) {
            C[s & 1] = pack_bf16x2(v0, v1);
            C[2 + (s & 1)] = pack_bf16x2(l0, l1);
        } else {
            C[2 * (s & 1)] = __float_as_uint(l0), C[2 * (s & 1) + 1] = __float_as_uint(l1);
        }
    }
}
// col_hi / col_c: this warp's 8 columns of the chunk in the tf32 area / the correction area
template <bool BF>
__device__ __forceinline__ void store_afrag(uint32_t lane_lo, uint32_t lane_hi, uint32_t col_hi, uint32_t col_c, const AFrag& f) {
    tc::tmem_st_16x256b_x1(lane_lo + col_hi, f.LoHi);
    tc::tmem_st_16x256b_x1(lane_hi + col_hi, f.HiHi);
    if (BF) {
        tc::tmem_st_16x128b_x2(lane_lo + col_c, f.LoC);
        tc::tmem_st_16x128b_x2(lane_hi + col_c, f.HiC);
    } else {
        tc::tmem_st_16x256b_x1(lane_lo + col_c, f.LoC);
        tc::tmem_st_16x256b_x1(lane_hi + col_c, f.HiC);
    }
}

// 32-byte vector access to one (point, feature pair) slot of an image block (dw_tc_kernel.cuh: blk_off)
__device__ __forceinline__ void ldg_v8(const float* g, float (&v)[8]) {
    asm volatile("ld.global.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(g));
}
__device__ __forceinline__ void stg_v8(float* g, const float (&v)[8]) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(g), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]),
                 "f"(v[5]), "f"(v[6]), "f"(v[7])
                 : "memory");
}
// float offset of the slot of (point p, features f0 = 8 g + 2 tq and f0 + 1) in a [block][128 x 32] image array
__device__ __forceinline__ size_t slot_off(long long p, int g, int tq) {
    return (size_t)(p >> 3) * 4096 + (size_t)g * 256 + (size_t)(p & 7) * 32 + 8 * tq;
}

// STASH: additionally write, for every hidden layer l >= 1, the pre-activation jets {tanh value | a0, z1.., z2} to the
// image blocks stash[l-1][point / 8] (what the reverse sweep and the dW contraction read back; nblk8 = blocks of this
// launch, stash_ls = floats per layer of the stash array -- larger than nblk8 * 4096 when the stash of ALL points is resident).
// Layer 0 is recomputed from x by both consumers.
// ACT >= 0 fixes the activation at compile time (the tanh nets of every BASELINE config: no runtime activation branches in
// the epilogue); ACT = -1 reads it from P.act.
template <int D, bool STASH, int ACT = -1>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTcThreads, 1)
jet_fwd_tc_kernel(TcParams P, const float* __restrict__ x, long long n, float* __restrict__ out_u, float* __restrict__ out_g,
                  float* __restrict__ out_h, float* __restrict__ stash, long long stash_ls, int nblk8, int num_pairs) {
    using tc::cluster_ctarank; using tc::cluster_sync; using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before;
    using tc::tc_fence_after; using tc::tmem_wait_ld; using tc::tmem_wait_st; using tc::mbar_arrive_cluster;
    using tc::mbar_wait_cluster; using tc::smem_desc; using tc::instr_desc; using tc::mma_ts; using tc::mma_commit_pair;
    using IO = FragIO<kNI>;
    constexpr int R = D + 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int L = P.n_hidden, nmat = L - 1;
    const int n_small = (D + L + 1) * 128 + 4;
    float* wimg_s = reinterpret_cast<float*>(smem);
    float* small = wimg_s + (size_t)nmat * 2 * kImgFloats;
    float* comb = small + n_small;
    uint64_t* a_ready = reinterpret_cast<uint64_t*>(comb + kEpiWarps * 32 * 4);
    uint64_t* d_full = a_ready + 4;
    uint64_t* w_full = d_full + 2;
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(w_full + 1);

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const uint32_t cta = cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

    if (tid == 0) {
        for (int c = 0; c < 4; ++c) mbar_init(&a_ready[c], 2 * kEpiWarps);
        mbar_init(&d_full[0], 1);
        mbar_init(&d_full[1], 1);
        mbar_init(w_full, 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(w_full, (uint32_t)nmat * 2 * kImgFloats * 4);
        for (int l = 0; l < nmat; ++l)
            for (int hl = 0; hl < 2; ++hl)
                tma_bulk_g2s(wimg_s + ((size_t)l * 2 + hl) * kImgFloats, P.wimg + (((size_t)l * 2 + cta) * 2 + hl) * kImgFloats,
                             kImgFloats * 4, w_full);
        mbar_wait(w_full, 0);   // both halves must be resident before the leader's first MMA (cluster_sync below)
    }
    if (w == kEpiWarps) tmem_alloc<2>(tbase_s, 512);
    for (int i = tid; i < n_small; i += kTcThreads) small[i] = P.small[i];
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* W0t = small;
    const float* bias = small + D * 128;
    const float* wout = small + (D + L) * 128;
    const float bout = small[(D + L + 1) * 128];
    const int act = ACT >= 0 ? ACT : P.act;

    if (w < kEpiWarps) {
        // Three code paths per tile -- input layer (from x, no accumulator), middle layers, last hidden layer (output dot
        // product, nothing handed on) -- so none of them carries the others' predicated-off instructions.
        const int q = w & 3, sub = w >> 2, pl = lane >> 2, tq = lane & 3;
        const uint32_t lane_lo = tb + ((uint32_t)(32 * q) << 16), lane_hi = lane_lo + (16u << 16);
        const int fb = kCw * sub + 2 * tq;                 // this thread's feature pair inside every 32-column chunk
        uint32_t dc0 = 0u, dc1 = 0u;                       // completed phases of d_full[0] / d_full[1]
        // split the four stream slots of features (f, f + 1) and hand the chunk to the MMA thread
        auto hand_over = [&](int c, const float (&a)[2][R]) {
            AFrag fr;
            make_afrag<R, kBfFwd>(a, fr);
            const uint32_t col = (uint32_t)(32 * c + kCw * sub);
            store_afrag<kBfFwd>(lane_lo, lane_hi, kColAHi + col, kColALo + col, fr);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&a_ready[c], 0);
        };
        for (int it = 0;; ++it) {
            const long long T = cluster_id + (long long)it * num_clusters;
            if (T >= num_pairs) break;
            const long long p = (T * 2 + cta) * kTcPts + 8 * q + pl;
            const bool valid = p < n;
            float xin[D];
#pragma unroll
            for (int i = 0; i < D; ++i) xin[i] = valid ? fmaf(P.xa[i], __ldg(x + p * D + i), P.xb[i]) : 0.f;
            // whole image blocks are written: the tail rows (x = 0) stay finite
            float* sp = stash + slot_off(p, sub, tq);
            const bool do_stash = STASH && (p >> 3) < nblk8;
            // ---- input layer: z0 = W_0 x + b_0, dz0/dx_k = W_0[:,k], no second derivatives ----
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                const int f = 32 * c + fb;
                const float2 b2 = *reinterpret_cast<const float2*>(bias + f);
                float2 w2[D];
#pragma unroll
                for (int k = 0; k < D; ++k) w2[k] = *reinterpret_cast<const float2*>(W0t + k * 128 + f);
                float a[2][R];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float z0 = e ? b2.y : b2.x;
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        const float wk = e ? w2[k].y : w2[k].x;
                        z0 = fmaf(wk, xin[k], z0);
                        a[e][1 + k] = wk * P.xa[k];      // chain rule of x_eff = xa x + xb (xa = 1 outside mirrored value calls)
                    }
                    a[e][0] = z0;
                    a[e][1 + D] = 0.f;
                    tc_act_fwd<D, ACT>(a[e], act);
                }
                hand_over(c, a);
            }
            // ---- hidden layers 1 .. L-1 ----
            float o[R];
#pragma unroll
            for (int s = 0; s < R; ++s) o[s] = 0.f;
#pragma unroll 1
            for (int l = 1; l < L; ++l) {
                const uint32_t dcol = kColD0 + 128u * (uint32_t)(l & 1);
                if (l & 1) { mbar_wait(&d_full[1], dc1 & 1u); ++dc1; } else { mbar_wait(&d_full[0], dc0 & 1u); ++dc0; }
                tc_fence_after();
                const float* bl = bias + l * 128;
                const bool last = l == L - 1;
                float* spl = sp + (size_t)(l - 1) * stash_ls;
                uint32_t nLo[4], nHi[4];
                IO::ld(lane_lo + dcol + kCw * sub, nLo);
                IO::ld(lane_hi + dcol + kCw * sub, nHi);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int f = 32 * c + fb;
                    tmem_wait_ld();
                    uint32_t LoHi[4], HiHi[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                    if (c < 3) {   // the next chunk's accumulator columns travel while this chunk is computed
                        IO::ld(lane_lo + dcol + 32 * (c + 1) + kCw * sub, nLo);
                        IO::ld(lane_hi + dcol + 32 * (c + 1) + kCw * sub, nHi);
                    }
                    const float2 b2 = *reinterpret_cast<const float2*>(bl + f);
                    float a[2][R], sv8[8];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
#pragma unroll
                        for (int s = 0; s < R; ++s) a[e][s] = P.debias * __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                        a[e][0] += e ? b2.y : b2.x;
#pragma unroll
                        for (int s = 1; s < 4; ++s) sv8[4 * e + s] = s < R ? a[e][s] : 0.f;
                        sv8[4 * e] = tc_act_fwd<D, ACT>(a[e], act);
                    }
                    if (do_stash) stg_v8(spl + (size_t)c * 1024, sv8);
                    if (!last) {
                        hand_over(c, a);
                    } else {
                        const float2 wo = *reinterpret_cast<const float2*>(wout + f);
#pragma unroll
                        for (int s = 0; s < R; ++s) o[s] = fmaf(wo.y, a[1][s], fmaf(wo.x, a[0][s], o[s]));
                    }
                }
            }
            // output layer: reduce over the 4 lanes of a point, then over the warps of the quadrant
#pragma unroll
            for (int s = 0; s < R; ++s) {
                o[s] += __shfl_xor_sync(0xffffffffu, o[s], 1);
                o[s] += __shfl_xor_sync(0xffffffffu, o[s], 2);
            }
            if (tq == 0) {
#pragma unroll
                for (int s = 0; s < R; ++s) comb[(w * 8 + pl) * 4 + s] = o[s];
            }
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
            if (sub == 0 && tq == 0 && valid && out_u) {
#pragma unroll
                for (int k = 1; k < kWpq; ++k) {
                    const float* oc = comb + ((w + 4 * k) * 8 + pl) * 4;
#pragma unroll
                    for (int s = 0; s < R; ++s) o[s] += oc[s];
                }
                out_u[p] = o[0] + bout;
#pragma unroll
                for (int i = 0; i < D; ++i)
                    if (out_g) out_g[p * D + i] = o[1 + i];
                if (out_h) out_h[p] = o[1 + D];
            }
        }
    } else if (cta == 0 && lane == 0) {
        // ---- MMA issue thread (leader CTA) ----
        constexpr uint32_t idesc = instr_desc(2, 256, 128, 0, 0);
        const uint32_t wbase = smem_u32(wimg_s);
        uint32_t acnt = 0;
        for (int it = 0;; ++it) {
            const long long T = cluster_id + (long long)it * num_clusters;
            if (T >= num_pairs) break;
            for (int l = 1; l < L; ++l, ++acnt) {
                const uint32_t dcol = tb + kColD0 + 128u * (uint32_t)(l & 1);
                const uint32_t bhi0 = wbase + (uint32_t)(l - 1) * 2 * kImgFloats * 4, blo0 = bhi0 + kImgFloats * 4;
                const uint64_t dhi = smem_desc(bhi0, 128, 4096), dlo = smem_desc(blo0, 128, 4096);
                if (P.interleave >= 1) {
                    // hi*hi of a chunk right after its two correction products
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        mbar_wait(&a_ready[c], acnt & 1u);
                        tc_fence_after();
                        mma_chunk_3x<kBfFwd>(dcol, tb + kColAHi + 32 * c, tb + kColALo + 32 * c, dhi + (uint64_t)(64 * c), dlo + (uint64_t)(64 * c), idesc,
                                     c > 0 ? 1u : 0u);
                    }
                } else {
                    // the two small cross products of every chunk first, the 16 hi*hi products last: the accumulator is rounded
                    // toward zero after every MMA, so the correction MMAs land while it is ~2^-11 of its final size and only
                    // 16 roundings count
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        mbar_wait(&a_ready[c], acnt & 1u);
                        tc_fence_after();
                        mma_chunk_corr<kBfFwd>(dcol, tb + kColAHi + 32 * c, tb + kColALo + 32 * c, dhi + (uint64_t)(64 * c), dlo + (uint64_t)(64 * c), idesc,
                                       c > 0 ? 1u : 0u);
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) mma_chunk_hh(dcol, tb + kColAHi + 32 * c, dhi + (uint64_t)(64 * c), idesc, 1u);
                }
                mma_commit_pair(&d_full[l & 1], 3);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (w == kEpiWarps) tmem_dealloc<2>(tb, 512);
}

// ------------------------------------------------------------------------------------------
// Two tiles in flight per CTA pair (jet_fwd2_tc_kernel, rev2_role).
//
// With one tile per pair the tensor pipe idles whenever the epilogue warps work alone: the input layer, the last layer
// and, every layer, the first chunk plus the two hand-off latencies (epilogue -> MMA thread -> epilogue); measured on
// B200 the 144 MMAs of a tile are 52 % of its 17.7 k cycles.  Here the 16 epilogue warps alternate between two tiles X
// and Y at LAYER granularity -- L0(X) L0(Y) L1(X) L1(Y) ... -- so the MMAs of one tile run under the epilogue of the
// other.  What makes that fit the 512 TMEM columns:
//   * three accumulators that rotate: MMA step g (= one tile, one layer) writes buffer g % 3; the epilogue that reads
//     buffer (g - 3) % 3 is the one that feeds step g - 1, so it has finished when step g starts;
//   * ONE activation ring of two 32-column chunks (hi | lo, 64 TMEM columns each) shared by both tiles: only one tile
//     hands activations over at any time, the MMA thread consumes the chunks in the order they were produced, and a
//     chunk slot is reused as soon as the MMAs that read it have retired (a_free, a tcgen05.commit per chunk).
// The hi*hi MMAs of a chunk follow its two correction products (all three read the same chunk slot).
// ------------------------------------------------------------------------------------------
constexpr uint32_t k2ColD = 0, k2ColA = 384;      // accumulators 3 x 128 columns; ring slots 2 x (32 hi + 32 lo)

#ifdef GPE_TC_TRACE
// timing probe (debug builds only): clock64 stamps of the first chunks of cluster 0 -- [chunk][0..3] = MMA thread woke /
// MMA thread committed / epilogue warp 0 got the ring slot / epilogue warp 0 arrived
__device__ long long g_tc_trace[4096];
#define GPE_TRACE(cond, idx, slot) do { if ((cond) && (idx) < 1000) g_tc_trace[4 * (idx) + (slot)] = clock64(); } while (0)
#else
#define GPE_TRACE(cond, idx, slot) do { } while (0)
#endif

// MMA issue thread of the two-tile kernels: `steps` = (tile, layer) MMA steps of this pair in program order
template <bool BF>
__device__ __forceinline__ void mma2_issue(uint32_t tb, uint32_t wbase, int nmat, bool reverse, int iters, uint64_t* a_ready,
                                           uint64_t* a_free, uint64_t* d_full) {
    using tc::tc_fence_after; using tc::smem_desc; using tc::instr_desc; using tc::mma_commit_pair;
    constexpr uint32_t idesc = instr_desc(2, 256, 128, 0, 0);
    uint32_t m = 0, g = 0;
    for (int it = 0; it < iters; ++it) {
        for (int k = 0; k < nmat; ++k) {
            const int mat = reverse ? nmat - 1 - k : k;      // forward: W_1 .. W_{L-1}; reverse: W_{L-1}^T .. W_1^T
            const uint32_t bhi0 = wbase + (uint32_t)mat * 2 * kImgFloats * 4;
            const uint64_t dhi = smem_desc(bhi0, 128, 4096), dlo = smem_desc(bhi0 + kImgFloats * 4, 128, 4096);
#pragma unroll 1
            for (int t = 0; t < 2; ++t, ++g) {
                const uint32_t dcol = tb + k2ColD + 128u * (g % 3u);
#pragma unroll
                for (int c = 0; c < 4; ++c, ++m) {
                    const uint32_t slot = m & 1u;
                    mbar_wait(&a_ready[slot], (m >> 1) & 1u);
                    tc_fence_after();
                    GPE_TRACE(blockIdx.x == 0, m, 0);
                    const uint32_t ahi = tb + k2ColA + 64u * slot;
                    mma_chunk_3x<BF>(dcol, ahi, ahi + 32u, dhi + (uint64_t)(64 * c), dlo + (uint64_t)(64 * c), idesc, c > 0 ? 1u : 0u);
                    mma_commit_pair(&a_free[slot], 3);
                    GPE_TRACE(blockIdx.x == 0, m, 1);
                }
                mma_commit_pair(&d_full[g % 3u], 3);
            }
        }
    }
}

template <int D, bool STASH, int ACT = -1>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTcThreads, 1)
jet_fwd2_tc_kernel(TcParams P, const float* __restrict__ x, long long n, float* __restrict__ out_u, float* __restrict__ out_g,
                   float* __restrict__ out_h, float* __restrict__ stash, long long stash_ls, int nblk8, int num_pairs) {
    using tc::cluster_ctarank; using tc::cluster_sync; using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before;
    using tc::tc_fence_after; using tc::tmem_wait_ld; using tc::tmem_wait_st; using tc::mbar_arrive_cluster;
    using IO = FragIO<kNI>;
    constexpr int R = D + 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int L = P.n_hidden, nmat = L - 1;
    const int n_small = (D + L + 1) * 128 + 4;
    float* wimg_s = reinterpret_cast<float*>(smem);
    float* small = wimg_s + (size_t)nmat * 2 * kImgFloats;
    float* comb0 = small + n_small;
    uint64_t* a_ready = reinterpret_cast<uint64_t*>(comb0 + kEpiWarps * 32 * 4);
    uint64_t* a_free = a_ready + 2;
    uint64_t* d_full = a_free + 2;
    uint64_t* w_full = d_full + 3;
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(w_full + 1);
    float* comb1 = reinterpret_cast<float*>(smem + tc_smem_base_bytes(D, P.n_hidden));   // the reverse sweep's dW_0 area: free here

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31;
    const uint32_t cta = cluster_ctarank();
    const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
    // this pair's tiles: cluster_id, cluster_id + num_clusters, ...; two per iteration (the second may be past the end)
    const int my_tiles = num_pairs > cluster_id ? (num_pairs - cluster_id + num_clusters - 1) / num_clusters : 0;
    const int iters = (my_tiles + 1) / 2;

    if (tid == 0) {
        for (int c = 0; c < 2; ++c) {
            mbar_init(&a_ready[c], 2 * kEpiWarps);
            mbar_init(&a_free[c], 1);
        }
        for (int c = 0; c < 3; ++c) mbar_init(&d_full[c], 1);
        mbar_init(w_full, 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(w_full, (uint32_t)nmat * 2 * kImgFloats * 4);
        for (int l = 0; l < nmat; ++l)
            for (int hl = 0; hl < 2; ++hl)
                tma_bulk_g2s(wimg_s + ((size_t)l * 2 + hl) * kImgFloats, P.wimg + (((size_t)l * 2 + cta) * 2 + hl) * kImgFloats,
                             kImgFloats * 4, w_full);
        mbar_wait(w_full, 0);   // both halves must be resident before the leader's first MMA (cluster_sync below)
    }
    if (w == kEpiWarps) tmem_alloc<2>(tbase_s, 512);
    for (int i = tid; i < n_small; i += kTcThreads) small[i] = P.small[i];
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* W0t = small;
    const float* bias = small + D * 128;
    const float* wout = small + (D + L) * 128;
    const float bout = small[(D + L + 1) * 128];
    const int act = ACT >= 0 ? ACT : P.act;

    if (w < kEpiWarps) {
        const int q = w & 3, sub = w >> 2, pl = lane >> 2, tq = lane & 3;
        const uint32_t lane_lo = tb + ((uint32_t)(32 * q) << 16), lane_hi = lane_lo + (16u << 16);
        const int fb = kCw * sub + 2 * tq;                 // this thread's feature pair inside every 32-column chunk
        uint32_t nch = 0;                                  // chunks handed over so far (ring position)
        // split the four stream slots of features (f, f + 1), wait for the ring slot, hand the chunk to the MMA thread
        auto hand_over = [&](const float (&a)[2][R]) {
            AFrag fr;
            make_afrag<R, kBfFwd>(a, fr);
            const uint32_t slot = nch & 1u;
            if (nch >= 2) {                                // the MMAs that read this slot two chunks ago have retired
                mbar_wait(&a_free[slot], ((nch >> 1) - 1) & 1u);
                tc_fence_after();
            }
            GPE_TRACE(blockIdx.x == 0 && tid == 0, nch, 2);
            const uint32_t col = k2ColA + 64u * slot + (uint32_t)(kCw * sub);
            store_afrag<kBfFwd>(lane_lo, lane_hi, col, col + 32, fr);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&a_ready[slot], 0);
            GPE_TRACE(blockIdx.x == 0 && tid == 0, nch, 3);
            ++nch;
        };
        // Program order of the epilogue warps (the MMA steps keep their order: input layers of an iteration first):
        //   prologue   L0(X0) L0(Y0)
        //   iteration  L1(X) L1(Y) .. L_{L-2}(X) L_{L-2}(Y)   then   L0(X') L_last(X) L0(Y') L_last(Y)
        // i.e. the input layers of the NEXT pair of tiles are handed over between the last-layer epilogues of this pair, so
        // the tensor pipe has their MMAs to run while the epilogue produces outputs that feed no MMA.  Accumulator g mod 3
        // is free in that order: L0(X') starts step g' (buffer of step g'-3, read long ago); L_last(X) then frees the
        // buffer of step g'-2, which step g'+1 (fed by L0(Y')) takes; L_last(Y) reads the buffer of step g'-1.
        struct Tile {
            long long p;
            bool valid, st_ok;
            float* sp;
        };
        auto tile_state = [&](int it, int t) {
            const long long T = cluster_id + (long long)(2 * it + t) * num_clusters;   // >= num_pairs: a dummy tile
            Tile s;
            s.p = (T * 2 + cta) * kTcPts + 8 * q + pl;
            s.valid = T < num_pairs && s.p < n;
            s.st_ok = STASH && T < num_pairs && (s.p >> 3) < nblk8;   // whole image blocks are written: tail rows stay finite
            s.sp = stash + slot_off(s.st_ok ? s.p : 0, sub, tq);
            return s;
        };
        // input layer: z0 = W_0 x + b_0, dz0/dx_k = W_0[:,k], no second derivatives
        auto input_layer = [&](const Tile& ts) {
            float xin[D];
#pragma unroll
            for (int i = 0; i < D; ++i) xin[i] = ts.valid ? fmaf(P.xa[i], __ldg(x + ts.p * D + i), P.xb[i]) : 0.f;
#pragma unroll 1
            for (int c = 0; c < 4; ++c) {
                const int f = 32 * c + fb;
                const float2 b2 = *reinterpret_cast<const float2*>(bias + f);
                float2 w2[D];
#pragma unroll
                for (int k = 0; k < D; ++k) w2[k] = *reinterpret_cast<const float2*>(W0t + k * 128 + f);
                float a[2][R];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float z0 = e ? b2.y : b2.x;
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        const float wk = e ? w2[k].y : w2[k].x;
                        z0 = fmaf(wk, xin[k], z0);
                        a[e][1 + k] = wk * P.xa[k];      // chain rule of x_eff = xa x + xb (xa = 1 outside mirrored value calls)
                    }
                    a[e][0] = z0;
                    a[e][1 + D] = 0.f;
                    tc_act_fwd<D, ACT>(a[e], act);
                }
                hand_over(a);
            }
        };
        // hidden layer l >= 1 of tile t of iteration it: z_l arrives in the accumulator of MMA step g
        auto hidden_layer = [&](int it, int l, int t, const Tile& ts) {
            const float* bl = bias + l * 128;
            const bool last = l == L - 1;
            const uint32_t g = (uint32_t)it * 2u * (uint32_t)nmat + 2u * (uint32_t)(l - 1) + (uint32_t)t;
            const uint32_t dcol = k2ColD + 128u * (g % 3u);
            mbar_wait(&d_full[g % 3u], (g / 3u) & 1u);
            tc_fence_after();
            float* spl = ts.sp + (size_t)(l - 1) * stash_ls;
            float o[R];
#pragma unroll
            for (int s = 0; s < R; ++s) o[s] = 0.f;
            uint32_t nLo[4], nHi[4];
            IO::ld(lane_lo + dcol + kCw * sub, nLo);
            IO::ld(lane_hi + dcol + kCw * sub, nHi);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int f = 32 * c + fb;
                tmem_wait_ld();
                uint32_t LoHi[4], HiHi[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                if (c < 3) {   // the next chunk's accumulator columns travel while this chunk is computed
                    IO::ld(lane_lo + dcol + 32 * (c + 1) + kCw * sub, nLo);
                    IO::ld(lane_hi + dcol + 32 * (c + 1) + kCw * sub, nHi);
                }
                const float2 b2 = *reinterpret_cast<const float2*>(bl + f);
                float a[2][R], sv8[8];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
#pragma unroll
                    for (int s = 0; s < R; ++s) a[e][s] = P.debias * __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                    a[e][0] += e ? b2.y : b2.x;
#pragma unroll
                    for (int s = 1; s < 4; ++s) sv8[4 * e + s] = s < R ? a[e][s] : 0.f;
                    sv8[4 * e] = tc_act_fwd<D, ACT>(a[e], act);
                }
                if (ts.st_ok) stg_v8(spl + (size_t)c * 1024, sv8);
                if (!last) {
                    hand_over(a);
                } else {
                    const float2 wo = *reinterpret_cast<const float2*>(wout + f);
#pragma unroll
                    for (int s = 0; s < R; ++s) o[s] = fmaf(wo.y, a[1][s], fmaf(wo.x, a[0][s], o[s]));
                }
            }
            if (last) {
                // output layer: reduce over the 4 lanes of a point, then over the warps of the quadrant
                float* comb = t ? comb1 : comb0;
#pragma unroll
                for (int s = 0; s < R; ++s) {
                    o[s] += __shfl_xor_sync(0xffffffffu, o[s], 1);
                    o[s] += __shfl_xor_sync(0xffffffffu, o[s], 2);
                }
                if (tq == 0) {
#pragma unroll
                    for (int s = 0; s < R; ++s) comb[(w * 8 + pl) * 4 + s] = o[s];
                }
                asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
                if (sub == 0 && tq == 0 && ts.valid && out_u) {
#pragma unroll
                    for (int k = 1; k < kWpq; ++k) {
                        const float* oc = comb + ((w + 4 * k) * 8 + pl) * 4;
#pragma unroll
                        for (int s = 0; s < R; ++s) o[s] += oc[s];
                    }
                    out_u[ts.p] = o[0] + bout;
#pragma unroll
                    for (int i = 0; i < D; ++i)
                        if (out_g) out_g[ts.p * D + i] = o[1 + i];
                    if (out_h) out_h[ts.p] = o[1 + D];
                }
            }
        };
        Tile cur0{}, cur1{};
        if (iters > 0) {
            cur0 = tile_state(0, 0);
            cur1 = tile_state(0, 1);
            input_layer(cur0);
            input_layer(cur1);
        }
        for (int it = 0; it < iters; ++it) {
#pragma unroll 1
            for (int l = 1; l < L - 1; ++l) {
                hidden_layer(it, l, 0, cur0);
                hidden_layer(it, l, 1, cur1);
            }
            if (it + 1 < iters) {
                const Tile nx0 = tile_state(it + 1, 0), nx1 = tile_state(it + 1, 1);
                input_layer(nx0);
                hidden_layer(it, L - 1, 0, cur0);
                input_layer(nx1);
                hidden_layer(it, L - 1, 1, cur1);
                cur0 = nx0;
                cur1 = nx1;
            } else {
                hidden_layer(it, L - 1, 0, cur0);
                hidden_layer(it, L - 1, 1, cur1);
            }
        }
    } else if (cta == 0 && lane == 0) {
        mma2_issue<kBfFwd>(tb, smem_u32(wimg_s), nmat, false, iters, a_ready, a_free, d_full);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (w == kEpiWarps) tmem_dealloc<2>(tb, 512);
}

// ------------------------------------------------------------------------------------------
// Fused backward (bwd_fused_tc_kernel, below): reverse-sweep CTA pairs and dW CTAs run in ONE persistent launch and the
// zbar images travel from the former to the latter through a ring of tile slots that stays resident in L2 (every slot is
// overwritten before it can be evicted), so they never cost HBM traffic.  Flow control is a pair of monotonic counters in
// global memory:
//   prod[(l-1) * 2 nrc + 2 r + cta]   warp arrivals (kEpiWarps per finished half tile) of reverse CTA (r, cta) for layer l
//   cons[2 slot + cta]                consumer arrivals per half-tile slot: L-1 (one per layer class) per trip round the ring
// Every spin is bounded: a wait that outlasts kSpinLimit polls raises `abort`, which releases all other waits, so a
// protocol error ends the launch with a flag for the host instead of a hung GPU.
// ------------------------------------------------------------------------------------------
struct RingCtl {
    int ns;                  // tile slots (64 points each)
    int nrc;                 // reverse-sweep clusters
    unsigned int* prod;
    unsigned int* cons;
    unsigned int* abort;
};
constexpr unsigned int kSpinLimit = 1u << 23;   // polls (~ seconds); a whole 16M-point backward takes < 0.1 s
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned int* p, unsigned int v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// wait until *ctr >= target (or the launch is aborting)
__device__ __forceinline__ void ring_wait(const unsigned int* ctr, unsigned int target, unsigned int* abort) {
    unsigned int spins = 0;
    while (ld_acquire_gpu(ctr) < target) {
        if ((++spins & 63u) == 0u) {
            if (ld_acquire_gpu(abort) != 0u) return;
            if (spins > kSpinLimit) { atomicExch(abort, 1u); return; }
            __nanosleep(64);
        }
    }
}
// L2 eviction-priority policies: the stash is read once (evict first), the ring should stay (evict last)
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void ldg_v8_hint(const float* g, float (&v)[8], uint64_t pol) {
    asm volatile("ld.global.L2::cache_hint.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8], %9;"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
                 : "l"(g), "l"(pol));
}
__device__ __forceinline__ void stg_v8_hint(float* g, const float (&v)[8], uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8}, %9;" ::"l"(g), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]),
                 "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]), "l"(pol)
                 : "memory");
}

// ------------------------------------------------------------------------------------------
// reverse sweep: given the adjoints (u_bar, g_bar, h_bar) of the per-point outputs and the stashed forward
// jets, propagate  abar_{l-1} = zbar_l . W_l  down the layers on the tensor cores (B = images of W_l^T) and
// write zbar_l (adjoint of the pre-activation jets, all streams) of every hidden matrix layer l >= 1 to the image blocks
// zb[l-1][point / 8].  The weight gradients dW_l = zbar_l^T a_{l-1} are a contraction over points: dw_tc_kernel.
// Also accumulates the output-layer gradient (dwout, dbout) and the input-layer gradient (dW_0, db_0: 3 x 128 numbers,
// summed in registers over the tiles of the CTA) into small_acc (layout of `small`).
// ------------------------------------------------------------------------------------------
// RING = false: one chunk of points per launch, zbar_l of the chunk written to zb[l-1][chunk block] (jet_rev_tc_kernel).
// RING = true : all points in one launch; tile T writes to ring slot T % ns and signals the dW CTAs (bwd_fused_tc_kernel).
// cluster_id / num_clusters: this pair's index among the reverse-sweep pairs of the launch.
template <int D, int ACT, bool RING>
__device__ __forceinline__ void rev_role(const TcParams& P, const float* __restrict__ x, long long n, const float* __restrict__ u_bar,
                                         const float* __restrict__ g_bar, const float* __restrict__ h_bar,
                                         const float* __restrict__ stash, float* __restrict__ zb, float* __restrict__ small_acc,
                                         long long stash_ls, int nblk8, int num_pairs, int pf_dist, int cluster_id, int num_clusters,
                                         const RingCtl& ring) {
    using tc::cluster_ctarank; using tc::cluster_sync; using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before;
    using tc::tc_fence_after; using tc::tmem_wait_ld; using tc::tmem_wait_st; using tc::mbar_arrive_cluster;
    using tc::mbar_wait_cluster; using tc::smem_desc; using tc::instr_desc; using tc::mma_ts; using tc::mma_commit_pair;
    using IO = FragIO<kNI>;
    constexpr int R = D + 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int L = P.n_hidden, nmat = L - 1;
    const int n_small = (D + L + 1) * 128 + 4;
    float* wimg_s = reinterpret_cast<float*>(smem);
    float* small = wimg_s + (size_t)nmat * 2 * kImgFloats;
    float* comb = small + n_small;
    uint64_t* a_ready = reinterpret_cast<uint64_t*>(comb + kEpiWarps * 32 * 4);
    uint64_t* d_full = a_ready + 4;
    uint64_t* w_full = d_full + 2;
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(w_full + 1);
    float* dw0s = reinterpret_cast<float*>(smem + tc_smem_base_bytes(D, P.n_hidden));

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, nthr = blockDim.x;
    const uint32_t cta = cluster_ctarank();

    if (tid == 0) {
        for (int c = 0; c < 4; ++c) mbar_init(&a_ready[c], 2 * kEpiWarps);
        mbar_init(&d_full[0], 1);
        mbar_init(&d_full[1], 1);
        mbar_init(w_full, 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(w_full, (uint32_t)nmat * 2 * kImgFloats * 4);
        for (int l = 0; l < nmat; ++l)
            for (int hl = 0; hl < 2; ++hl)
                tma_bulk_g2s(wimg_s + ((size_t)l * 2 + hl) * kImgFloats, P.wimg + (((size_t)l * 2 + cta) * 2 + hl) * kImgFloats,
                             kImgFloats * 4, w_full);
        mbar_wait(w_full, 0);
    }
    if (w == kEpiWarps) tmem_alloc<2>(tbase_s, 512);
    for (int i = tid; i < (D + 1) * 128 * kDw0Stride; i += nthr) dw0s[i] = 0.f;
    for (int i = tid; i < n_small; i += nthr) small[i] = P.small[i];
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* W0t = small;
    const float* bias = small + D * 128;
    const float* wout = small + (D + L) * 128;
    const int act = ACT >= 0 ? ACT : P.act;

    if (w < kEpiWarps) {
        // Three code paths per tile: last hidden layer (adjoint from the output layer, no accumulator), middle layers,
        // input layer (forward jets recomputed from x, gradient summed in shared memory, nothing handed on).
        const int q = w & 3, sub = w >> 2, pl = lane >> 2, tq = lane & 3;
        const uint32_t lane_lo = tb + ((uint32_t)(32 * q) << 16), lane_hi = lane_lo + (16u << 16);
        const int fb = kCw * sub + 2 * tq;                 // this thread's feature pair inside every 32-column chunk
        uint32_t dc0 = 0u, dc1 = 0u;
        float dwout_acc[4][2];
#pragma unroll
        for (int a = 0; a < 4; ++a) dwout_acc[a][0] = dwout_acc[a][1] = 0.f;
        float dbout_acc = 0.f;
        float dw0_acc[4][2][D + 1];   // input-layer gradient of this thread's 8 features: dW_0[f][k], db_0[f]
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int k = 0; k <= D; ++k) dw0_acc[a][e][k] = 0.f;
        // zbar destination: the chunk buffer (floats per layer = blocks of the chunk) or the ring (ns tile slots of 8 blocks)
        const size_t lstride = RING ? (size_t)ring.ns * 8 * 4096 : (size_t)nblk8 * 4096;
        const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
        auto hand_over = [&](int c, const float (&a)[2][R]) {
            AFrag fr;
            make_afrag<R, kBfRev>(a, fr);
            const uint32_t col = (uint32_t)(32 * c + kCw * sub);
            store_afrag<kBfRev>(lane_lo, lane_hi, kColAHi + col, kColALo + col, fr);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&a_ready[c], 0);
        };
        auto store_zbar = [&](float* dst, const float (&ab)[2][R]) {
            float zv8[8];
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int s = 0; s < 4; ++s) zv8[4 * e + s] = s < R ? ab[e][s] : 0.f;
            if (RING) stg_v8_hint(dst, zv8, pol_last); else stg_v8(dst, zv8);
        };
        // all zbar_l stores of this warp for the current tile are out: one release-arrive per warp (kEpiWarps per half tile)
        auto signal_layer = [&](int l) {
            if (RING) {
                __syncwarp();
                if (lane == 0) {
                    __threadfence();
                    red_release_gpu_add(ring.prod + (size_t)(l - 1) * 2 * ring.nrc + 2 * cluster_id + cta, 1u);
                }
            }
        };
        for (int it = 0;; ++it) {
            const long long T = cluster_id + (long long)it * num_clusters;
            if (T >= num_pairs) break;
            const long long p = (T * 2 + cta) * kTcPts + 8 * q + pl;
            const bool valid = p < n;
            float xin[D], adj[R];
#pragma unroll
            for (int i = 0; i < D; ++i) xin[i] = valid ? __ldg(x + p * D + i) : 0.f;
            adj[0] = (valid && u_bar) ? __ldg(u_bar + p) : 0.f;
#pragma unroll
            for (int i = 0; i < D; ++i) adj[1 + i] = (valid && g_bar) ? __ldg(g_bar + p * D + i) : 0.f;
            adj[1 + D] = (valid && h_bar) ? __ldg(h_bar + p) : 0.f;
            if (sub == 0 && tq == 0) dbout_acc += adj[0];
            // L2 bulk prefetch of the stash images one layer ahead of their use (4 image blocks = 64 KB contiguous per
            // tile and layer).  Not further ahead: at ~4.5 TB/s the whole L2 turns over in ~25 us.
            auto prefetch_stash = [&](int l, long long Tt) {
                const long long bn = ((Tt * 2 + cta) * kTcPts) >> 3;
                if (Tt < num_pairs && bn < nblk8)
                    tc::prefetch_l2_bulk(stash + (size_t)(l - 1) * stash_ls + (size_t)bn * 4096, (uint32_t)(nblk8 - bn < 4 ? nblk8 - bn : 4) * 16384u);
            };
            auto rolling_prefetch = [&](int l) {   // the stash image consumed pf_dist (tile, layer) steps from now
                if (tid == 0) {
                    const int tgt = (L - 1 - l) + pf_dist;
                    prefetch_stash(L - 1 - tgt % (L - 1), T + (long long)(tgt / (L - 1)) * num_clusters);
                }
            };
            const bool in_blk = (p >> 3) < nblk8;   // whole blocks are read and written; rows of points >= n carry zero adjoints
            // stashed forward jets of (layer l, chunk c) for this thread: one 32-byte slot per 8-column group.  Loaded one
            // chunk ahead (and across the layer boundary) so the load latency hides behind the previous chunk.
            const float* sbase = stash + slot_off(in_blk ? p : 0, sub, tq);
            // RING: tile T lives in slot T % ns; wait until the L-1 layer classes have drained what the previous trip
            // round the ring left there (each adds 1 per half tile)
            const long long pz = RING ? (long long)(T % ring.ns) * (2 * kTcPts) + cta * kTcPts + 8 * q + pl : (in_blk ? p : 0);
            float* zbase = zb + slot_off(pz, sub, tq);
            if (RING) {
                if (lane == 0) ring_wait(ring.cons + 2 * (T % ring.ns) + cta, (unsigned int)((L - 1) * (T / ring.ns)), ring.abort);
                __syncwarp();
            }
            auto load_stash = [&](int l, int c, float (&dst)[8]) {
                const float* src = sbase + (size_t)(l - 1) * stash_ls + (size_t)c * 1024;
                if (RING) ldg_v8_hint(src, dst, pol_first); else ldg_v8(src, dst);
            };
            if (tid == 0 && it == 0)   // the first tile's images that the rolling prefetch below does not cover
                for (int j = 0; j < pf_dist && j < L - 1; ++j) prefetch_stash(L - 1 - j, T);
            float cur[8];
            load_stash(L - 1, 0, cur);
            // ---- last hidden layer l = L-1: abar = adj * wout; output-layer gradient ----
            rolling_prefetch(L - 1);
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                const int f = 32 * c + fb;
                float nxt[8];
                if (c < 3) load_stash(L - 1, c + 1, nxt);
                else if (L > 2) load_stash(L - 2, 0, nxt);
                const float2 wo = *reinterpret_cast<const float2*>(wout + f);
                float ab[2][R];
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    float st[R], a[R];
#pragma unroll
                    for (int s = 0; s < R; ++s) st[s] = cur[4 * e + s];
                    tc_act_from_stash<D, ACT>(st, a, act);
                    float dw = dwout_acc[c][e];
#pragma unroll
                    for (int s = 0; s < R; ++s) {
                        dw = fmaf(adj[s], a[s], dw);
                        ab[e][s] = adj[s] * (e ? wo.y : wo.x);
                    }
                    dwout_acc[c][e] = dw;
                    tc_act_bwd<D, ACT>(st, ab[e], act);
                }
                if (in_blk) store_zbar(zbase + (size_t)(L - 2) * lstride + (size_t)c * 1024, ab);
                hand_over(c, ab);
#pragma unroll
                for (int s = 0; s < 8; ++s) cur[s] = nxt[s];
            }
            signal_layer(L - 1);
            // ---- middle layers l = L-2 .. 1 ----
#pragma unroll 1
            for (int l = L - 2; l >= 1; --l) {
                rolling_prefetch(l);
                const uint32_t dcol = kColD0 + 128u * (uint32_t)(l & 1);
                if (l & 1) { mbar_wait(&d_full[1], dc1 & 1u); ++dc1; } else { mbar_wait(&d_full[0], dc0 & 1u); ++dc0; }
                tc_fence_after();
                uint32_t nLo[4], nHi[4];
                IO::ld(lane_lo + dcol + kCw * sub, nLo);
                IO::ld(lane_hi + dcol + kCw * sub, nHi);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float nxt[8];
                    if (c < 3) load_stash(l, c + 1, nxt);
                    else if (l > 1) load_stash(l - 1, 0, nxt);
                    tmem_wait_ld();
                    uint32_t LoHi[4], HiHi[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                    if (c < 3) {
                        IO::ld(lane_lo + dcol + 32 * (c + 1) + kCw * sub, nLo);
                        IO::ld(lane_hi + dcol + 32 * (c + 1) + kCw * sub, nHi);
                    }
                    float ab[2][R];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float st[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) {
                            st[s] = cur[4 * e + s];
                            ab[e][s] = __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                        }
                        tc_act_bwd<D, ACT>(st, ab[e], act);
                    }
                    if (in_blk) store_zbar(zbase + (size_t)(l - 1) * lstride + (size_t)c * 1024, ab);
                    hand_over(c, ab);
#pragma unroll
                    for (int s = 0; s < 8; ++s) cur[s] = nxt[s];
                }
                signal_layer(l);
            }
            // ---- input layer l = 0: forward jets from x; dW_0[f][k] = sum_p zbar_val x_k + zbar_dk, db_0[f] = sum_p zbar_val ----
            {
                mbar_wait(&d_full[0], dc0 & 1u);
                ++dc0;
                tc_fence_after();
                uint32_t nLo[4], nHi[4];
                IO::ld(lane_lo + kColD0 + kCw * sub, nLo);
                IO::ld(lane_hi + kColD0 + kCw * sub, nHi);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int f = 32 * c + fb;
                    tmem_wait_ld();
                    uint32_t LoHi[4], HiHi[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                    if (c < 3) {
                        IO::ld(lane_lo + kColD0 + 32 * (c + 1) + kCw * sub, nLo);
                        IO::ld(lane_hi + kColD0 + 32 * (c + 1) + kCw * sub, nHi);
                    }
                    const float2 b2 = *reinterpret_cast<const float2*>(bias + f);
                    float2 w2[D];
#pragma unroll
                    for (int k = 0; k < D; ++k) w2[k] = *reinterpret_cast<const float2*>(W0t + k * 128 + f);
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float st[R], ab[R];
                        float z0 = e ? b2.y : b2.x;
#pragma unroll
                        for (int k = 0; k < D; ++k) {
                            const float wk = e ? w2[k].y : w2[k].x;
                            z0 = fmaf(wk, xin[k], z0);
                            st[1 + k] = wk;
                        }
                        st[1 + D] = 0.f;
                        if (ACT == ACT_TANH || ACT == ACT_SHIFTED_TANH) {
                            st[0] = tanh_tc(z0);
                        } else {
                            float zz[R];
#pragma unroll
                            for (int s = 1; s < R; ++s) zz[s] = st[s];
                            zz[0] = z0;
                            act_jet_fwd<D, JET_LAPLACIAN>(zz, act, st[0]);
                        }
#pragma unroll
                        for (int s = 0; s < R; ++s) ab[s] = __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                        tc_act_bwd<D, ACT>(st, ab, act);
                        // register accumulators over the tiles of this thread (shared-memory float atomics are CAS loops)
#pragma unroll
                        for (int k = 0; k < D; ++k) dw0_acc[c][e][k] += fmaf(ab[0], xin[k], ab[1 + k]);
                        dw0_acc[c][e][D] += ab[0];
                    }
                }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");   // all D reads of this tile done
        }
        // flush the output-layer gradient: sum over the 8 points of the warp, one atomic per feature
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float v = dwout_acc[c][j];
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                if (pl == 0) atomicAdd(&small_acc[(D + L) * 128 + 32 * c + fb + j], v);
            }
        // input-layer gradient: sum over the 8 points of the warp, then over the 4 quadrant warps of a feature group through
        // shared memory (slot q), one global atomic per element and CTA (rows 0..D-1: dW_0^T, row D: b_0)
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int k = 0; k <= D; ++k) {
                    float v = dw0_acc[c][e][k];
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (pl == 0) dw0s[(size_t)(k * 128 + 32 * c + fb + e) * kDw0Stride + q] = v;
                }
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
        for (int i = tid; i < (D + 1) * 128; i += kEpiWarps * 32) {
            const float* dq = dw0s + (size_t)i * kDw0Stride;
            atomicAdd(&small_acc[i], (dq[0] + dq[1]) + (dq[2] + dq[3]));
        }
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 4);
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 8);
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 16);
        if (lane == 0 && sub == 0) atomicAdd(&small_acc[(D + L + 1) * 128], dbout_acc);
    } else if (w == kEpiWarps && cta == 0 && lane == 0) {
        constexpr uint32_t idesc = instr_desc(2, 256, 128, 0, 0);
        const uint32_t wbase = smem_u32(wimg_s);
        uint32_t acnt = 0;
        for (int it = 0;; ++it) {
            const long long T = cluster_id + (long long)it * num_clusters;
            if (T >= num_pairs) break;
            for (int l = L - 1; l >= 1; --l, ++acnt) {
                const uint32_t dcol = tb + kColD0 + 128u * (uint32_t)((l - 1) & 1);
                const uint32_t bhi0 = wbase + (uint32_t)(l - 1) * 2 * kImgFloats * 4, blo0 = bhi0 + kImgFloats * 4;
                const uint64_t dhi = smem_desc(bhi0, 128, 4096), dlo = smem_desc(blo0, 128, 4096);
                if (P.interleave >= 1) {
                    // hi*hi of a chunk right after its two correction products
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        mbar_wait(&a_ready[c], acnt & 1u);
                        tc_fence_after();
                        mma_chunk_3x<kBfRev>(dcol, tb + kColAHi + 32 * c, tb + kColALo + 32 * c, dhi + (uint64_t)(64 * c), dlo + (uint64_t)(64 * c), idesc,
                                     c > 0 ? 1u : 0u);
                    }
                } else {
                    // the two small cross products of every chunk first, the 16 hi*hi products last: the accumulator is rounded
                    // toward zero after every MMA, so the correction MMAs land while it is ~2^-11 of its final size and only
                    // 16 roundings count
#pragma unroll
                    for (int c = 0; c < 4; ++c) {
                        mbar_wait(&a_ready[c], acnt & 1u);
                        tc_fence_after();
                        mma_chunk_corr<kBfRev>(dcol, tb + kColAHi + 32 * c, tb + kColALo + 32 * c, dhi + (uint64_t)(64 * c), dlo + (uint64_t)(64 * c), idesc,
                                       c > 0 ? 1u : 0u);
                    }
#pragma unroll
                    for (int c = 0; c < 4; ++c) mma_chunk_hh(dcol, tb + kColAHi + 32 * c, dhi + (uint64_t)(64 * c), idesc, 1u);
                }
                mma_commit_pair(&d_full[(l - 1) & 1], 3);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (w == kEpiWarps) tmem_dealloc<2>(tb, 512);
}

// ------------------------------------------------------------------------------------------
// reverse sweep with two tiles in flight (see the two-tile notes above jet_fwd2_tc_kernel): same mathematics, stash and zbar
// layout as rev_role; the epilogue warps alternate between tiles X and Y layer by layer (top layer of X, top layer of Y,
// next layer of X, ...), three accumulators rotate and one two-slot activation ring is shared.  Nothing per tile is kept in
// registers across layers -- adjoints are read at the top layer, x at the input layer -- so the register budget is rev_role's.
// ------------------------------------------------------------------------------------------
template <int D, int ACT, bool RING>
__device__ __forceinline__ void rev2_role(const TcParams& P, const float* __restrict__ x, long long n, const float* __restrict__ u_bar,
                                          const float* __restrict__ g_bar, const float* __restrict__ h_bar,
                                          const float* __restrict__ stash, float* __restrict__ zb, float* __restrict__ small_acc,
                                          long long stash_ls, int nblk8, int num_pairs, int cluster_id, int num_clusters, const RingCtl& ring) {
    using tc::cluster_ctarank; using tc::cluster_sync; using tc::tmem_alloc; using tc::tmem_dealloc; using tc::tc_fence_before;
    using tc::tc_fence_after; using tc::tmem_wait_ld; using tc::tmem_wait_st; using tc::mbar_arrive_cluster;
    using IO = FragIO<kNI>;
    constexpr int R = D + 2;
    extern __shared__ __align__(1024) unsigned char smem[];
    const int L = P.n_hidden, nmat = L - 1;
    const int n_small = (D + L + 1) * 128 + 4;
    float* wimg_s = reinterpret_cast<float*>(smem);
    float* small = wimg_s + (size_t)nmat * 2 * kImgFloats;
    float* comb = small + n_small;
    uint64_t* a_ready = reinterpret_cast<uint64_t*>(comb + kEpiWarps * 32 * 4);
    uint64_t* a_free = a_ready + 2;
    uint64_t* d_full = a_free + 2;
    uint64_t* w_full = d_full + 3;
    uint32_t* tbase_s = reinterpret_cast<uint32_t*>(w_full + 1);
    float* dw0s = reinterpret_cast<float*>(smem + tc_smem_base_bytes(D, P.n_hidden));

    const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, nthr = blockDim.x;
    const uint32_t cta = cluster_ctarank();
    const int my_tiles = num_pairs > cluster_id ? (num_pairs - cluster_id + num_clusters - 1) / num_clusters : 0;
    const int iters = (my_tiles + 1) / 2;

    if (tid == 0) {
        for (int c = 0; c < 2; ++c) {
            mbar_init(&a_ready[c], 2 * kEpiWarps);
            mbar_init(&a_free[c], 1);
        }
        for (int c = 0; c < 3; ++c) mbar_init(&d_full[c], 1);
        mbar_init(w_full, 1);
        mbar_fence_init();
        mbar_arrive_expect_tx(w_full, (uint32_t)nmat * 2 * kImgFloats * 4);
        for (int l = 0; l < nmat; ++l)
            for (int hl = 0; hl < 2; ++hl)
                tma_bulk_g2s(wimg_s + ((size_t)l * 2 + hl) * kImgFloats, P.wimg + (((size_t)l * 2 + cta) * 2 + hl) * kImgFloats,
                             kImgFloats * 4, w_full);
        mbar_wait(w_full, 0);
    }
    if (w == kEpiWarps) tmem_alloc<2>(tbase_s, 512);
    for (int i = tid; i < n_small; i += nthr) small[i] = P.small[i];
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* W0t = small;
    const float* bias = small + D * 128;
    const float* wout = small + (D + L) * 128;
    const int act = ACT >= 0 ? ACT : P.act;

    if (w < kEpiWarps) {
        const int q = w & 3, sub = w >> 2, pl = lane >> 2, tq = lane & 3;
        const uint32_t lane_lo = tb + ((uint32_t)(32 * q) << 16), lane_hi = lane_lo + (16u << 16);
        const int fb = kCw * sub + 2 * tq;                 // this thread's feature pair inside every 32-column chunk
        uint32_t nch = 0;                                  // chunks handed over so far (ring position)
        float dwout_acc[4][2];
#pragma unroll
        for (int a = 0; a < 4; ++a) dwout_acc[a][0] = dwout_acc[a][1] = 0.f;
        float dbout_acc = 0.f;
        float dw0_acc[4][2][D + 1];   // input-layer gradient of this thread's 8 features: dW_0[f][k], db_0[f]
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int k = 0; k <= D; ++k) dw0_acc[a][e][k] = 0.f;
        const size_t lstride = RING ? (size_t)ring.ns * 8 * 4096 : (size_t)nblk8 * 4096;
        const uint64_t pol_first = l2_policy_evict_first(), pol_last = l2_policy_evict_last();
        auto hand_over = [&](const float (&a)[2][R]) {
            AFrag fr;
            make_afrag<R, kBfRev>(a, fr);
            const uint32_t slot = nch & 1u;
            if (nch >= 2) {                                // the MMAs that read this slot two chunks ago have retired
                mbar_wait(&a_free[slot], ((nch >> 1) - 1) & 1u);
                tc_fence_after();
            }
            GPE_TRACE(blockIdx.x == 0 && tid == 0, nch, 2);
            const uint32_t col = k2ColA + 64u * slot + (uint32_t)(kCw * sub);
            store_afrag<kBfRev>(lane_lo, lane_hi, col, col + 32, fr);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(&a_ready[slot], 0);
            GPE_TRACE(blockIdx.x == 0 && tid == 0, nch, 3);
            ++nch;
        };
        auto store_zbar = [&](float* dst, const float (&ab)[2][R]) {
            float zv8[8];
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int s = 0; s < 4; ++s) zv8[4 * e + s] = s < R ? ab[e][s] : 0.f;
            if (RING) stg_v8_hint(dst, zv8, pol_last); else stg_v8(dst, zv8);
        };
        // per-tile quantities, recomputed where needed (tile t of iteration it)
        auto tile_of = [&](int it, int t) -> long long { return cluster_id + (long long)(2 * it + t) * num_clusters; };
        auto point_of = [&](long long T) -> long long { return (T * 2 + cta) * kTcPts + 8 * q + pl; };
        auto stash_ptr = [&](long long T) -> const float* {
            const long long p = point_of(T);
            return stash + slot_off((T < num_pairs && (p >> 3) < nblk8) ? p : 0, sub, tq);
        };
        auto load_stash = [&](const float* sb, int l, int c, float (&dst)[8]) {
            const float* src = sb + (size_t)(l - 1) * stash_ls + (size_t)c * 1024;
            if (RING) ldg_v8_hint(src, dst, pol_first); else ldg_v8(src, dst);
        };
        // L2 bulk prefetch of the stash image of (tile T, layer l): 4 image blocks = 64 KB contiguous per CTA
        auto prefetch_stash = [&](int l, long long T) {
            const long long bn = ((T * 2 + cta) * kTcPts) >> 3;
            if (l >= 1 && T < num_pairs && bn < nblk8)
                tc::prefetch_l2_bulk(stash + (size_t)(l - 1) * stash_ls + (size_t)bn * 4096, (uint32_t)(nblk8 - bn < 4 ? nblk8 - bn : 4) * 16384u);
        };
        float cur[8];
        if (iters > 0) {
            if (tid == 0) { prefetch_stash(L - 1, tile_of(0, 0)); prefetch_stash(L - 1, tile_of(0, 1)); }
            load_stash(stash_ptr(tile_of(0, 0)), L - 1, 0, cur);
        }
        // Program order of the epilogue warps (the MMA steps and the stash reads keep their order; the input layer hands
        // nothing over and reads no stash):
        //   prologue   top(X0) top(Y0)
        //   iteration  l = L-2 .. 1: mid(l, X) mid(l, Y)   then   top(X') in(X) top(Y') in(Y)
        // so the first MMA steps of the NEXT pair run while the input-layer epilogues (which feed no MMA) are computed.
        // Accumulator g mod 3 is free in that order exactly as in jet_fwd2_tc_kernel.
        struct Tile {
            long long T, p;
            bool live, valid, in_blk;
            const float* sb;
            float* zbase;
        };
        auto tile_state = [&](int it, int t) {
            Tile s;
            s.T = tile_of(it, t);
            s.p = point_of(s.T);
            s.live = s.T < num_pairs;                      // a dummy tile past the end: computed, never stored
            s.valid = s.live && s.p < n;
            s.in_blk = s.live && (s.p >> 3) < nblk8;       // whole blocks are read and written
            s.sb = stash_ptr(s.T);
            const long long pz = RING ? (long long)(s.T % ring.ns) * (2 * kTcPts) + cta * kTcPts + 8 * q + pl : (s.in_blk ? s.p : 0);
            s.zbase = zb + slot_off(pz, sub, tq);
            return s;
        };
        // a layer with stash (l >= 1) of tile t of iteration it
        auto stash_layer = [&](int it, int l, int t) {
            const Tile ts = tile_state(it, t);
            const long long T = ts.T, p = ts.p;
            const bool live = ts.live, valid = ts.valid, in_blk = ts.in_blk;
            const float* sb = ts.sb;
            float* zbase = ts.zbase;
            // the stash this thread needs next, in program order: (same tile, next chunk) .. (other tile, same layer) ..
            // (first tile, next layer) .. (first tile of the next iteration, top layer)
            const long long Tn = t == 0 ? tile_of(it, 1) : (l > 1 ? tile_of(it, 0) : tile_of(it + 1, 0));
            const int ln = t == 0 ? l : (l > 1 ? l - 1 : L - 1);
            const bool has_next = t == 0 || l > 1 || it + 1 < iters;
            const float* sbn = stash_ptr(Tn);
            if (tid == 0) prefetch_stash(l > 1 ? l - 1 : L - 1, l > 1 ? T : tile_of(it + 1, t));   // this tile's next image
            if (l == L - 1) {
                // ---- top hidden layer: abar = adj * wout; output-layer gradient ----
                float adj[R];
                adj[0] = (valid && u_bar) ? __ldg(u_bar + p) : 0.f;
#pragma unroll
                for (int i = 0; i < D; ++i) adj[1 + i] = (valid && g_bar) ? __ldg(g_bar + p * D + i) : 0.f;
                adj[1 + D] = (valid && h_bar) ? __ldg(h_bar + p) : 0.f;
                if (sub == 0 && tq == 0) dbout_acc += adj[0];
                if (RING && live) {   // tile T lives in ring slot T % ns: wait until the dW CTAs have drained its previous tenant
                    if (lane == 0) ring_wait(ring.cons + 2 * (T % ring.ns) + cta, (unsigned int)((L - 1) * (T / ring.ns)), ring.abort);
                    __syncwarp();
                }
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int f = 32 * c + fb;
                    float nxt[8];
                    if (c < 3) load_stash(sb, l, c + 1, nxt);
                    else if (has_next) load_stash(sbn, ln, 0, nxt);
                    const float2 wo = *reinterpret_cast<const float2*>(wout + f);
                    float ab[2][R];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float st[R], a[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) st[s] = cur[4 * e + s];
                        tc_act_from_stash<D, ACT>(st, a, act);
                        float dw = dwout_acc[c][e];
#pragma unroll
                        for (int s = 0; s < R; ++s) {
                            dw = fmaf(adj[s], a[s], dw);
                            ab[e][s] = adj[s] * (e ? wo.y : wo.x);
                        }
                        dwout_acc[c][e] = dw;
                        tc_act_bwd<D, ACT>(st, ab[e], act);
                    }
                    if (in_blk) store_zbar(zbase + (size_t)(l - 1) * lstride + (size_t)c * 1024, ab);
                    hand_over(ab);
#pragma unroll
                    for (int s = 0; s < 8; ++s) cur[s] = nxt[s];
                }
            } else {
                const uint32_t g = (uint32_t)it * 2u * (uint32_t)nmat + 2u * (uint32_t)(L - 2 - l) + (uint32_t)t;
                const uint32_t dcol = k2ColD + 128u * (g % 3u);
                mbar_wait(&d_full[g % 3u], (g / 3u) & 1u);
                tc_fence_after();
                uint32_t nLo[4], nHi[4];
                IO::ld(lane_lo + dcol + kCw * sub, nLo);
                IO::ld(lane_hi + dcol + kCw * sub, nHi);
                // ---- middle layers ----
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    float nxt[8];
                    if (c < 3) load_stash(sb, l, c + 1, nxt);
                    else if (has_next) load_stash(sbn, ln, 0, nxt);
                    tmem_wait_ld();
                    uint32_t LoHi[4], HiHi[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                    if (c < 3) {
                        IO::ld(lane_lo + dcol + 32 * (c + 1) + kCw * sub, nLo);
                        IO::ld(lane_hi + dcol + 32 * (c + 1) + kCw * sub, nHi);
                    }
                    float ab[2][R];
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float st[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) {
                            st[s] = cur[4 * e + s];
                            ab[e][s] = __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                        }
                        tc_act_bwd<D, ACT>(st, ab[e], act);
                    }
                    if (in_blk) store_zbar(zbase + (size_t)(l - 1) * lstride + (size_t)c * 1024, ab);
                    hand_over(ab);
#pragma unroll
                    for (int s = 0; s < 8; ++s) cur[s] = nxt[s];
                }
            }
            if (RING && live) {   // zbar_l of this half tile is out: one release-arrive per warp
                __syncwarp();
                if (lane == 0) {
                    __threadfence();
                    red_release_gpu_add(ring.prod + (size_t)(l - 1) * 2 * ring.nrc + 2 * cluster_id + cta, 1u);
                }
            }
        };
        // the input layer of tile t of iteration it: abar_0 arrives in the accumulator of the tile's last MMA step
        auto input_layer = [&](int it, int t) {
            const Tile ts = tile_state(it, t);
            const long long p = ts.p;
            const bool valid = ts.valid;
            const uint32_t g = (uint32_t)it * 2u * (uint32_t)nmat + 2u * (uint32_t)(L - 2) + (uint32_t)t;
            const uint32_t dcol = k2ColD + 128u * (g % 3u);
            mbar_wait(&d_full[g % 3u], (g / 3u) & 1u);
            tc_fence_after();
            uint32_t nLo[4], nHi[4];
            IO::ld(lane_lo + dcol + kCw * sub, nLo);
            IO::ld(lane_hi + dcol + kCw * sub, nHi);
                // ---- input layer: forward jets from x; dW_0[f][k] = sum_p zbar_val x_k + zbar_dk, db_0[f] = sum_p zbar_val ----
                float xin[D];
#pragma unroll
                for (int i = 0; i < D; ++i) xin[i] = valid ? __ldg(x + p * D + i) : 0.f;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int f = 32 * c + fb;
                    tmem_wait_ld();
                    uint32_t LoHi[4], HiHi[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) LoHi[j] = nLo[j], HiHi[j] = nHi[j];
                    if (c < 3) {
                        IO::ld(lane_lo + dcol + 32 * (c + 1) + kCw * sub, nLo);
                        IO::ld(lane_hi + dcol + 32 * (c + 1) + kCw * sub, nHi);
                    }
                    const float2 b2 = *reinterpret_cast<const float2*>(bias + f);
                    float2 w2[D];
#pragma unroll
                    for (int k = 0; k < D; ++k) w2[k] = *reinterpret_cast<const float2*>(W0t + k * 128 + f);
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        float st[R], ab[R];
                        float z0 = e ? b2.y : b2.x;
#pragma unroll
                        for (int k = 0; k < D; ++k) {
                            const float wk = e ? w2[k].y : w2[k].x;
                            z0 = fmaf(wk, xin[k], z0);
                            st[1 + k] = wk;
                        }
                        st[1 + D] = 0.f;
                        if (ACT == ACT_TANH || ACT == ACT_SHIFTED_TANH) {
                            st[0] = tanh_tc(z0);
                        } else {
                            float zz[R];
#pragma unroll
                            for (int s = 1; s < R; ++s) zz[s] = st[s];
                            zz[0] = z0;
                            act_jet_fwd<D, JET_LAPLACIAN>(zz, act, st[0]);
                        }
#pragma unroll
                        for (int s = 0; s < R; ++s) ab[s] = __uint_as_float(s < 2 ? LoHi[2 * (s & 1) + e] : HiHi[2 * (s & 1) + e]);
                        tc_act_bwd<D, ACT>(st, ab, act);
#pragma unroll
                        for (int k = 0; k < D; ++k) dw0_acc[c][e][k] += fmaf(ab[0], xin[k], ab[1 + k]);
                        dw0_acc[c][e][D] += ab[0];
                    }
                }
        };
        if (iters > 0) {
#pragma unroll 1
            for (int t = 0; t < 2; ++t) stash_layer(0, L - 1, t);
        }
        for (int it = 0; it < iters; ++it) {
#pragma unroll 1
            for (int l = L - 2; l >= 1; --l) {
#pragma unroll 1
                for (int t = 0; t < 2; ++t) stash_layer(it, l, t);
            }
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                if (it + 1 < iters) stash_layer(it + 1, L - 1, t);
                input_layer(it, t);
            }
        }
        // flush the output-layer gradient: sum over the 8 points of the warp, one atomic per feature
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                float v = dwout_acc[c][j];
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                if (pl == 0) atomicAdd(&small_acc[(D + L) * 128 + 32 * c + fb + j], v);
            }
        // input-layer gradient: sum over the 8 points of the warp, then over the 4 quadrant warps of a feature group through
        // shared memory (slot q), one global atomic per element and CTA (rows 0..D-1: dW_0^T, row D: b_0)
#pragma unroll
        for (int c = 0; c < 4; ++c)
#pragma unroll
            for (int e = 0; e < 2; ++e)
#pragma unroll
                for (int k = 0; k <= D; ++k) {
                    float v = dw0_acc[c][e][k];
                    v += __shfl_xor_sync(0xffffffffu, v, 4);
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    if (pl == 0) dw0s[(size_t)(k * 128 + 32 * c + fb + e) * kDw0Stride + q] = v;
                }
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiWarps * 32) : "memory");
        for (int i = tid; i < (D + 1) * 128; i += kEpiWarps * 32) {
            const float* dq = dw0s + (size_t)i * kDw0Stride;
            atomicAdd(&small_acc[i], (dq[0] + dq[1]) + (dq[2] + dq[3]));
        }
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 4);
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 8);
        dbout_acc += __shfl_xor_sync(0xffffffffu, dbout_acc, 16);
        if (lane == 0 && sub == 0) atomicAdd(&small_acc[(D + L + 1) * 128], dbout_acc);
    } else if (w == kEpiWarps && cta == 0 && lane == 0) {
        mma2_issue<kBfRev>(tb, smem_u32(wimg_s), nmat, true, iters, a_ready, a_free, d_full);
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync();
    if (w == kEpiWarps) tmem_dealloc<2>(tb, 512);
}

template <int D, int ACT = -1>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTcThreads, 1)
jet_rev2_tc_kernel(TcParams P, const float* __restrict__ x, long long n, const float* __restrict__ u_bar,
                   const float* __restrict__ g_bar, const float* __restrict__ h_bar, const float* __restrict__ stash,
                   float* __restrict__ zb, float* __restrict__ small_acc, long long stash_ls, int nblk8, int num_pairs, int pf_dist) {
    rev2_role<D, ACT, false>(P, x, n, u_bar, g_bar, h_bar, stash, zb, small_acc, stash_ls, nblk8, num_pairs, (int)(blockIdx.x >> 1),
                             (int)(gridDim.x >> 1), RingCtl{});
}

template <int D, int ACT = -1>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTcThreads, 1)
jet_rev_tc_kernel(TcParams P, const float* __restrict__ x, long long n, const float* __restrict__ u_bar,
                  const float* __restrict__ g_bar, const float* __restrict__ h_bar, const float* __restrict__ stash,
                  float* __restrict__ zb, float* __restrict__ small_acc, long long stash_ls, int nblk8, int num_pairs, int pf_dist) {
    rev_role<D, ACT, false>(P, x, n, u_bar, g_bar, h_bar, stash, zb, small_acc, stash_ls, nblk8, num_pairs, pf_dist, (int)(blockIdx.x >> 1),
                            (int)(gridDim.x >> 1), RingCtl{});
}

}  // namespace tcp
}  // namespace gpe

#include "dw_tc_kernel.cuh"
