// Shared device code for the Taylor-mode ("jet") MLP kernels -- sm_100a only.
//
// The reference evaluates u, grad u and the Laplacian with nested torch.autograd.grad calls
// (Gross_Pitaevskii_1D_power_Test.ipynb cell 6:90-104, Gross-Pitaevskii/src/gross_pitaevskii_2D.py:183-188,
// Gross-Pitaevskii/src/helmholtz_2D.py:78-83).  Here every collocation point carries R "streams"
// (value, first derivatives, second derivatives) through the tanh MLP in one sweep
// (SURVEY.md App. B).  A linear layer acts on all streams with the same weight matrix, so a
// tile of points is one [rows x W] x [W x W] contraction with rows = points x streams.
//
// CTA layout: 4 consumer warps (128 threads) + 1 producer warp.
//   * consumer thread (ty, tx): rows  ty*TR .. ty*TR+TR-1  (TP points x R streams),
//                               cols  4*tx..4*tx+3  and  WP/2+4*tx..WP/2+4*tx+3
//   * activations live in shared memory row-major [row][WP+4]; the accumulator tile
//     TR x 8 lives in registers; the activation jets are thread-local.
//   * the producer warp streams the weight matrix of the current layer from L2 in KC-row
//     chunks with cp.async.bulk (TMA bulk copy, SASS UBLKCP) through a kStages-deep
//     mbarrier ring.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gpe {

constexpr int kConsumers = 128;            // consumer threads per CTA
constexpr int kThreads = kConsumers + 32;  // + producer warp
constexpr int KC = 16;                     // weight rows (k) per pipeline stage
constexpr int kStages = 4;
constexpr int kMaxLayers = 8;              // hidden layers
constexpr int kPad = 4;                    // smem row padding (floats)

// ACT_RESBLOCK_TANH: the residual network of R/box_to_gaussian_pinn_simulation.py:52-62,100-130 --
// layer 0 ShiftedTanh, then blocks tanh(lin2(tanh(lin1(a))) + a): hidden layers 2k+1 / 2k+2 are lin1 / lin2 of block
// k, all tanh, and the jets of a_{2k} are added to the pre-activation jets of layer 2k+2 (the skip is linear, so
// every stream passes through it unchanged).
enum Act { ACT_TANH = 0, ACT_SHIFTED_TANH = 1, ACT_LEAKY_RELU = 2, ACT_RESBLOCK_TANH = 3 };
__host__ __device__ constexpr int layer_act(int act, int l) {
    return act == ACT_RESBLOCK_TANH ? (l == 0 ? ACT_SHIFTED_TANH : ACT_TANH) : act;
}
enum Jet { JET_VALUE = 0, JET_DIAG = 1, JET_FULL = 2, JET_LAPLACIAN = 3 };

// number of streams per point
__host__ __device__ constexpr int jet_rows(int d, int jet) {
    return jet == JET_VALUE ? 1 : jet == JET_DIAG ? 1 + 2 * d : jet == JET_FULL ? (d == 2 ? 6 : 3) : (1 + d + 1);
}
// second-derivative outputs per point
__host__ __device__ constexpr int jet_hess(int d, int jet) {
    return jet == JET_VALUE ? 0 : jet == JET_DIAG ? d : jet == JET_FULL ? (d == 2 ? 3 : 1) : 1;
}
// points per thread
__host__ __device__ constexpr int jet_tp(int rows) { return rows == 1 ? 8 : rows == 3 ? 4 : rows == 4 ? 3 : 2; }

struct MlpParams {
    int n_hidden;          // L
    int width;             // W (true hidden width, <= WP)
    int act;
    const float* wt_fwd;   // [L-1][WP][WP]  wt[l][k=in][n=out]
    const float* w_bwd;    // [L-1][WP][WP]  w[l][k=out][n=in]
    const float* small;    // W0t[D][WP], b[L][WP], wout[WP], bout[4]
};

struct XForm {             // x_eff = a * x + b  per input coordinate (mirror points)
    float a[2], b[2];
};

// ------------------------------------------------------------------------------------------
// PTX helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t done;
    const uint32_t addr = smem_u32(bar);
    do {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    } while (!done);
}
// TMA bulk copy global -> shared, completion signalled on an mbarrier (SASS: UBLKCP)
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumers) : "memory"); }

// ------------------------------------------------------------------------------------------
// weight pipeline: a ring of kStages chunks [KC][WP]
// ------------------------------------------------------------------------------------------
template <int WP>
struct WeightPipe {
    float* stages;       // kStages * KC * WP floats
    uint64_t* full;      // [kStages]
    uint64_t* empty;     // [kStages]
    uint32_t it;         // running chunk counter (same sequence on both sides)

    static constexpr uint32_t kChunkBytes = KC * WP * sizeof(float);
    static constexpr int kChunksPerMatrix = WP / KC;

    __device__ __forceinline__ void init_barriers() {   // one thread
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kConsumers / 32);
        }
        mbar_fence_init();
    }
    // producer (one lane): stream one padded WP x WP matrix
    __device__ __forceinline__ void produce_matrix(const float* gmat) {
        for (int c = 0; c < kChunksPerMatrix; ++c, ++it) {
            const uint32_t s = it % kStages, ph = (it / kStages) & 1u;
            mbar_wait(&empty[s], ph ^ 1u);
            mbar_arrive_expect_tx(&full[s], kChunkBytes);
            tma_bulk_g2s(stages + (size_t)s * KC * WP, gmat + (size_t)c * KC * WP, kChunkBytes, &full[s]);
        }
    }
    __device__ __forceinline__ const float* consumer_acquire() {
        const uint32_t s = it % kStages, ph = (it / kStages) & 1u;
        mbar_wait(&full[s], ph);
        return stages + (size_t)s * KC * WP;
    }
    __device__ __forceinline__ void consumer_release() {
        const uint32_t s = it % kStages;
        __syncwarp();
        if ((threadIdx.x & 31) == 0) mbar_arrive(&empty[s]);
        ++it;
    }
};

// ------------------------------------------------------------------------------------------
// register-blocked FFMA contraction:  acc[TR][8] = A[rows][0..WP) * B[0..WP)[cols]
//   A: shared [row][WP+kPad] (row-major, k contiguous), B streamed in [KC][WP] chunks
// ------------------------------------------------------------------------------------------
template <int TR, int WP>
__device__ __forceinline__ void gemm_rows(float (&acc)[TR][8], const float* __restrict__ As, int row0, int tx,
                                          WeightPipe<WP>& pipe) {
    constexpr int LDA = WP + kPad;
#pragma unroll
    for (int r = 0; r < TR; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) acc[r][c] = 0.f;
    const float* arow = As + (size_t)row0 * LDA;
#pragma unroll 1
    for (int ch = 0; ch < WP / KC; ++ch) {
        const float* Bs = pipe.consumer_acquire();
#pragma unroll 1
        for (int kk = 0; kk < KC; kk += 4) {
            float4 a[TR];
#pragma unroll
            for (int r = 0; r < TR; ++r) a[r] = *reinterpret_cast<const float4*>(arow + r * LDA + ch * KC + kk);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 b0 = *reinterpret_cast<const float4*>(Bs + (kk + j) * WP + 4 * tx);
                const float4 b1 = *reinterpret_cast<const float4*>(Bs + (kk + j) * WP + WP / 2 + 4 * tx);
#pragma unroll
                for (int r = 0; r < TR; ++r) {
                    const float av = j == 0 ? a[r].x : j == 1 ? a[r].y : j == 2 ? a[r].z : a[r].w;
                    acc[r][0] = fmaf(av, b0.x, acc[r][0]);
                    acc[r][1] = fmaf(av, b0.y, acc[r][1]);
                    acc[r][2] = fmaf(av, b0.z, acc[r][2]);
                    acc[r][3] = fmaf(av, b0.w, acc[r][3]);
                    acc[r][4] = fmaf(av, b1.x, acc[r][4]);
                    acc[r][5] = fmaf(av, b1.y, acc[r][5]);
                    acc[r][6] = fmaf(av, b1.z, acc[r][6]);
                    acc[r][7] = fmaf(av, b1.w, acc[r][7]);
                }
            }
        }
        pipe.consumer_release();
    }
}

// store a thread's TR x 8 register tile to shared [row][WP+kPad]
template <int TR, int WP>
__device__ __forceinline__ void store_tile(float* __restrict__ S, int row0, int tx, const float (&v)[TR][8]) {
    constexpr int LDA = WP + kPad;
#pragma unroll
    for (int r = 0; r < TR; ++r) {
        float* p = S + (size_t)(row0 + r) * LDA;
        *reinterpret_cast<float4*>(p + 4 * tx) = make_float4(v[r][0], v[r][1], v[r][2], v[r][3]);
        *reinterpret_cast<float4*>(p + WP / 2 + 4 * tx) = make_float4(v[r][4], v[r][5], v[r][6], v[r][7]);
    }
}

// ------------------------------------------------------------------------------------------
// activation jets (thread-local).  Stream order inside a point:
//   s=0 value; s=1..D first derivatives; then DIAG: D pure second derivatives,
//   FULL(D=2): xx, yy, xy;  LAPLACIAN: one stream carrying sum_i d_ii.
// ------------------------------------------------------------------------------------------
struct ActDeriv {
    float a0, s1, s2, s3;
};

__device__ __forceinline__ ActDeriv act_eval(float z0, int act, bool third) {
    ActDeriv o;
    if (act == ACT_LEAKY_RELU) {
        o.s1 = z0 > 0.f ? 1.f : 0.01f;
        o.a0 = z0 * o.s1;
        o.s2 = 0.f;
        o.s3 = 0.f;
    } else {
        const float t = tanhf(z0);
        o.s1 = fmaf(-t, t, 1.f);
        o.s2 = -2.f * t * o.s1;
        o.s3 = third ? -2.f * o.s1 * fmaf(-3.f * t, t, 1.f) : 0.f;
        o.a0 = act == ACT_SHIFTED_TANH ? t + 1.f : t;   // eps = 2.2e-16 vanishes in float32
    }
    return o;
}
// derivatives from the *stashed* activation value (tanh: t itself; leaky: a0, whose sign is z0's)
__device__ __forceinline__ ActDeriv act_from_stash(float st, int act) {
    ActDeriv o;
    if (act == ACT_LEAKY_RELU) {
        o.s1 = st > 0.f ? 1.f : 0.01f;
        o.a0 = st;
        o.s2 = 0.f;
        o.s3 = 0.f;
    } else {
        o.s1 = fmaf(-st, st, 1.f);
        o.s2 = -2.f * st * o.s1;
        o.s3 = -2.f * o.s1 * fmaf(-3.f * st, st, 1.f);
        o.a0 = act == ACT_SHIFTED_TANH ? st + 1.f : st;
    }
    return o;
}

// pre-activation jets z[R] -> post-activation jets (in place).  `stash0` receives the value to
// remember for the reverse sweep (t for tanh, a0 for leaky); derivative streams are stashed by
// the caller *before* this call (they are the pre-activation z).
template <int D, int JET>
__device__ __forceinline__ void act_jet_fwd(float* z, int act, float& stash0) {
    const ActDeriv a = act_eval(z[0], act, false);
    stash0 = act == ACT_LEAKY_RELU ? a.a0 : (act == ACT_SHIFTED_TANH ? a.a0 - 1.f : a.a0);
    if (JET != JET_VALUE) {
        if (JET == JET_DIAG) {
#pragma unroll
            for (int i = 0; i < D; ++i) z[1 + D + i] = fmaf(a.s1, z[1 + D + i], a.s2 * z[1 + i] * z[1 + i]);
        } else if (JET == JET_FULL) {
            if (D == 2) z[5] = fmaf(a.s1, z[5], a.s2 * z[1] * z[2]);
#pragma unroll
            for (int i = 0; i < D; ++i) z[1 + D + i] = fmaf(a.s1, z[1 + D + i], a.s2 * z[1 + i] * z[1 + i]);
        } else {  // LAPLACIAN
            float q = 0.f;
#pragma unroll
            for (int i = 0; i < D; ++i) q = fmaf(z[1 + i], z[1 + i], q);
            z[1 + D] = fmaf(a.s1, z[1 + D], a.s2 * q);
        }
#pragma unroll
        for (int i = 0; i < D; ++i) z[1 + i] *= a.s1;
    }
    z[0] = a.a0;
}

// stash {st0, z1.., z2..} -> post-activation jets a[R]
template <int D, int JET>
__device__ __forceinline__ void act_jet_from_stash(const float* st, float* a, int act) {
    const ActDeriv d = act_from_stash(st[0], act);
    a[0] = d.a0;
    if (JET != JET_VALUE) {
#pragma unroll
        for (int i = 0; i < D; ++i) a[1 + i] = d.s1 * st[1 + i];
        if (JET == JET_DIAG) {
#pragma unroll
            for (int i = 0; i < D; ++i) a[1 + D + i] = fmaf(d.s1, st[1 + D + i], d.s2 * st[1 + i] * st[1 + i]);
        } else if (JET == JET_FULL) {
#pragma unroll
            for (int i = 0; i < D; ++i) a[1 + D + i] = fmaf(d.s1, st[1 + D + i], d.s2 * st[1 + i] * st[1 + i]);
            if (D == 2) a[5] = fmaf(d.s1, st[5], d.s2 * st[1] * st[2]);
        } else {
            float q = 0.f;
#pragma unroll
            for (int i = 0; i < D; ++i) q = fmaf(st[1 + i], st[1 + i], q);
            a[1 + D] = fmaf(d.s1, st[1 + D], d.s2 * q);
        }
    }
}

// reverse of the activation jets: abar[R] (adjoint of post-activation) -> zbar[R] in place
template <int D, int JET>
__device__ __forceinline__ void act_jet_bwd(const float* st, float* ab, int act) {
    const ActDeriv d = act_from_stash(st[0], act);
    if (JET == JET_VALUE) {
        ab[0] *= d.s1;
        return;
    }
    float z0b = d.s1 * ab[0];
    if (JET == JET_DIAG || JET == JET_FULL) {
        float z1b[D];
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const float z1 = st[1 + i], z2 = st[1 + D + i], a1b = ab[1 + i], a2b = ab[1 + D + i];
            z0b = fmaf(d.s2 * z1, a1b, z0b);
            z0b = fmaf(fmaf(d.s3 * z1, z1, d.s2 * z2), a2b, z0b);
            z1b[i] = fmaf(d.s1, a1b, 2.f * d.s2 * z1 * a2b);
            ab[1 + D + i] = d.s1 * a2b;
        }
        if (JET == JET_FULL && D == 2) {
            const float axy = ab[5];
            z0b = fmaf(fmaf(d.s3 * st[1], st[2], d.s2 * st[5]), axy, z0b);
            z1b[0] = fmaf(d.s2 * st[2], axy, z1b[0]);
            z1b[D - 1] = fmaf(d.s2 * st[1], axy, z1b[D - 1]);
            ab[5] = d.s1 * axy;
        }
#pragma unroll
        for (int i = 0; i < D; ++i) ab[1 + i] = z1b[i];
    } else {  // LAPLACIAN
        const float a2b = ab[1 + D];
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < D; ++i) {
            const float z1 = st[1 + i];
            q = fmaf(z1, z1, q);
            z0b = fmaf(d.s2 * z1, ab[1 + i], z0b);
            ab[1 + i] = fmaf(d.s1, ab[1 + i], 2.f * d.s2 * z1 * a2b);
        }
        z0b = fmaf(fmaf(d.s3, q, d.s2 * st[1 + D]), a2b, z0b);
        ab[1 + D] = d.s1 * a2b;
    }
    ab[0] = z0b;
}

// shared-memory footprint helpers (host + device)
template <int WP>
struct SmemLayout {
    static constexpr int LDA = WP + kPad;
    __host__ __device__ static constexpr size_t act_bytes(int rows) { return (size_t)rows * LDA * sizeof(float); }
    __host__ __device__ static constexpr size_t stage_bytes() { return (size_t)kStages * KC * WP * sizeof(float); }
    __host__ __device__ static constexpr size_t small_floats(int d) { return (size_t)(d + kMaxLayers + 1) * WP + 4; }
};

}  // namespace gpe
