// Forward ("jet") and reverse kernels of the fused Taylor-mode MLP -- sm_100a.
// Replaces, for one tile of collocation points at a time and without ever writing an [N,W]
// activation to HBM:
//   * nn.Sequential forward               (NBroot cell 6:30-54, S/gross_pitaevskii_2D.py:59-81, S/helmholtz_2D.py:40-56)
//   * the nested autograd.grad derivatives (NBroot cell 6:90-104, S/gross_pitaevskii_2D.py:183-188)
//   * loss.backward() through that graph   (NBroot cell 10:100, S/gross_pitaevskii_2D.py:376,422)
#pragma once
#include "jet_common.cuh"

namespace gpe {

template <int D, int JET, int WP>
struct TileCfg {
    static constexpr int R = jet_rows(D, JET);
    static constexpr int H = jet_hess(D, JET);
    static constexpr int TP = jet_tp(R);
    static constexpr int TR = R * TP;
    static constexpr int TX = WP / 8;
    static constexpr int TY = kConsumers / TX;
    static constexpr int M = TY * TR;      // rows per CTA tile
    static constexpr int PTS = TY * TP;    // points per CTA tile
    static constexpr int LDA = WP + kPad;
    static constexpr int STASH_F4 = TR * 2;  // float4 per thread per layer
    __host__ __device__ static size_t smem_fwd(int L, bool res = false) {
        return (size_t)M * LDA * 4 + (size_t)kStages * KC * WP * 4 + ((size_t)(D + L + 1) * WP + 4) * 4 + 2 * kStages * 8 +
               (res ? STASH_F4 * kConsumers * 16 : 0);
    }
    __host__ __device__ static size_t smem_bwd(int L, bool res = false) {
        return smem_fwd(L, false) + (size_t)M * LDA * 4 + ((size_t)(D + L + 1) * WP + 4) * 4 + (res ? STASH_F4 * kConsumers * 16 : 0);
    }
    __host__ __device__ static size_t stash_floats_per_cta(int L) { return (size_t)L * STASH_F4 * 4 * kConsumers; }
    // residual nets (smem_*(L, true)): a thread-private copy of the tile -- the block input going forward, zbar of lin2
    // going back
};

// thread-private, conflict-free [TR*2][kConsumers] float4 image of a thread's TR x 8 tile
template <int TR>
__device__ __forceinline__ void skip_save(float4* sk, const float (&v)[TR][8]) {
#pragma unroll
    for (int r = 0; r < TR; ++r) {
        sk[(size_t)(2 * r) * kConsumers] = make_float4(v[r][0], v[r][1], v[r][2], v[r][3]);
        sk[(size_t)(2 * r + 1) * kConsumers] = make_float4(v[r][4], v[r][5], v[r][6], v[r][7]);
    }
}
template <int TR>
__device__ __forceinline__ void skip_add(const float4* sk, float (&v)[TR][8]) {
#pragma unroll
    for (int r = 0; r < TR; ++r) {
        const float4 a = sk[(size_t)(2 * r) * kConsumers], b = sk[(size_t)(2 * r + 1) * kConsumers];
        v[r][0] += a.x, v[r][1] += a.y, v[r][2] += a.z, v[r][3] += a.w;
        v[r][4] += b.x, v[r][5] += b.y, v[r][6] += b.z, v[r][7] += b.w;
    }
}

__device__ __forceinline__ int col_of(int tx, int j, int WP) { return j < 4 ? 4 * tx + j : WP / 2 + 4 * tx + (j - 4); }

// layer 0: z0 = W0 x + b0, z1_i = W0[:,i], z2 = 0
template <int D, int JET, int WP, int TP, int R>
__device__ __forceinline__ void first_layer(float (&acc)[TP * R][8], const float (&xin)[TP][D], const float* W0t,
                                            const float* bias, int tx) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = col_of(tx, j, WP);
        float w[D];
#pragma unroll
        for (int i = 0; i < D; ++i) w[i] = W0t[i * WP + c];
        const float b = bias[c];
#pragma unroll
        for (int tp = 0; tp < TP; ++tp) {
            float z = b;
#pragma unroll
            for (int i = 0; i < D; ++i) z = fmaf(w[i], xin[tp][i], z);
            acc[tp * R][j] = z;
            if (JET != JET_VALUE) {
#pragma unroll
                for (int i = 0; i < D; ++i) acc[tp * R + 1 + i][j] = w[i];
#pragma unroll
                for (int s = 1 + D; s < R; ++s) acc[tp * R + s][j] = 0.f;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// forward: per point u, grad u, second derivatives
// ------------------------------------------------------------------------------------------
template <int D, int JET, int WP, bool RES = false>
__global__ void __launch_bounds__(kThreads)
jet_fwd_kernel(MlpParams P, const float* __restrict__ x, long long n, XForm xf, float* __restrict__ out_u,
               float* __restrict__ out_g, float* __restrict__ out_h, int num_tiles) {
    using C = TileCfg<D, JET, WP>;
    constexpr int R = C::R, TP = C::TP, TR = C::TR, TX = C::TX, H = C::H;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* As = reinterpret_cast<float*>(smem_raw);
    float* stages = As + (size_t)C::M * C::LDA;
    float* small = stages + (size_t)kStages * KC * WP;
    const int L = P.n_hidden;
    const int n_small = (D + L + 1) * WP + 4;
    uint64_t* bars = reinterpret_cast<uint64_t*>(small + n_small);
    float4* skip = reinterpret_cast<float4*>(bars + 2 * kStages) + threadIdx.x;

    const int tid = threadIdx.x;
    WeightPipe<WP> pipe{stages, bars, bars + kStages, 0u};
    if (tid == 0) pipe.init_barriers();
    for (int i = tid; i < n_small; i += kThreads) small[i] = P.small[i];
    __syncthreads();

    if (tid >= kConsumers) {  // ---- producer warp: stream the hidden weight matrices
        if (tid == kConsumers)
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x)
                for (int l = 1; l < L; ++l) pipe.produce_matrix(P.wt_fwd + (size_t)(l - 1) * WP * WP);
        return;
    }

    const int tx = tid % TX, ty = tid / TX, row0 = ty * TR;
    const float* W0t = small;
    const float* bias = small + D * WP;
    const float* wout = small + (D + L) * WP;
    const float bout = small[(D + L + 1) * WP];
    const int act = P.act;

    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const long long p0 = (long long)tile * C::PTS + ty * TP;
        float xin[TP][D];
#pragma unroll
        for (int tp = 0; tp < TP; ++tp) {
            const long long p = p0 + tp;
#pragma unroll
            for (int i = 0; i < D; ++i) xin[tp][i] = p < n ? fmaf(xf.a[i], x[p * D + i], xf.b[i]) : 0.f;
        }
        float acc[TR][8];
        first_layer<D, JET, WP, TP, R>(acc, xin, W0t, bias, tx);
        for (int l = 0;; ++l) {
            if (RES && !(l & 1) && l >= 2) skip_add<TR>(skip, acc);  // z_l += a_{l-2}
            const int act_l = RES ? layer_act(ACT_RESBLOCK_TANH, l) : act;
            // activation jets, thread-local
#pragma unroll
            for (int tp = 0; tp < TP; ++tp)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float z[R];
#pragma unroll
                    for (int s = 0; s < R; ++s) z[s] = acc[tp * R + s][j];
                    float st0;
                    act_jet_fwd<D, JET>(z, act_l, st0);
#pragma unroll
                    for (int s = 0; s < R; ++s) acc[tp * R + s][j] = z[s];
                }
            if (l == L - 1) break;
            if (RES && !(l & 1)) skip_save<TR>(skip, acc);  // block input
            consumer_sync();  // everybody finished reading As (previous layer / previous tile)
            store_tile<TR, WP>(As, row0, tx, acc);
            consumer_sync();
            gemm_rows<TR, WP>(acc, As, row0, tx, pipe);
            const float* bl = bias + (l + 1) * WP;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float b = bl[col_of(tx, j, WP)];
#pragma unroll
                for (int tp = 0; tp < TP; ++tp) acc[tp * R][j] += b;
            }
        }
        // output layer: one dot product per row, reduced over the TX lanes that share the rows
        float o[TR];
#pragma unroll
        for (int r = 0; r < TR; ++r) {
            float s = 0.f;
#pragma unroll
            for (int j = 0; j < 8; ++j) s = fmaf(acc[r][j], wout[col_of(tx, j, WP)], s);
            o[r] = s;
        }
#pragma unroll
        for (int off = TX / 2; off >= 1; off >>= 1)
#pragma unroll
            for (int r = 0; r < TR; ++r) o[r] += __shfl_xor_sync(0xffffffffu, o[r], off);
        if (tx == 0) {
#pragma unroll
            for (int tp = 0; tp < TP; ++tp) {
                const long long p = p0 + tp;
                if (p < n) {
                    out_u[p] = o[tp * R] + bout;
                    if (JET != JET_VALUE) {
#pragma unroll
                        for (int i = 0; i < D; ++i) out_g[p * D + i] = o[tp * R + 1 + i];
#pragma unroll
                        for (int k = 0; k < H; ++k) out_h[p * H + k] = o[tp * R + 1 + D + k];
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// reverse: parameter gradients of  sum_n ( u_bar u + <g_bar, grad u> + <h_bar, second derivs> )
// The forward is recomputed per tile; pre-activation jets are stashed in a per-CTA, L2-resident
// scratch (thread-private, coalesced float4).  dW of the hidden layers accumulates in a per-CTA
// global accumulator (vector RED), reduced over CTAs by finish_grads_kernel.
// ------------------------------------------------------------------------------------------
template <int WP, int M>
__device__ __forceinline__ void dw_gemm_accumulate(const float* __restrict__ Zb, const float* __restrict__ As,
                                                   float* __restrict__ dWl, int tid) {
    constexpr int TX = WP / 8, TY = kConsumers / TX, LDA = WP + kPad;
    constexpr int OT = (WP / TY) < 8 ? (WP / TY) : 8;
    constexpr int PASSES = WP / (TY * OT);
    const int txi = tid % TX, tyo = tid / TX;
#pragma unroll 1
    for (int pass = 0; pass < PASSES; ++pass) {
        const int o0 = (pass * TY + tyo) * OT;
        float acc[OT][8];
#pragma unroll
        for (int oo = 0; oo < OT; ++oo)
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[oo][c] = 0.f;
#pragma unroll 2
        for (int row = 0; row < M; ++row) {
            float zo[OT];
#pragma unroll
            for (int q = 0; q < OT / 4; ++q) {
                const float4 v = *reinterpret_cast<const float4*>(Zb + row * LDA + o0 + 4 * q);
                zo[4 * q] = v.x, zo[4 * q + 1] = v.y, zo[4 * q + 2] = v.z, zo[4 * q + 3] = v.w;
            }
            const float4 a0 = *reinterpret_cast<const float4*>(As + row * LDA + 4 * txi);
            const float4 a1 = *reinterpret_cast<const float4*>(As + row * LDA + WP / 2 + 4 * txi);
#pragma unroll
            for (int oo = 0; oo < OT; ++oo) {
                acc[oo][0] = fmaf(zo[oo], a0.x, acc[oo][0]);
                acc[oo][1] = fmaf(zo[oo], a0.y, acc[oo][1]);
                acc[oo][2] = fmaf(zo[oo], a0.z, acc[oo][2]);
                acc[oo][3] = fmaf(zo[oo], a0.w, acc[oo][3]);
                acc[oo][4] = fmaf(zo[oo], a1.x, acc[oo][4]);
                acc[oo][5] = fmaf(zo[oo], a1.y, acc[oo][5]);
                acc[oo][6] = fmaf(zo[oo], a1.z, acc[oo][6]);
                acc[oo][7] = fmaf(zo[oo], a1.w, acc[oo][7]);
            }
        }
#pragma unroll
        for (int oo = 0; oo < OT; ++oo) {
            float* dst = dWl + (size_t)(o0 + oo) * WP;
            atomicAdd(reinterpret_cast<float4*>(dst + 4 * txi), make_float4(acc[oo][0], acc[oo][1], acc[oo][2], acc[oo][3]));
            atomicAdd(reinterpret_cast<float4*>(dst + WP / 2 + 4 * txi),
                      make_float4(acc[oo][4], acc[oo][5], acc[oo][6], acc[oo][7]));
        }
    }
}

template <int D, int JET, int WP, bool RES = false>
__global__ void __launch_bounds__(kThreads)
jet_bwd_kernel(MlpParams P, const float* __restrict__ x, long long n, XForm xf, const float* __restrict__ u_bar,
               const float* __restrict__ g_bar, const float* __restrict__ h_bar, float* __restrict__ stash_all,
               float* __restrict__ dWacc_all, float* __restrict__ small_acc, int num_tiles) {
    using C = TileCfg<D, JET, WP>;
    constexpr int R = C::R, TP = C::TP, TR = C::TR, TX = C::TX, H = C::H, LDA = C::LDA;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* As = reinterpret_cast<float*>(smem_raw);
    float* stages = As + (size_t)C::M * LDA;
    float* small = stages + (size_t)kStages * KC * WP;
    const int L = P.n_hidden;
    const int n_small = (D + L + 1) * WP + 4;
    uint64_t* bars = reinterpret_cast<uint64_t*>(small + n_small);
    float* Zb = reinterpret_cast<float*>(bars + 2 * kStages);
    float* gacc = Zb + (size_t)C::M * LDA;  // same layout as `small`: dW0t[D][WP], db[L][WP], dwout[WP], dbout
    float4* skip = reinterpret_cast<float4*>(gacc + n_small) + threadIdx.x;

    const int tid = threadIdx.x;
    WeightPipe<WP> pipe{stages, bars, bars + kStages, 0u};
    if (tid == 0) pipe.init_barriers();
    for (int i = tid; i < n_small; i += kThreads) {
        small[i] = P.small[i];
        gacc[i] = 0.f;
    }
    __syncthreads();

    if (tid >= kConsumers) {
        if (tid == kConsumers)
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                for (int l = 1; l < L; ++l) pipe.produce_matrix(P.wt_fwd + (size_t)(l - 1) * WP * WP);
                for (int l = L - 1; l >= 1; --l) pipe.produce_matrix(P.w_bwd + (size_t)(l - 1) * WP * WP);
            }
        return;
    }

    const int tx = tid % TX, ty = tid / TX, row0 = ty * TR;
    const float* W0t = small;
    const float* bias = small + D * WP;
    const float* wout = small + (D + L) * WP;
    const int act = P.act;
    float4* stash = reinterpret_cast<float4*>(stash_all + (size_t)blockIdx.x * C::stash_floats_per_cta(L));
    float* dWacc = dWacc_all + (size_t)blockIdx.x * (size_t)(L - 1) * WP * WP;
    float dwout_acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) dwout_acc[j] = 0.f;
    float dbout_acc = 0.f;

    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const long long p0 = (long long)tile * C::PTS + ty * TP;
        float xin[TP][D];
#pragma unroll
        for (int tp = 0; tp < TP; ++tp) {
            const long long p = p0 + tp;
#pragma unroll
            for (int i = 0; i < D; ++i) xin[tp][i] = p < n ? fmaf(xf.a[i], x[p * D + i], xf.b[i]) : 0.f;
        }
        float acc[TR][8];
        // ---------------- forward recompute with stash ----------------
        first_layer<D, JET, WP, TP, R>(acc, xin, W0t, bias, tx);
        for (int l = 0;; ++l) {
            if (RES && !(l & 1) && l >= 2) skip_add<TR>(skip, acc);  // z_l += a_{l-2}
            const int act_l = RES ? layer_act(ACT_RESBLOCK_TANH, l) : act;
            float4* sl = stash + (size_t)l * C::STASH_F4 * kConsumers + tid;
#pragma unroll
            for (int tp = 0; tp < TP; ++tp)
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    float sv[R][4];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        const int j = hf * 4 + jj;
                        float z[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) z[s] = acc[tp * R + s][j];
#pragma unroll
                        for (int s = 1; s < R; ++s) sv[s][jj] = z[s];
                        act_jet_fwd<D, JET>(z, act_l, sv[0][jj]);
#pragma unroll
                        for (int s = 0; s < R; ++s) acc[tp * R + s][j] = z[s];
                    }
#pragma unroll
                    for (int s = 0; s < R; ++s)
                        sl[(size_t)((tp * R + s) * 2 + hf) * kConsumers] = make_float4(sv[s][0], sv[s][1], sv[s][2], sv[s][3]);
                }
            if (l == L - 1) break;
            if (RES && !(l & 1)) skip_save<TR>(skip, acc);  // block input
            consumer_sync();
            store_tile<TR, WP>(As, row0, tx, acc);
            consumer_sync();
            gemm_rows<TR, WP>(acc, As, row0, tx, pipe);
            const float* bl = bias + (l + 1) * WP;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float b = bl[col_of(tx, j, WP)];
#pragma unroll
                for (int tp = 0; tp < TP; ++tp) acc[tp * R][j] += b;
            }
        }
        // ---------------- output layer adjoint ----------------
        {
            float adj[TR];
#pragma unroll
            for (int tp = 0; tp < TP; ++tp) {
                const long long p = p0 + tp;
                const bool ok = p < n;
                adj[tp * R] = (ok && u_bar) ? u_bar[p] : 0.f;
                if (JET != JET_VALUE) {
#pragma unroll
                    for (int i = 0; i < D; ++i) adj[tp * R + 1 + i] = (ok && g_bar) ? g_bar[p * D + i] : 0.f;
#pragma unroll
                    for (int k = 0; k < H; ++k) adj[tp * R + 1 + D + k] = (ok && h_bar) ? h_bar[p * H + k] : 0.f;
                }
                if (tx == 0) dbout_acc += adj[tp * R];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float w = wout[col_of(tx, j, WP)];
                float s = dwout_acc[j];
#pragma unroll
                for (int r = 0; r < TR; ++r) {
                    s = fmaf(adj[r], acc[r][j], s);
                    acc[r][j] = adj[r] * w;   // abar of the last hidden layer
                }
                dwout_acc[j] = s;
            }
        }
        // ---------------- hidden layers, top down ----------------
        for (int l = L - 1;; --l) {
            if (RES && !(l & 1) && l < L - 1) skip_add<TR>(skip, acc);  // abar_l += zbar_{l+2} (the skip)
            const int act_l = RES ? layer_act(ACT_RESBLOCK_TANH, l) : act;
            // zbar from abar (in place) using the stash of layer l
            const float4* sl = stash + (size_t)l * C::STASH_F4 * kConsumers + tid;
#pragma unroll
            for (int tp = 0; tp < TP; ++tp)
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    float4 sv[R];
#pragma unroll
                    for (int s = 0; s < R; ++s) sv[s] = __ldcg(sl + (size_t)((tp * R + s) * 2 + hf) * kConsumers);
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        const int j = hf * 4 + jj;
                        float st[R], ab[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) {
                            st[s] = jj == 0 ? sv[s].x : jj == 1 ? sv[s].y : jj == 2 ? sv[s].z : sv[s].w;
                            ab[s] = acc[tp * R + s][j];
                        }
                        act_jet_bwd<D, JET>(st, ab, act_l);
#pragma unroll
                        for (int s = 0; s < R; ++s) acc[tp * R + s][j] = ab[s];
                    }
                }
            if (RES && !(l & 1) && l >= 2) skip_save<TR>(skip, acc);  // zbar_l also flows to abar_{l-2}
            // bias gradient (+ first-layer weight gradient)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int c = col_of(tx, j, WP);
                float s = 0.f;
#pragma unroll
                for (int tp = 0; tp < TP; ++tp) s += acc[tp * R][j];
                atomicAdd(&gacc[(D + l) * WP + c], s);
                if (l == 0) {
#pragma unroll
                    for (int i = 0; i < D; ++i) {
                        float w = 0.f;
#pragma unroll
                        for (int tp = 0; tp < TP; ++tp) {
                            w = fmaf(acc[tp * R][j], xin[tp][i], w);
                            if (JET != JET_VALUE) w += acc[tp * R + 1 + i][j];
                        }
                        atomicAdd(&gacc[i * WP + c], w);
                    }
                }
            }
            if (l == 0) break;
            consumer_sync();  // previous readers of Zb / As are done
            store_tile<TR, WP>(Zb, row0, tx, acc);
            // a^{l-1} from the stash of layer l-1 -> As
            const float4* sp = stash + (size_t)(l - 1) * C::STASH_F4 * kConsumers + tid;
#pragma unroll
            for (int tp = 0; tp < TP; ++tp)
#pragma unroll
                for (int hf = 0; hf < 2; ++hf) {
                    float4 sv[R];
#pragma unroll
                    for (int s = 0; s < R; ++s) sv[s] = __ldcg(sp + (size_t)((tp * R + s) * 2 + hf) * kConsumers);
                    float av[R][4];
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) {
                        float st[R], a[R];
#pragma unroll
                        for (int s = 0; s < R; ++s) st[s] = jj == 0 ? sv[s].x : jj == 1 ? sv[s].y : jj == 2 ? sv[s].z : sv[s].w;
                        act_jet_from_stash<D, JET>(st, a, RES ? layer_act(ACT_RESBLOCK_TANH, l - 1) : act);
#pragma unroll
                        for (int s = 0; s < R; ++s) av[s][jj] = a[s];
                    }
#pragma unroll
                    for (int s = 0; s < R; ++s)
                        *reinterpret_cast<float4*>(As + (size_t)(row0 + tp * R + s) * LDA + (hf ? WP / 2 : 0) + 4 * tx) =
                            make_float4(av[s][0], av[s][1], av[s][2], av[s][3]);
                }
            consumer_sync();
            dw_gemm_accumulate<WP, C::M>(Zb, As, dWacc + (size_t)(l - 1) * WP * WP, tid);
            gemm_rows<TR, WP>(acc, Zb, row0, tx, pipe);  // abar^{l-1} = zbar^l W_l
        }
    }
    // ---------------- flush the small accumulators ----------------
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(&gacc[(D + L) * WP + col_of(tx, j, WP)], dwout_acc[j]);
    if (tx == 0) atomicAdd(&gacc[(D + L + 1) * WP], dbout_acc);
    consumer_sync();
    for (int i = tid; i < n_small; i += kConsumers) {
        const float v = gacc[i];
        if (v != 0.f) atomicAdd(&small_acc[i], v);
    }
}

}  // namespace gpe
