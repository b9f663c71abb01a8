// Fused loss epilogue over the per-point jets -- sm_100a.
//
// Replaces the elementwise + torch.mean/torch.sum part of the reference loss methods
//   pde_loss            NBroot cell 6:106-127, R/harmonic_pinn_simulation.py:174-196,
//                       S/gross_pitaevskii_1D_Different_Optimizers.py:318-346, S/gross_pitaevskii_2D.py:189-213
//   loss_pde            S/helmholtz_2D.py:84-89
//   riesz_loss          S/gross_pitaevskii_2D.py:139-151, S/gross_pitaevskii_1D_Different_Optimizers.py:266-282
//   normalization_loss  NBroot cell 6:157-162, S/gross_pitaevskii_1D_Different_Optimizers.py:423
//   symmetry_loss       NBroot cell 6:137-155, S/gross_pitaevskii_1D_Different_Optimizers.py:348-375
// by ONE pass that forms the full solution jets  u = base + s F(x) N(x)  per collocation point,
// the feature vector  f = [1, u, lap u, V, V u, nl(u), q]  and accumulates its Gram matrix
// sum_n f_i f_j (28 entries) plus sum |grad u|^2, sum N^2, sum (N -/+ N_mirror)^2 in fp64
// (per-thread fp64 accumulators -> warp shuffle -> shared memory -> one fp64 atomic per CTA
// and moment).  Every loss family of the reference -- Rayleigh or energy-quotient mu, the
// pointwise residual, the reference's [N]x[N,1] -> [N,N] broadcast, Riesz
// energy, normalisation, symmetry -- is a smooth function of those 31 numbers, evaluated on the
// host side in fp64; its derivative c = dL/dm comes back into gram_adjoints_kernel, which turns it
// into the per-point adjoints (u_bar, g_bar, h_bar) that feed the reverse jet kernel.
#include "../../include/gpe_b200.h"
#include "jet_host.h"

namespace gpe {
namespace {

constexpr int NF = GPE_NFEAT;       // 7
constexpr int NM = GPE_NMOM;        // 32 (31 used)
constexpr int kLossThreads = 256;

struct LossArgs {
    int d, nl_kind, bfactor, sym_sign, h_cols, lap_cols;
    float nl_power, scale;
    long long n;
    const float *x, *u, *g, *h, *base, *V, *q, *nm;
};

struct PointJets {      // full-solution jets and the pieces needed to chain back to the network jets
    double f[NF];
    double gx, gy, G;   // full gradient, |grad u|^2
    double n0, sd;      // raw network value, symmetry difference (N - sign * N_mirror)
    double F0, F1, F2;  // boundary factor and its derivatives (1-D)
    double dnl;         // d nl / d u
};

__device__ __forceinline__ double ipow(double b, int e) {
    double r = 1.0;
    while (e > 0) {
        if (e & 1) r *= b;
        b *= b;
        e >>= 1;
    }
    return r;
}

__device__ __forceinline__ void point_features(const LossArgs& A, long long p, PointJets& J) {
    const double s = (double)A.scale;
    const double n0 = (double)A.u[p];
    double n1x = 0.0, n1y = 0.0, nlap = 0.0;
    if (A.g) {
        n1x = (double)A.g[p * A.d];
        if (A.d == 2) n1y = (double)A.g[p * 2 + 1];
    }
    if (A.h)
        for (int i = 0; i < A.lap_cols; ++i) nlap += (double)A.h[p * A.h_cols + i];
    J.F0 = 1.0, J.F1 = 0.0, J.F2 = 0.0;
    if (A.bfactor == 1) {   // F = sin(pi x)   (R/box_pinn_simulation.py:119-130)
        const double pi = 3.14159265358979323846;
        const double xx = (double)A.x[p * A.d];
        double sn, cs;
        sincos(pi * xx, &sn, &cs);
        J.F0 = sn, J.F1 = pi * cs, J.F2 = -pi * pi * sn;
    }
    double u = s * J.F0 * n0;
    double gx = s * (J.F0 * n1x + J.F1 * n0);
    double gy = s * n1y;
    double lap = s * (J.F0 * nlap + 2.0 * J.F1 * n1x + J.F2 * n0);
    if (A.base) {           // caller-supplied base jets [n, 2+d] = (value, grad[d], laplacian)
        const float* b = A.base + p * (2 + A.d);
        u += (double)b[0];
        gx += (double)b[1];
        if (A.d == 2) gy += (double)b[2];
        lap += (double)b[1 + A.d];
    }
    const double V = A.V ? (double)A.V[p] : 0.0;
    const double q = A.q ? (double)A.q[p] : 0.0;
    double nl = 0.0, dnl = 0.0;
    if (A.nl_kind == 1) {           // u ** p with integer p (sign lost for even p, as in the reference)
        const int pw = (int)A.nl_power;
        const double um1 = pw >= 1 ? ipow(u, pw - 1) : 0.0;
        nl = pw >= 1 ? um1 * u : 1.0;
        dnl = pw * um1;
    } else if (A.nl_kind == 2) {    // |u|^(p-1) u
        const double au = fabs(u);
        const double pw = (double)A.nl_power;
        const double am1 = pw == 3.0 ? au * au : pow(au, pw - 1.0);
        nl = am1 * u;
        dnl = pw * am1;
    }
    J.f[0] = 1.0, J.f[1] = u, J.f[2] = lap, J.f[3] = V, J.f[4] = V * u, J.f[5] = nl, J.f[6] = q;
    J.gx = gx, J.gy = gy, J.G = gx * gx + gy * gy;
    J.n0 = n0;
    J.sd = A.nm ? n0 - (double)A.sym_sign * (double)A.nm[p] : 0.0;
    J.dnl = dnl;
}

__global__ void __launch_bounds__(kLossThreads) gram_moments_kernel(LossArgs A, double* __restrict__ m_out) {
    double acc[NM];
#pragma unroll
    for (int k = 0; k < NM; ++k) acc[k] = 0.0;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < A.n; p += (long long)gridDim.x * blockDim.x) {
        PointJets J;
        point_features(A, p, J);
        int k = 0;
#pragma unroll
        for (int i = 0; i < NF; ++i)
#pragma unroll
            for (int j = i; j < NF; ++j) acc[k++] += J.f[i] * J.f[j];
        acc[GPE_MOM_GRAD2] += J.G;
        acc[GPE_MOM_NN2] += J.n0 * J.n0;
        acc[GPE_MOM_SYM2] += J.sd * J.sd;
    }
    __shared__ double red[kLossThreads / 32][NM];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < NM; ++k) {
        double v = acc[k];
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        if (lane == 0) red[warp][k] = v;
    }
    __syncthreads();
    if (threadIdx.x < NM) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < kLossThreads / 32; ++w) v += red[w][threadIdx.x];
        if (v != 0.0) atomicAdd(&m_out[threadIdx.x], v);
    }
}

__global__ void __launch_bounds__(kLossThreads)
gram_adjoints_kernel(LossArgs A, const double* __restrict__ c, float* __restrict__ u_bar, float* __restrict__ g_bar,
                     float* __restrict__ h_bar, float* __restrict__ nm_bar) {
    __shared__ double S[NF][NF];
    __shared__ double cx[3];
    if (threadIdx.x < NF * NF) {
        const int i = threadIdx.x / NF, j = threadIdx.x % NF;
        const int a = i < j ? i : j, b = i < j ? j : i;
        const int k = a * NF - a * (a - 1) / 2 + (b - a);   // index of (a,b), a<=b, row-major upper triangle
        S[i][j] = (i == j ? 2.0 : 1.0) * c[k];
    }
    if (threadIdx.x < 3) cx[threadIdx.x] = c[GPE_MOM_GRAD2 + threadIdx.x];
    __syncthreads();
    const double s = (double)A.scale;
    for (long long p = (long long)blockIdx.x * blockDim.x + threadIdx.x; p < A.n; p += (long long)gridDim.x * blockDim.x) {
        PointJets J;
        point_features(A, p, J);
        double fb[NF];
#pragma unroll
        for (int i = 1; i < NF; ++i) {
            double v = 0.0;
#pragma unroll
            for (int j = 0; j < NF; ++j) v += S[i][j] * J.f[j];
            fb[i] = v;
        }
        const double ub = fb[1] + J.f[3] * fb[4] + J.dnl * fb[5];   // d/du through u, V u, nl(u)
        const double ab = fb[2];                                    // d/d(lap u)
        const double gbx = 2.0 * cx[0] * J.gx, gby = 2.0 * cx[0] * J.gy;
        // chain to the network jets:  u = b + s F N,  u' = b' + s (F N' + F' N),  lap = .. + s (F N'' + 2 F' N' + F'' N)
        double nb = s * (J.F0 * ub + J.F1 * gbx + J.F2 * ab);
        const double n1bx = s * (J.F0 * gbx + 2.0 * J.F1 * ab), n1by = s * gby;
        const double n2b = s * J.F0 * ab;
        nb += 2.0 * cx[1] * J.n0 + 2.0 * cx[2] * J.sd;
        u_bar[p] = (float)nb;
        if (g_bar) {
            g_bar[p * A.d] = (float)n1bx;
            if (A.d == 2) g_bar[p * 2 + 1] = (float)n1by;
        }
        if (h_bar)
            for (int i = 0; i < A.h_cols; ++i) h_bar[p * A.h_cols + i] = i < A.lap_cols ? (float)n2b : 0.f;
        if (nm_bar) nm_bar[p] = (float)(-(double)A.sym_sign * 2.0 * cx[2] * J.sd);
    }
}

int loss_args(const gpe_loss_desc* desc, int64_t n, const float* x, const float* u, const float* g, const float* h,
              const float* base, const float* V, const float* q, const float* nm, LossArgs* A) {
    if (!desc) { set_error("null loss descriptor"); return -1; }
    if (desc->d != 1 && desc->d != 2) { set_error("d must be 1 or 2 (got %d)", desc->d); return -1; }
    if (desc->nl_kind < 0 || desc->nl_kind > 2) { set_error("unknown nonlinearity kind %d", desc->nl_kind); return -1; }
    if (desc->bfactor < 0 || desc->bfactor > 1 || (desc->bfactor == 1 && desc->d != 1)) {
        set_error("boundary factor %d unsupported for d=%d", desc->bfactor, desc->d);
        return -1;
    }
    if (desc->h_cols < 0 || desc->h_cols > 3 || desc->lap_cols < 0 || desc->lap_cols > desc->h_cols) {
        set_error("bad h_cols/lap_cols %d/%d", desc->h_cols, desc->lap_cols);
        return -1;
    }
    if (n < 0) { set_error("negative n"); return -1; }
    if (n > 0 && !u) { set_error("null u"); return -1; }
    if (desc->bfactor == 1 && !x) { set_error("x required for the boundary factor"); return -1; }
    if (desc->sym_sign != 0 && desc->sym_sign != 1 && desc->sym_sign != -1) { set_error("sym_sign must be -1, 0 or 1"); return -1; }
    if ((desc->sym_sign != 0) != (nm != nullptr)) { set_error("mirror values and sym_sign must be given together"); return -1; }
    if (desc->h_cols > 0 && !h && n > 0) { set_error("null h with h_cols=%d", desc->h_cols); return -1; }
    A->d = desc->d, A->nl_kind = desc->nl_kind, A->bfactor = desc->bfactor, A->sym_sign = desc->sym_sign;
    A->h_cols = desc->h_cols, A->lap_cols = desc->lap_cols, A->nl_power = desc->nl_power, A->scale = desc->scale;
    A->n = n, A->x = x, A->u = u, A->g = g, A->h = desc->h_cols ? h : nullptr, A->base = base, A->V = V, A->q = q, A->nm = nm;
    return 0;
}

int loss_grid(long long n) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const long long want = (n + kLossThreads - 1) / kLossThreads;
    const long long cap = (long long)sms * 4;     // a multiple of the SM count; grid-stride beyond
    return (int)(want < cap ? (want < 1 ? 1 : want) : cap);
}

// ---------------------------------------------------------------- fused scalar algebra of the loss terms
// Every loss term of the reference (residual mean squares with a Rayleigh / energy-quotient mu inside, Riesz energies, the
// [N,N]-broadcast factorisations, normalisation and symmetry terms) is a smooth function of the <= 3 x 32 moments.  The
// host describes that function once as a straight-line program (gpe_b200/expr.py: nodes = constants, inputs, + - * /, neg,
// sqrt); this kernel evaluates it in float64 together with its Jacobian by forward-mode differentiation -- thread j carries
// d/d input_j through the program -- so a loss method costs two launches (moments, this) instead of ~60 fp64 ATen
// kernels plus their autograd graph, and its backward is one [T] x [T, n_in] product.
constexpr int kProgMaxNodes = 1024;
enum { PROG_CONST = 0, PROG_INPUT = 1, PROG_ADD = 2, PROG_SUB = 3, PROG_MUL = 4, PROG_DIV = 5, PROG_NEG = 6, PROG_SQRT = 7 };

__global__ void __launch_bounds__(128) scalar_program_kernel(const int* __restrict__ ops, const double* __restrict__ cval, int n_nodes,
                                                             const double* __restrict__ in, int n_in, const int* __restrict__ outs,
                                                             int n_out, double* __restrict__ vals, double* __restrict__ jac) {
    // every thread evaluates the whole program (values are the same in all of them) and its own tangent column: no
    // synchronisation; the node arrays live in local memory (a few hundred nodes: L1-resident)
    double val[kProgMaxNodes], tan[kProgMaxNodes];
    for (int j = threadIdx.x; j < n_in || j == (int)threadIdx.x; j += blockDim.x) {
        for (int i = 0; i < n_nodes; ++i) {
            const int op = ops[3 * i], a = ops[3 * i + 1], b = ops[3 * i + 2];
            double v, t;
            switch (op) {
                case PROG_CONST: v = cval[i]; t = 0.0; break;
                case PROG_INPUT: v = in[a]; t = (a == j) ? 1.0 : 0.0; break;
                case PROG_ADD: v = val[a] + val[b]; t = tan[a] + tan[b]; break;
                case PROG_SUB: v = val[a] - val[b]; t = tan[a] - tan[b]; break;
                case PROG_MUL: v = val[a] * val[b]; t = tan[a] * val[b] + val[a] * tan[b]; break;
                case PROG_DIV: v = val[a] / val[b]; t = (tan[a] - v * tan[b]) / val[b]; break;
                case PROG_NEG: v = -val[a]; t = -tan[a]; break;
                default: v = sqrt(val[a]); t = tan[a] / (2.0 * v); break;
            }
            val[i] = v;
            tan[i] = t;
        }
        for (int k = 0; k < n_out; ++k) {
            if (j < n_in) jac[(size_t)k * n_in + j] = tan[outs[k]];
            if (j == 0) vals[k] = val[outs[k]];
        }
    }
}

}  // namespace
}  // namespace gpe

using namespace gpe;

extern "C" {

int gpe_gram_moments(const gpe_loss_desc* desc, int64_t n, const float* x, const float* u, const float* g, const float* h,
                     const float* base, const float* V, const float* q, const float* nm, double* m, gpe_stream_t stream) {
    gpe::NvtxRange range("gpe:gram_moments");
    LossArgs A;
    const int rc = loss_args(desc, n, x, u, g, h, base, V, q, nm, &A);
    if (rc) return rc;
    if (!m) { set_error("null moment buffer"); return -1; }
    if (n == 0) return 0;
    const int grid = loss_grid(n);
    if (grid < 0) { set_error("no CUDA device (this library has no CPU fallback)"); return -2; }
    gram_moments_kernel<<<grid, kLossThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(A, m);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("gram_moments_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}

int gpe_gram_adjoints(const gpe_loss_desc* desc, int64_t n, const float* x, const float* u, const float* g, const float* h,
                      const float* base, const float* V, const float* q, const float* nm, const double* c, float* u_bar,
                      float* g_bar, float* h_bar, float* nm_bar, gpe_stream_t stream) {
    gpe::NvtxRange range("gpe:gram_adjoints");
    LossArgs A;
    const int rc = loss_args(desc, n, x, u, g, h, base, V, q, nm, &A);
    if (rc) return rc;
    if (!c || (n > 0 && !u_bar)) { set_error("null coefficient / u_bar buffer"); return -1; }
    if (n > 0 && ((g && !g_bar) || (A.h && !h_bar) || (nm && !nm_bar))) { set_error("missing adjoint output buffer"); return -1; }
    if (n == 0) return 0;
    const int grid = loss_grid(n);
    if (grid < 0) { set_error("no CUDA device (this library has no CPU fallback)"); return -2; }
    gram_adjoints_kernel<<<grid, kLossThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(A, c, u_bar, g ? g_bar : nullptr,
                                                                                         A.h ? h_bar : nullptr, nm ? nm_bar : nullptr);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("gram_adjoints_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}

}  // extern "C"

extern "C" int gpe_scalar_program(const int32_t* ops, const double* consts, int n_nodes, const double* inputs, int n_in,
                                  const int32_t* outs, int n_out, double* vals, double* jac, gpe_stream_t stream) {
    gpe::NvtxRange range("gpe:scalar_program");
    if (!ops || !consts || !inputs || !outs || !vals || !jac) { gpe::set_error("gpe_scalar_program: null pointer argument"); return -1; }
    if (n_nodes < 1 || n_nodes > gpe::kProgMaxNodes) { gpe::set_error("gpe_scalar_program: %d nodes (1..%d)", n_nodes, gpe::kProgMaxNodes); return -1; }
    if (n_in < 1 || n_out < 1) { gpe::set_error("gpe_scalar_program: n_in and n_out must be positive"); return -1; }
    const int threads = n_in >= 128 ? 128 : (n_in + 31) / 32 * 32;
    gpe::scalar_program_kernel<<<1, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ops, consts, n_nodes, inputs, n_in, outs, n_out, vals, jac);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { gpe::set_error("scalar_program_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}


// ---- curriculum stage bookkeeping: one launch per epoch ------------------------------------------------------------------
// What R/harmonic_pinn_simulation.py:385-423 does on the HOST after every epoch -- scheduler.step(loss) of
// CosineAnnealingWarmRestarts (the reference passes the loss where the scheduler expects an epoch), the loss / mu /
// constraint logs, best-loss patience counter, the early-stop test and the snapshot of the parameters of the stop epoch --
// on device scalars, so an epoch stays one CUDA-graph replay.  One CTA: thread 0 does the scalar part, all threads copy the
// snapshot in the (single) epoch that trips the stop.
namespace gpe {
namespace {
__global__ void stage_bookkeeping_kernel(gpe_stage_state st, gpe_stage_consts c, gpe_stage_snapshot snap) {
    __shared__ int first_s;
    if (threadIdx.x == 0) {
        const float val = *st.total, best = *st.best;
        const long long ep = *st.epoch;
        // CosineAnnealingWarmRestarts(T_0, T_mult, eta_min).step(val): fp32, the order of gpe_b200.training._loss_as_epoch_lr
        const float T0 = c.t_0, tm1 = c.t_mult - 1.f;
        float n = floorf(log2f(fmaxf(val / T0 * tm1 + 1.f, 1.f)));
        const bool restart = val >= T0;
        const float p2 = exp2f(n);
        const float t_cur = restart ? val - T0 * (p2 - 1.f) / tm1 : val;
        const float t_i = restart ? T0 * p2 : T0;
        *st.lr = c.eta_min + (c.base_lr - c.eta_min) * (1.f + cosf(3.14159265358979323846f * t_cur / t_i)) / 2.f;
        if (ep >= 0 && ep < c.ring_width) {
            st.rings[ep] = val;
            st.rings[c.ring_width + ep] = *st.mu;
            st.rings[2 * c.ring_width + ep] = *st.constraint;
        }
        const bool improved = val < best;
        const long long stale = improved ? 0 : *st.stale + 1;
        *st.stale = stale;
        *st.best = (val != val || best != best) ? (val != val ? val : best) : fminf(best, val);   // torch.minimum: NaN propagates
        const bool hit = (val <= c.tol) || (stale >= c.patience);
        const bool first = hit && !*st.done;
        if (first) *st.final_epoch = ep;
        if (hit) *st.done = 1;
        *st.epoch = ep + 1;
        first_s = first ? 1 : 0;
    }
    __syncthreads();
    if (first_s) {
        for (int t = 0; t < snap.n; ++t)
            for (long long i = threadIdx.x; i < snap.numel[t]; i += blockDim.x) snap.dst[t][i] = snap.src[t][i];
    }
}
}  // namespace
}  // namespace gpe

extern "C" int gpe_stage_bookkeeping(const gpe_stage_state* st, const gpe_stage_consts* c, const gpe_stage_snapshot* snap, gpe_stream_t stream) {
    if (!st || !c || !snap) { gpe::set_error("gpe_stage_bookkeeping: null argument"); return -1; }
    if (!st->total || !st->mu || !st->constraint || !st->lr || !st->epoch || !st->final_epoch || !st->done || !st->best || !st->stale || !st->rings) {
        gpe::set_error("gpe_stage_bookkeeping: null state pointer"); return -1;
    }
    if (snap->n < 0 || snap->n > GPE_STAGE_MAX_TENSORS) { gpe::set_error("gpe_stage_bookkeeping: %d snapshot tensors (0..%d)", snap->n, GPE_STAGE_MAX_TENSORS); return -1; }
    for (int t = 0; t < snap->n; ++t)
        if (!snap->src[t] || !snap->dst[t] || snap->numel[t] < 0) { gpe::set_error("gpe_stage_bookkeeping: bad snapshot tensor %d", t); return -1; }
    if (c->ring_width < 0 || !(c->t_0 > 0.f) || !(c->t_mult > 1.f)) { gpe::set_error("gpe_stage_bookkeeping: bad constants"); return -1; }
    gpe::stage_bookkeeping_kernel<<<1, 1024, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*st, *c, *snap);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { gpe::set_error("stage_bookkeeping_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}
