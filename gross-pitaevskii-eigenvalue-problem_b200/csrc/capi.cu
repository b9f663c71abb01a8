// C ABI of libgpe_b200.so (see include/gpe_b200.h).  No torch types; device pointers in,
// device pointers out, everything enqueued on the caller's stream.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "../../include/gpe_b200.h"
#include "jet_host.h"
#include "tc_wimg.cuh"

namespace gpe {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

namespace {

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

// ---------------------------------------------------------------- weight packing
// wt_fwd[l][k=in][n=out] and w_bwd[l][k=out][n=in], zero-padded to WP; small = W0t | b | wout | bout
// One launch per call packs everything a call needs (a call used to cost up to three pack launches and a memset, which
// is what a 10k-point step is made of): the FFMA matrices, the small parameters, and -- tensor path -- the hi/lo weight
// images of the forward (wimg_f) and / or the reverse sweep (wimg_r: transposed), and it zeroes the accumulators.
__global__ void pack_weights_kernel(PtrArray W, PtrArray B, int D, int L, int Wd, int WP, float* __restrict__ wt_fwd,
                                    float* __restrict__ w_bwd, float* __restrict__ small, float* __restrict__ small_acc,
                                    float* __restrict__ wimg_f, float* __restrict__ wimg_r, double* __restrict__ acc64, long long n_acc64) {
    const long long n_mat = (long long)(L - 1) * WP * WP;
    const int n_small = (D + L + 1) * WP + 4;
    const long long total = n_mat + n_small;
    const long long stride = (long long)gridDim.x * blockDim.x, tid0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (wimg_f) tcp::pack_wimg_elems(W, L, Wd, 0, wimg_f, tid0, stride);
    if (wimg_r) tcp::pack_wimg_elems(W, L, Wd, 1, wimg_r, tid0, stride);
    for (long long i = tid0; i < n_acc64; i += stride) acc64[i] = 0.0;
    for (long long idx = tid0; idx < total; idx += stride) {
        if (idx < n_mat) {
            const int l = (int)(idx / ((long long)WP * WP));
            const int rem = (int)(idx % ((long long)WP * WP));
            const int k = rem / WP, nn = rem % WP;
            const bool in = k < Wd && nn < Wd;
            const float* w = W.p[l + 1];
            wt_fwd[idx] = in ? w[(size_t)nn * Wd + k] : 0.f;
            w_bwd[idx] = in ? w[(size_t)k * Wd + nn] : 0.f;
        } else {
            const int i = (int)(idx - n_mat);
            const int row = i / WP, c = i % WP;
            float v = 0.f;
            if (row < D) {
                if (c < Wd) v = W.p[0][(size_t)c * D + row];
            } else if (row < D + L) {
                if (c < Wd) v = B.p[row - D][c];
            } else if (row == D + L) {
                if (c < Wd) v = W.p[L][c];
            } else if (c == 0) {
                v = B.p[L][0];
            }
            small[i] = v;
            if (small_acc) small_acc[i] = 0.f;
        }
    }
}

// sum the per-CTA hidden-layer accumulators and scatter everything into PyTorch-shaped grads
// dWacc: per-CTA fp32 accumulators of the FFMA path (n_cta of them), or -- acc64 != nullptr -- the single fp64
// accumulator of the tensor path (the dW contraction adds every CTA's partial sums into it with fp64 atomics, so the
// sum over CTAs and launches carries no fp32 rounding)
// `abort_flag` (fused backward only): non-zero when the launch's flow-control watchdog fired -- the gradient is then
// poisoned with NaN so the failure cannot pass unnoticed.
__global__ void finish_grads_kernel(const float* __restrict__ dWacc, const double* __restrict__ acc64, int n_cta,
                                    const float* __restrict__ small_acc, int D, int L, int Wd, int WP, MutPtrArray dW, MutPtrArray dB,
                                    const unsigned int* __restrict__ abort_flag) {
    const float poison = (abort_flag && *abort_flag) ? __int_as_float(0x7fc00000) : 0.f;
    const long long n_hidden_el = (long long)(L - 1) * Wd * Wd;
    const long long n_rest = (long long)Wd * D + (long long)L * Wd + Wd + 1;
    const size_t cta_stride = (size_t)(L - 1) * WP * WP;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n_hidden_el + n_rest;
         idx += (long long)gridDim.x * blockDim.x) {
        if (idx < n_hidden_el) {
            const int l = (int)(idx / ((long long)Wd * Wd));
            const int rem = (int)(idx % ((long long)Wd * Wd));
            const int o = rem / Wd, i = rem % Wd;
            const size_t at = (size_t)l * WP * WP + (size_t)o * WP + i;
            double s = 0.0;
            if (acc64) {
                if (n_cta > 0) s = acc64[at];
            } else {
                // independent loads first (8 in flight per thread), then the ordered fp64 sum: the per-CTA partials of a small
                // launch are read at L2 latency, one after the other otherwise (18 us for 63 CTAs)
                int c = 0;
                for (; c + 8 <= n_cta; c += 8) {
                    float v[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) v[j] = __ldg(dWacc + at + (size_t)(c + j) * cta_stride);
#pragma unroll
                    for (int j = 0; j < 8; ++j) s += (double)v[j];
                }
                for (; c < n_cta; ++c) s += (double)dWacc[at + (size_t)c * cta_stride];
            }
            dW.p[l + 1][(size_t)o * Wd + i] = (float)s + poison;
        } else {
            long long j = idx - n_hidden_el;
            if (j < (long long)Wd * D) {
                const int c = (int)(j / D), i = (int)(j % D);
                dW.p[0][j] = small_acc[i * WP + c] + poison;
            } else if ((j -= (long long)Wd * D) < (long long)L * Wd) {
                const int l = (int)(j / Wd), c = (int)(j % Wd);
                dB.p[l][c] = small_acc[(D + l) * WP + c];
            } else if ((j -= (long long)L * Wd) < Wd) {
                dW.p[L][j] = small_acc[(D + L) * WP + j];
            } else {
                dB.p[L][0] = small_acc[(D + L + 1) * WP];
            }
        }
    }
}

// ---------------------------------------------------------------- plan cache
struct PlanKey {
    int d, jet, wp, L, dev, res;
};
struct PlanEntry {
    PlanKey k;
    JetPlan p;
};
std::mutex g_plan_mu;
PlanEntry g_plans[64];
int g_n_plans = 0;

bool use_tensor_path(const gpe_mlp_desc* desc);
bool tc_shape_ok(const gpe_mlp_desc* desc);
int padded_width(int w) { return w <= 64 ? 64 : w <= 128 ? 128 : w <= 256 ? 256 : 0; }

int get_plan(const gpe_mlp_desc* desc, JetPlan* out, int* wp_out) {
    if (!desc) { set_error("null descriptor"); return -1; }
    if (desc->d != 1 && desc->d != 2) { set_error("d must be 1 or 2 (got %d)", desc->d); return -1; }
    if (desc->n_hidden < 1 || desc->n_hidden > kMaxLayers) { set_error("n_hidden must be in [1,%d]", kMaxLayers); return -1; }
    if (desc->act < 0 || desc->act > 3) { set_error("unknown activation %d", desc->act); return -1; }
    const int res = desc->act == ACT_RESBLOCK_TANH;
    if (res && !(desc->n_hidden & 1)) {
        set_error("residual-block network: n_hidden must be 1 + 2 * blocks (got %d)", desc->n_hidden);
        return -1;
    }
    if (desc->jet < 0 || desc->jet > 3) { set_error("unknown jet kind %d", desc->jet); return -1; }
    if (desc->precision < GPE_PREC_FP32 || desc->precision > GPE_PREC_AUTO) { set_error("unknown precision %d", desc->precision); return -1; }
    if (desc->precision == GPE_PREC_TF32X3 && !tc_shape_ok(desc)) {
        set_error("GPE_PREC_TF32X3: no tcgen05 kernel for d=%d width=%d n_hidden=%d jet=%d act=%d (33 <= width <= 128 -- LeakyReLU: 65 <= width "
                  "--, 2..4 hidden layers, Laplacian-type jets); use GPE_PREC_AUTO or GPE_PREC_FP32", desc->d, desc->width, desc->n_hidden, desc->jet, desc->act);
        return -1;
    }
    // the tensor path packs every width up to 128 into its 128-wide layout
    const int wp = (desc->width >= 1 && desc->width <= 64 && use_tensor_path(desc)) ? 128 : padded_width(desc->width);
    if (desc->width < 1 || !wp) { set_error("hidden width %d unsupported (1..256)", desc->width); return -1; }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device (this library has no CPU fallback)"); return -2; }
    *wp_out = wp;
    std::lock_guard<std::mutex> lk(g_plan_mu);
    for (int i = 0; i < g_n_plans; ++i) {
        const PlanKey& k = g_plans[i].k;
        if (k.d == desc->d && k.jet == desc->jet && k.wp == wp && k.L == desc->n_hidden && k.dev == dev && k.res == res) {
            *out = g_plans[i].p;
            return 0;
        }
    }
    JetPlan p;
    const int rc = wp == 64    ? plan_wp64(desc->d, desc->jet, desc->n_hidden, res, &p)
                   : wp == 128 ? plan_wp128(desc->d, desc->jet, desc->n_hidden, res, &p)
                               : plan_wp256(desc->d, desc->jet, desc->n_hidden, res, &p);
    if (rc) return rc;
    if (g_n_plans < 64) g_plans[g_n_plans++] = PlanEntry{PlanKey{desc->d, desc->jet, wp, desc->n_hidden, dev, res}, p};
    *out = p;
    return 0;
}

// GPE_TENSOR=0 forces the FFMA kernels (the tcgen05 path is the default where it applies)
int g_tc_enabled = -1;
bool tc_enabled() {
    if (g_tc_enabled < 0) {
        const char* e = getenv("GPE_TENSOR");
        g_tc_enabled = (e && e[0] == '0') ? 0 : 1;
    }
    return g_tc_enabled != 0;
}

// Which kernels serve a descriptor: gpe_mlp_desc.precision decides PER CALL -- GPE_PREC_FP32: the FFMA kernels;
// GPE_PREC_TF32X3: the tcgen05 kernels (get_plan rejects shapes they do not cover); GPE_PREC_AUTO: tcgen05 where a kernel
// exists, unless the process default (GPE_TENSOR=0 / gpe_set_tensor_path(0)) says FFMA.
bool tc_shape_ok(const gpe_mlp_desc* desc) {
    // W <= 64 (zero-padded into the 128-wide kernels) only with the smooth activations: a piecewise-linear activation turns the
    // tensor path's ~1e-6 perturbation of a pre-activation next to the kink into an O(1/N) change of the gradient
    const bool smooth = desc->act == ACT_TANH || desc->act == ACT_SHIFTED_TANH;
    return desc->act != ACT_RESBLOCK_TANH && (desc->width > 64 || smooth) && tc_eligible(desc->d, desc->jet, desc->width, desc->n_hidden);
}
bool use_tensor_path(const gpe_mlp_desc* desc) {
    if (desc->precision == GPE_PREC_FP32) return false;
    if (desc->precision == GPE_PREC_TF32X3) return tc_shape_ok(desc);
    return tc_enabled() && tc_shape_ok(desc);
}

// Value-only forwards (the mirrored points of symmetry_loss, monitoring calls) of a net the tensor kernels cover: from
// kValueTcMinPoints points on, the tensor forward -- jets computed and dropped -- beats the FFMA value kernel (1M points,
// [1,128x4,1]: 1.5 vs 2.4 ms).  GPE_PREC_AUTO only; the backward of a value-only call stays on the FFMA kernels.
constexpr long long kValueTcMinPoints = 16384;
bool value_on_tensor(const gpe_mlp_desc* desc, long long n, gpe_mlp_desc* as_jets) {
    if (desc->jet != GPE_JET_VALUE || desc->precision != GPE_PREC_AUTO || n < kValueTcMinPoints) return false;
    gpe_mlp_desc dv = *desc;
    dv.jet = desc->d == 1 ? GPE_JET_DIAG : GPE_JET_LAPLACIAN;
    if (!use_tensor_path(&dv)) return false;
    if (as_jets) *as_jets = dv;
    return true;
}

struct WsLayout {
    size_t wt_fwd, w_bwd, small, wimg, small_acc, dWacc, stash, wimg_rev, zb, total_fwd, total_bwd;
    bool tc;
    int n_small, grid_fwd, grid_bwd, tiles;
};

WsLayout layout(const gpe_mlp_desc* desc, const JetPlan& p, int wp, long long n) {
    WsLayout w;
    const int L = desc->n_hidden;
    const size_t mat = (size_t)(L - 1) * wp * wp * sizeof(float);
    w.n_small = (desc->d + L + 1) * wp + 4;
    long long tiles = (n + p.pts_per_tile - 1) / p.pts_per_tile;
    w.tiles = (int)tiles;
    w.grid_fwd = (int)(tiles < p.max_ctas_fwd ? tiles : p.max_ctas_fwd);
    w.grid_bwd = (int)(tiles < p.max_ctas_bwd ? tiles : p.max_ctas_bwd);
    if (w.grid_fwd < 1) w.grid_fwd = 1;
    if (w.grid_bwd < 1) w.grid_bwd = 1;
    size_t off = 0;
    w.wt_fwd = off; off = align_up(off + mat);
    w.w_bwd = off; off = align_up(off + mat);
    w.small = off; off = align_up(off + w.n_small * sizeof(float));
    w.tc = use_tensor_path(desc);
    w.wimg = off;
    if (w.tc) off = align_up(off + tc_wimg_bytes(L));
    w.total_fwd = off;
    w.small_acc = off; off = align_up(off + w.n_small * sizeof(float));
    w.wimg_rev = w.zb = 0;
    if (w.tc) {   // tensor path: one global accumulator, chunked stash of forward jets and of their adjoints
        w.dWacc = off; off = align_up(off + 2 * mat);   // fp64
        w.wimg_rev = off; off = align_up(off + tc_wimg_bytes(L));
        w.stash = off; off = align_up(off + tc_stash_bytes(L, n));
        w.zb = off; off = align_up(off + tc_zb_bytes(L, n));
    } else {
        w.dWacc = off; off = align_up(off + (size_t)w.grid_bwd * mat);
        w.stash = off; off = align_up(off + (size_t)w.grid_bwd * p.stash_floats_per_cta * sizeof(float));
    }
    w.total_bwd = off;
    return w;
}

int check_launch(const char* what) {
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); return -4; }
    return 0;
}

int pack(const gpe_mlp_desc* desc, int wp, const WsLayout& w, const float* const* W, const float* const* b, char* ws,
         bool with_acc, cudaStream_t s, bool img_fwd = false, bool img_rev = false, bool zero_acc64 = false) {
    PtrArray Wp{}, Bp{};
    for (int l = 0; l <= desc->n_hidden; ++l) {
        if (!W[l] || !b[l]) { set_error("null weight/bias pointer at layer %d", l); return -1; }
        Wp.p[l] = W[l];
        Bp.p[l] = b[l];
    }
    const long long total = (long long)(desc->n_hidden - 1) * wp * wp + w.n_small;
    const int blocks = (int)((total + 255) / 256 < 592 ? (total + 255) / 256 : 592);
    const long long n_acc64 = zero_acc64 ? (long long)(desc->n_hidden - 1) * wp * wp : 0;
    pack_weights_kernel<<<blocks, 256, 0, s>>>(Wp, Bp, desc->d, desc->n_hidden, desc->width, wp,
                                               reinterpret_cast<float*>(ws + w.wt_fwd), reinterpret_cast<float*>(ws + w.w_bwd),
                                               reinterpret_cast<float*>(ws + w.small),
                                               with_acc ? reinterpret_cast<float*>(ws + w.small_acc) : nullptr,
                                               img_fwd ? reinterpret_cast<float*>(ws + w.wimg) : nullptr,
                                               img_rev ? reinterpret_cast<float*>(ws + w.wimg_rev) : nullptr,
                                               zero_acc64 ? reinterpret_cast<double*>(ws + w.dWacc) : nullptr, n_acc64);
    return check_launch("pack_weights_kernel");
}

XForm make_xform(const gpe_xform* xf) {
    XForm x;
    for (int i = 0; i < 2; ++i) {
        x.a[i] = xf ? xf->a[i] : 1.f;
        x.b[i] = xf ? xf->b[i] : 0.f;
    }
    return x;
}

// ---------------------------------------------------------------- FFMA probe
__global__ void __launch_bounds__(256) ffma_probe_kernel(int mode, int iters, float a, float b, float* sink) {
    if (mode == 0) {
        float acc[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = (float)(threadIdx.x + i);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) acc[i] = fmaf(acc[i], a, b);
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; ++i) s += acc[i];
        if (s == 123.456f) sink[0] = s;
    } else {
        unsigned long long acc[8], av, bv;
        asm("mov.b64 %0, {%1, %1};" : "=l"(av) : "f"(a));
        asm("mov.b64 %0, {%1, %1};" : "=l"(bv) : "f"(b));
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float lo = (float)(threadIdx.x + i), hi = lo + 0.5f;
            asm("mov.b64 %0, {%1, %2};" : "=l"(acc[i]) : "f"(lo), "f"(hi));
        }
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[i]) : "l"(av), "l"(bv));
        }
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float lo, hi;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[i]));
            s += lo + hi;
        }
        if (s == 123.456f) sink[0] = s;
    }
}

}  // namespace
}  // namespace gpe

using namespace gpe;

extern "C" {

const char* gpe_version(void) { return "gpe_b200 0.1 (sm_100a)"; }
const char* gpe_last_error(void) { return g_err; }

int gpe_workspace_bytes(const gpe_mlp_desc* desc, int64_t n, size_t* fwd_bytes, size_t* bwd_bytes) {
    JetPlan p;
    int wp;
    const int rc = get_plan(desc, &p, &wp);
    if (rc) return rc;
    if (n < 0) { set_error("negative n"); return -1; }
    const WsLayout w = layout(desc, p, wp, n);
    size_t fwd = w.total_fwd;
    gpe_mlp_desc dv;
    if (value_on_tensor(desc, n, &dv)) {       // the forward may run on the tensor kernels: their layout (128-wide, weight images)
        JetPlan pv;
        int wpv;
        if (int rcv = get_plan(&dv, &pv, &wpv)) return rcv;
        const size_t fv = layout(&dv, pv, wpv, n).total_fwd;
        fwd = fv > fwd ? fv : fwd;
    }
    if (fwd_bytes) *fwd_bytes = fwd;
    if (bwd_bytes) *bwd_bytes = w.total_bwd;
    return 0;
}

static int jet_fwd_impl(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                        const gpe_xform* xform, float* u, float* g, float* h, float* resident, size_t resident_bytes, void* ws,
                        size_t ws_bytes, gpe_stream_t stream) {
    JetPlan p;
    int wp;
    int rc = get_plan(desc, &p, &wp);
    if (rc) return rc;
    if (n < 0) { set_error("negative n"); return -1; }
    if (n == 0) return 0;
    if (!W || !b || !x || !u || !ws) { set_error("null pointer argument"); return -1; }
    if (desc->jet != GPE_JET_VALUE && (!g || !h)) { set_error("g and h outputs required for derivative jets"); return -1; }
    if (xform && desc->jet != GPE_JET_VALUE) { set_error("xform is only valid with GPE_JET_VALUE"); return -1; }
    gpe_mlp_desc dv;
    if (!resident && value_on_tensor(desc, n, &dv)) {
        JetPlan pv;
        int wpv;
        if ((rc = get_plan(&dv, &pv, &wpv))) return rc;
        const WsLayout wv = layout(&dv, pv, wpv, n);
        if (ws_bytes < wv.total_fwd) { set_error("workspace too small: %zu < %zu", ws_bytes, wv.total_fwd); return -1; }
        char* vbase = static_cast<char*>(ws);
        cudaStream_t vs = reinterpret_cast<cudaStream_t>(stream);
        NvtxRange vrange("gpe:jet_fwd(tensor, value only)");
        if ((rc = pack(&dv, wpv, wv, W, b, vbase, false, vs, true, false, false))) return rc;
        const XForm vxf = make_xform(xform);
        return launch_fwd_tc(dv.d, dv.n_hidden, dv.act, reinterpret_cast<float*>(vbase + wv.wimg), reinterpret_cast<float*>(vbase + wv.small), x, n, u,
                             nullptr, nullptr, nullptr, vs, vxf.a, vxf.b);
    }
    const WsLayout w = layout(desc, p, wp, n);
    if (ws_bytes < w.total_fwd) { set_error("workspace too small: %zu < %zu", ws_bytes, w.total_fwd); return -1; }
    char* base = static_cast<char*>(ws);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    NvtxRange range(w.tc ? "gpe:jet_fwd(tensor)" : "gpe:jet_fwd(ffma)");
    if ((rc = pack(desc, wp, w, W, b, base, false, s, w.tc, false, false))) return rc;
    MlpParams P{desc->n_hidden, desc->width, desc->act, reinterpret_cast<float*>(base + w.wt_fwd),
                reinterpret_cast<float*>(base + w.w_bwd), reinterpret_cast<float*>(base + w.small)};
    const XForm xf = make_xform(xform);
    if (w.tc) {
        float* wimg = reinterpret_cast<float*>(base + w.wimg);
        if (resident && resident_bytes < tc_resident_stash_bytes(desc->n_hidden, n)) {
            set_error("stash too small: %zu < %zu", resident_bytes, tc_resident_stash_bytes(desc->n_hidden, n));
            return -1;
        }
        return launch_fwd_tc(desc->d, desc->n_hidden, desc->act, wimg, P.small, x, n, u, g, h, resident, s);
    }
    if (resident) { set_error("this descriptor has no stash (gpe_stash_bytes = 0): use gpe_mlp_jet_fwd"); return -1; }
    rc = wp == 64    ? launch_fwd_wp64(desc->d, desc->jet, P, x, n, xf, u, g, h, w.grid_fwd, w.tiles, s)
         : wp == 128 ? launch_fwd_wp128(desc->d, desc->jet, P, x, n, xf, u, g, h, w.grid_fwd, w.tiles, s)
                     : launch_fwd_wp256(desc->d, desc->jet, P, x, n, xf, u, g, h, w.grid_fwd, w.tiles, s);
    if (rc) return rc;
    return check_launch("jet_fwd_kernel");
}

int gpe_mlp_jet_fwd(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                    const gpe_xform* xform, float* u, float* g, float* h, void* ws, size_t ws_bytes,
                    gpe_stream_t stream) {
    return jet_fwd_impl(desc, W, b, x, n, xform, u, g, h, nullptr, 0, ws, ws_bytes, stream);
}

// ---- forward without any allocation: the workspace is a static device array --------------------------------------------
// (no counterpart in the reference; added for callers that cannot call cudaMalloc.)
// The array lives in every CUDA context that loads this library until the process exits.  In HOST code `g_static_ws` is only
// a handle for the symbol, never a device address: the address is asked for once per device (with that device current)
// through cudaGetSymbolAddress and that pointer is what the kernels get.
static __device__ __align__(256) unsigned char g_static_ws[GPE_STATIC_WS_BYTES];

static int static_ws_ptr(void** out) {
    constexpr int kMaxDev = 64;
    static std::mutex mu;
    static void* ptrs[kMaxDev] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { set_error("cudaGetDevice: %s", cudaGetErrorString(e)); return -3; }
    if (dev < 0 || dev >= kMaxDev) { set_error("device index %d outside the static-workspace table (%d)", dev, kMaxDev); return -1; }
    std::lock_guard<std::mutex> lk(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_static_ws);
        if (e != cudaSuccess || !p) { set_error("cudaGetSymbolAddress(g_static_ws): %s", cudaGetErrorString(e)); return -3; }
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return 0;
}

int gpe_mlp_jet_fwd_noalloc(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                            const gpe_xform* xform, float* u, float* g, float* h, gpe_stream_t stream) {
    size_t need = 0;
    int rc = gpe_workspace_bytes(desc, n, &need, nullptr);
    if (rc) return rc;
    if (need > (size_t)GPE_STATIC_WS_BYTES) {
        set_error("gpe_mlp_jet_fwd_noalloc: this network needs %zu bytes of workspace, the static array holds %zu (GPE_STATIC_WS_BYTES); "
                  "use gpe_mlp_jet_fwd with a caller workspace", need, (size_t)GPE_STATIC_WS_BYTES);
        return -1;
    }
    if (n > (int64_t)GPE_STATIC_MAX_POINTS) {
        set_error("gpe_mlp_jet_fwd_noalloc: n = %lld exceeds GPE_STATIC_MAX_POINTS = %lld", (long long)n, (long long)GPE_STATIC_MAX_POINTS);
        return -1;
    }
    void* ws = nullptr;
    if ((rc = static_ws_ptr(&ws))) return rc;
    // every byte of the workspace the forward reads (packed weights, zero padding, weight images) is rewritten by the pack
    // kernel of this call, on `stream`, before the sweep reads it: nothing of the previous call's contents survives into it
    return jet_fwd_impl(desc, W, b, x, n, xform, u, g, h, nullptr, 0, ws, (size_t)GPE_STATIC_WS_BYTES, stream);
}

int gpe_stash_bytes(const gpe_mlp_desc* desc, int64_t n, size_t* bytes) {
    JetPlan p;
    int wp;
    const int rc = get_plan(desc, &p, &wp);
    if (rc) return rc;
    if (n < 0 || !bytes) { set_error("gpe_stash_bytes: bad arguments"); return -1; }
    *bytes = (use_tensor_path(desc) && n > 0) ? tc_resident_stash_bytes(desc->n_hidden, n) : 0;
    return 0;
}

int gpe_mlp_jet_fwd_stash(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                          float* u, float* g, float* h, void* stash, size_t stash_bytes, void* ws, size_t ws_bytes,
                          gpe_stream_t stream) {
    if (!stash) { set_error("null stash"); return -1; }
    return jet_fwd_impl(desc, W, b, x, n, nullptr, u, g, h, static_cast<float*>(stash), stash_bytes, ws, ws_bytes, stream);
}

static int jet_bwd_impl(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                        const gpe_xform* xform, const float* u_bar, const float* g_bar, const float* h_bar, const float* resident,
                        size_t resident_bytes, float* const* dW, float* const* db, void* ws, size_t ws_bytes, gpe_stream_t stream) {
    JetPlan p;
    int wp;
    int rc = get_plan(desc, &p, &wp);
    if (rc) return rc;
    if (n < 0) { set_error("negative n"); return -1; }
    if (!W || !b || !dW || !db || !ws || (n > 0 && !x)) { set_error("null pointer argument"); return -1; }
    if (xform && desc->jet != GPE_JET_VALUE) { set_error("xform is only valid with GPE_JET_VALUE"); return -1; }
    const WsLayout w = layout(desc, p, wp, n);
    if (ws_bytes < w.total_bwd) { set_error("workspace too small: %zu < %zu", ws_bytes, w.total_bwd); return -1; }
    char* base = static_cast<char*>(ws);
    cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
    NvtxRange range(w.tc ? "gpe:jet_bwd(tensor)" : "gpe:jet_bwd(ffma)");
    if ((rc = pack(desc, wp, w, W, b, base, true, s, w.tc && n > 0, w.tc && n > 0, w.tc))) return rc;
    const int L = desc->n_hidden;
    const size_t mat = (size_t)(L - 1) * wp * wp * sizeof(float);
    if (mat && !w.tc) {   // FFMA path: per-CTA fp32 accumulators (the tensor path's fp64 accumulator is zeroed by the pack launch)
        const cudaError_t e = cudaMemsetAsync(base + w.dWacc, 0, (size_t)w.grid_bwd * mat, s);
        if (e != cudaSuccess) { set_error("memset: %s", cudaGetErrorString(e)); return -4; }
    }
    MlpParams P{L, desc->width, desc->act, reinterpret_cast<float*>(base + w.wt_fwd), reinterpret_cast<float*>(base + w.w_bwd),
                reinterpret_cast<float*>(base + w.small)};
    const XForm xf = make_xform(xform);
    const unsigned int* abort_flag = nullptr;
    if (n > 0 && w.tc) {
        float* wimg_f = reinterpret_cast<float*>(base + w.wimg);
        float* wimg_r = reinterpret_cast<float*>(base + w.wimg_rev);
        if (resident && resident_bytes < tc_resident_stash_bytes(L, n)) {
            set_error("stash too small: %zu < %zu", resident_bytes, tc_resident_stash_bytes(L, n));
            return -1;
        }
        rc = launch_bwd_tc(desc->d, L, desc->act, wimg_f, wimg_r, P.small, x, n, u_bar, g_bar, h_bar,
                           reinterpret_cast<float*>(base + w.stash), resident, reinterpret_cast<float*>(base + w.zb),
                           reinterpret_cast<double*>(base + w.dWacc), reinterpret_cast<float*>(base + w.small_acc), s, &abort_flag);
        if (rc) return rc;
    } else if (n > 0) {
        if (resident) { set_error("this descriptor has no stash (gpe_stash_bytes = 0): use gpe_mlp_jet_bwd"); return -1; }
        float* st = reinterpret_cast<float*>(base + w.stash);
        float* da = reinterpret_cast<float*>(base + w.dWacc);
        float* sa = reinterpret_cast<float*>(base + w.small_acc);
        rc = wp == 64    ? launch_bwd_wp64(desc->d, desc->jet, P, x, n, xf, u_bar, g_bar, h_bar, st, da, sa, w.grid_bwd, w.tiles, s)
             : wp == 128 ? launch_bwd_wp128(desc->d, desc->jet, P, x, n, xf, u_bar, g_bar, h_bar, st, da, sa, w.grid_bwd, w.tiles, s)
                         : launch_bwd_wp256(desc->d, desc->jet, P, x, n, xf, u_bar, g_bar, h_bar, st, da, sa, w.grid_bwd, w.tiles, s);
        if (rc) return rc;
        if ((rc = check_launch("jet_bwd_kernel"))) return rc;
    }
    MutPtrArray dWp{}, dBp{};
    for (int l = 0; l <= L; ++l) {
        if (!dW[l] || !db[l]) { set_error("null gradient pointer at layer %d", l); return -1; }
        dWp.p[l] = dW[l];
        dBp.p[l] = db[l];
    }
    const long long total = (long long)(L - 1) * desc->width * desc->width + (long long)desc->width * (desc->d + L + 1) + 1;
    const int blocks = (int)((total + 255) / 256);
    finish_grads_kernel<<<blocks, 256, 0, s>>>(reinterpret_cast<float*>(base + w.dWacc),
                                               w.tc ? reinterpret_cast<double*>(base + w.dWacc) : nullptr,
                                               n > 0 ? (w.tc ? 1 : w.grid_bwd) : 0, reinterpret_cast<float*>(base + w.small_acc), desc->d, L, desc->width, wp, dWp, dBp,
                                               abort_flag);
    return check_launch("finish_grads_kernel");
}

int gpe_mlp_jet_bwd(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                    const gpe_xform* xform, const float* u_bar, const float* g_bar, const float* h_bar,
                    float* const* dW, float* const* db, void* ws, size_t ws_bytes, gpe_stream_t stream) {
    return jet_bwd_impl(desc, W, b, x, n, xform, u_bar, g_bar, h_bar, nullptr, 0, dW, db, ws, ws_bytes, stream);
}

int gpe_mlp_jet_bwd_stashed(const gpe_mlp_desc* desc, const float* const* W, const float* const* b, const float* x, int64_t n,
                            const float* u_bar, const float* g_bar, const float* h_bar, const void* stash, size_t stash_bytes,
                            float* const* dW, float* const* db, void* ws, size_t ws_bytes, gpe_stream_t stream) {
    if (!stash) { set_error("null stash"); return -1; }
    return jet_bwd_impl(desc, W, b, x, n, nullptr, u_bar, g_bar, h_bar, static_cast<const float*>(stash), stash_bytes, dW, db, ws,
                        ws_bytes, stream);
}

int gpe_set_tensor_path(int enable) {
    const int prev = tc_enabled() ? 1 : 0;
    g_tc_enabled = enable ? 1 : 0;
    return prev;
}

int gpe_tc_profile(int enable) {
    tc_profile_enable(enable != 0);
    return 0;
}

int gpe_tc_timings(double* ms, int n) {
    if (!ms || n < 1) { set_error("gpe_tc_timings: bad arguments"); return -1; }
    return tc_profile_read(ms, n);
}

int gpe_ffma_probe(int mode, int iters, float* sink, double* flops, gpe_stream_t stream) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) { set_error("no CUDA device"); return -2; }
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = sms * 8;
    ffma_probe_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(mode, iters, 1.0000001f, 1e-7f, sink);
    if (flops) *flops = (double)grid * 256.0 * (double)iters * 16.0 * 2.0;
    return check_launch("ffma_probe_kernel");
}

}  // extern "C"
