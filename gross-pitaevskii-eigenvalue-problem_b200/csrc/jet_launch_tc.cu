// Launch unit of the tensor-core (tcgen05) path.  Hidden width padded to 128, up to 3 resident hidden matrices.
#include "jet_host.h"
#include <cstdlib>
#include <mutex>
#include <utility>
#include <vector>
#include "jet_tc_kernels.cuh"

namespace gpe {

bool tc_eligible(int d, int jet, int width, int L) {
    const bool lap = jet == JET_LAPLACIAN || (d == 1 && (jet == JET_DIAG || jet == JET_FULL));
    // every width up to 128 is zero-padded to the 128-wide images: measured on B200 the padded tensor kernels beat the FFMA
    // kernels of W <= 64 at every size (scripts/w64_probe.py, [2,64x3,1]: 4000 points 0.11 vs 0.12 ms, 1M points 5.3 vs 12.0 ms)
    // (not below 33: a handful of features with a small, cancelling output sees the 2e-7 absolute error of the ex2/rcp tanh;
    // capi.cu additionally keeps kinked activations of W <= 64 on the FFMA kernels)
    return lap && width >= 33 && width <= 128 && L >= 2 && L - 1 <= tcp::kTcMaxMat;
}

size_t tc_wimg_bytes(int L) { return (size_t)(L - 1) * 2 * 2 * tcp::kImgFloats * sizeof(float); }

int tc_pack_wimg(const PtrArray& W, int L, int Wd, int transpose, float* wimg, cudaStream_t s) {
    tcp::pack_wimg_kernel<<<(L - 1) * 64, 256, 0, s>>>(W, L, Wd, transpose, wimg);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("pack_wimg_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}

static int tc_interleave();

// Per-device facts, looked up once per device (several devices may be driven from one process).
constexpr int kMaxDev = 64;
static int cur_dev() {
    int dev = 0;
    cudaGetDevice(&dev);
    return dev < 0 || dev >= kMaxDev ? 0 : dev;
}
static int tc_sms() {
    static int sms[kMaxDev] = {};
    const int dev = cur_dev();
    if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    return sms[dev];
}
// cudaFuncAttributeMaxDynamicSharedMemorySize is set once per (kernel, device): the largest footprint any eligible network
// needs (kTcMaxMat resident matrices), so the launches themselves carry no attribute call.
static std::mutex g_attr_mu;
static std::vector<std::pair<const void*, int>> g_attr_done;   // (kernel, device) pairs already configured
template <typename K>
static int ensure_smem(K kernel, size_t smem, const char* what) {
    const std::pair<const void*, int> key(reinterpret_cast<const void*>(kernel), cur_dev());
    std::lock_guard<std::mutex> lk(g_attr_mu);
    for (const auto& k : g_attr_done)
        if (k == key) return 0;
    const cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_error("%s: shared-memory attribute (%zu B): %s", what, smem, cudaGetErrorString(e)); return -3; }
    g_attr_done.push_back(key);
    return 0;
}

// Two tiles in flight per CTA pair (A/B knobs, read per call; GPE_TC_TWO_TILES=0 / GPE_TC_TWO_TILES_REV=0 select the one-tile
// kernels).  Measured on B200 at 4M points, 1000 W cap (scripts/power_probe.py): forward 7.5 -> 5.9 ms, stash forward
// 10.6 -> 8.1 ms, reverse sweep 13.6 -> 12.8 ms (with the bf16 correction products; with three tf32 products the reverse
// sweep was epilogue-bound and gained nothing).
static bool tc_two_tiles() {
    const char* ev = getenv("GPE_TC_TWO_TILES");
    return !(ev && ev[0] == '0');
}
static bool tc_two_tiles_rev() {
    const char* ev = getenv("GPE_TC_TWO_TILES_REV");
    return !(ev && ev[0] == '0');
}

template <int D, bool STASH, int ACT>
static int fwd_tc_act(const tcp::TcParams& P, const float* x, long long n, float* u, float* g, float* h, float* stash, cudaStream_t s) {
    const int sms = tc_sms();
    const size_t smem = tcp::tc_smem_bytes(D, P.n_hidden);
    const bool two = tc_two_tiles();
    auto kern = two ? tcp::jet_fwd2_tc_kernel<D, STASH, ACT> : tcp::jet_fwd_tc_kernel<D, STASH, ACT>;
    if (int rc = ensure_smem(kern, tcp::tc_smem_bytes(D, tcp::kTcMaxMat + 1), "jet_fwd_tc_kernel")) return rc;
    cudaError_t e;
    const long long pairs = (n + 2 * tcp::kTcPts - 1) / (2 * tcp::kTcPts);
    long long clusters = sms / 2;
    if (pairs < clusters) clusters = pairs;
    kern<<<(unsigned)(2 * clusters), tcp::kTcThreads, smem, s>>>(P, x, n, u, g, h, stash, (long long)((n + 7) / 8) * tcp::kBlkFloats, STASH ? (int)((n + 7) / 8) : 0, (int)pairs);
    e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("jet_fwd_tc_kernel launch: %s (grid %lld, smem %zu)", cudaGetErrorString(e), 2 * clusters, smem); return -4; }
    return 0;
}

// tanh (every BASELINE config) gets kernels with the activation fixed at compile time (GPE_TC_ACT_GENERIC=1: A/B knob)
static bool tc_act_generic() {
    static int v = -1;
    if (v < 0) { const char* ev = getenv("GPE_TC_ACT_GENERIC"); v = (ev && ev[0] == '1') ? 1 : 0; }
    return v == 1;
}
template <int D, bool STASH>
static int fwd_tc(const tcp::TcParams& P, const float* x, long long n, float* u, float* g, float* h, float* stash, cudaStream_t s) {
    return (P.act == ACT_TANH && !tc_act_generic()) ? fwd_tc_act<D, STASH, ACT_TANH>(P, x, n, u, g, h, stash, s) : fwd_tc_act<D, STASH, -1>(P, x, n, u, g, h, stash, s);
}

// optional per-kernel timing of the tensor-path backward (bench.py roofline): events on the launch stream
static bool g_tc_profile = false;
struct TcEv { cudaEvent_t a, b; int kind; };
static std::vector<TcEv> g_tc_events;
static std::mutex g_tc_mu;

void tc_profile_enable(bool on) {
    std::lock_guard<std::mutex> lk(g_tc_mu);
    g_tc_profile = on;
    for (auto& e : g_tc_events) { cudaEventDestroy(e.a); cudaEventDestroy(e.b); }
    g_tc_events.clear();
}
// accumulated milliseconds per kernel class {stash forward, reverse sweep, dW contraction}; clears the record
int tc_profile_read(double* ms, int n) {
    std::lock_guard<std::mutex> lk(g_tc_mu);
    for (int i = 0; i < n; ++i) ms[i] = 0.0;
    for (auto& e : g_tc_events) {
        if (cudaEventSynchronize(e.b) != cudaSuccess) { set_error("tc profile: event sync failed"); return -4; }
        float t = 0.f;
        cudaEventElapsedTime(&t, e.a, e.b);
        if (e.kind < n) ms[e.kind] += t;
        cudaEventDestroy(e.a); cudaEventDestroy(e.b);
    }
    g_tc_events.clear();
    return 0;
}
struct TcScope {
    bool on; TcEv ev; cudaStream_t s;
    TcScope(int kind, cudaStream_t st) : on(g_tc_profile), s(st) {
        if (on) { cudaEventCreate(&ev.a); cudaEventCreate(&ev.b); ev.kind = kind; cudaEventRecord(ev.a, s); }
    }
    ~TcScope() {
        if (on) { cudaEventRecord(ev.b, s); std::lock_guard<std::mutex> lk(g_tc_mu); g_tc_events.push_back(ev); }
    }
};

// tcgen05 accumulation truncates: every MMA rounds the fp32 accumulator toward zero, so a hidden-layer contraction (32-48 MMAs)
// comes out low by a few 2^-24 -- always in the same direction, which the residual of the loss (differences of nearly equal
// terms) amplifies.  The forward epilogue multiplies what it reads from tensor memory by 1 + beta 2^-24.  Measured against the
// fp64 oracle (scripts/debias_sweep.py), gradient of the GPE2D golden / jets of random networks:
//   three tf32 products   beta 0: 7.8e-6   beta 4: 1.0e-6 (optimum; jets 1.3-1.6x better; per-case optima between 4 and 10)
//   bf16 corrections      beta 0: 1.19e-5  beta 8: 1.4e-6 (jets of [1,128x4,1] 2.7x better, the 2-D nets unchanged)
// beta = GPE_TC_DEBIAS_FWD (read per call), default kTcDebiasFwd.  The reverse sweep's adjoints gain nothing measurable.
static float tc_debias() {
    const char* ev = getenv("GPE_TC_DEBIAS_FWD");
    const float beta = ev ? (float)atof(ev) : tcp::kTcDebiasFwd;
    return 1.f + beta * 5.9604644775390625e-8f;
}
static int tc_env_mode(const char* name, int dflt);
static int tc_interleave() {
    static int v = -1;
    if (v < 0) v = tc_env_mode("GPE_TC_INTERLEAVE", 0);
    return v;
}

// MMA issue order (TcParams::interleave).  Forward / stash-forward kernel: GPE_TC_INTERLEAVE_FWD = 0 | 1, default 1;
// reverse sweep: GPE_TC_INTERLEAVE = 0 | 1, default 0 (a non-zero value also sets the forward's unless that is given).
static int tc_env_mode(const char* name, int dflt) {
    const char* ev = getenv(name);
    return (ev && ev[0] >= '0' && ev[0] <= '2') ? ev[0] - '0' : dflt;
}
static int tc_interleave_fwd() {
    static int v = -1;
    if (v < 0) { const int r = tc_interleave(); v = tc_env_mode("GPE_TC_INTERLEAVE_FWD", 1) | r; }
    return v;
}

template <int D, int ACT>
static int bwd_tc_act(const tcp::TcParams& Pf, const tcp::TcParams& Pr, const float* x, long long n, const float* ub, const float* gb,
                  const float* hb, float* stash, const float* resident, float* zb, double* dWacc, float* small_acc, long long chunk, cudaStream_t s,
                  const unsigned int** abort_flag) {
    const int sms = tc_sms(), L = Pf.n_hidden;
    const size_t smem = tcp::tc_smem_bytes(D, L), smem_dw = tcp::dw_smem_bytes();
    const size_t smem_max = tcp::tc_smem_bytes(D, tcp::kTcMaxMat + 1);
    auto fwd_kern = tc_two_tiles() ? tcp::jet_fwd2_tc_kernel<D, true, ACT> : tcp::jet_fwd_tc_kernel<D, true, ACT>;
    if (int rc = ensure_smem(fwd_kern, smem_max, "jet_fwd_tc_kernel")) return rc;
    auto rev_kern = tc_two_tiles_rev() ? tcp::jet_rev2_tc_kernel<D, ACT> : tcp::jet_rev_tc_kernel<D, ACT>;
    if (int rc = ensure_smem(rev_kern, smem_max, "jet_rev_tc_kernel")) return rc;
    if (int rc = ensure_smem(tcp::dw_tc_kernel<D, ACT>, smem_dw, "dw_tc_kernel")) return rc;
    cudaError_t e;
    static int rev_pf = -1;   // GPE_REV_PF: L2 prefetch distance of the reverse sweep, in (tile, layer) steps
    if (rev_pf < 0) { const char* ev = getenv("GPE_REV_PF"); rev_pf = ev ? atoi(ev) : 1; if (rev_pf < 1) rev_pf = 1; }
    // ---- fused backward: one persistent launch, zbar handed from the sweep to dW through an L2-resident ring ----
    // GPE_BWD_FUSED=0 keeps the per-chunk kernels; GPE_FUSED_NRC = reverse-sweep clusters (of sms/2), GPE_RING_TILES = ring
    // slots of 64 points.  Needs the resident stash and enough points to fill the machine.
    // (read per call, so a test can switch between the two backward schemes inside one process)
    int fused_env = 1, nrc_env = 0, ring_env = 0;
    {
        const char* ev = getenv("GPE_BWD_FUSED"); fused_env = (ev && ev[0] == '1') ? 1 : 0;
        ev = getenv("GPE_FUSED_NRC"); nrc_env = ev ? atoi(ev) : 0;
        ev = getenv("GPE_RING_TILES"); ring_env = ev ? atoi(ev) : 0;
    }
    if (abort_flag) *abort_flag = nullptr;
    const int total_clusters = sms / 2;
    if (resident && fused_env && n >= 65536 && total_clusters >= 8) {
        const size_t smem_f = smem > smem_dw ? smem : smem_dw;
        const size_t smem_fmax = smem_max > smem_dw ? smem_max : smem_dw;
        const bool two_rev = tc_two_tiles_rev();
        auto fused_kern = two_rev ? tcp::bwd_fused_tc_kernel<D, ACT, true> : tcp::bwd_fused_tc_kernel<D, ACT, false>;
        if (int rc = ensure_smem(fused_kern, smem_fmax, "bwd_fused_tc_kernel")) return rc;
        // reverse sweep : dW contraction ~ 5 : 3 in SM time when run one after the other; the sweep keeps two thirds
        int nrc = nrc_env > 0 ? nrc_env : (total_clusters * 4 + 3) / 7;
        if (nrc > total_clusters - (L - 1 + 1) / 2) nrc = total_clusters - (L - 1 + 1) / 2;   // at least L-1 dW CTAs
        if (nrc < 1) nrc = 1;
        const long long pairs = (n + 2 * tcp::kTcPts - 1) / (2 * tcp::kTcPts);
        int ns = ring_env > 0 ? ring_env : (two_rev ? 160 : 96);
        if (ns < (two_rev ? 2 : 1) * nrc + 16) ns = (two_rev ? 2 : 1) * nrc + 16;   // every sweep pair holds one (two) slot(s) at a time
        const long long cap_tiles = chunk / (2 * tcp::kTcPts);
        // the ring lives in the zbar chunk buffer; the counters sit behind it
        const size_t n_ctr = (size_t)(L - 1) * 2 * nrc + 2 * (size_t)ns + 4;
        const size_t ring_floats = (size_t)(L - 1) * ns * 8 * tcp::kBlkFloats;
        if (ns < cap_tiles && ring_floats * 4 + n_ctr * 4 <= tc_zb_bytes(L, n)) {
            unsigned int* ctr = reinterpret_cast<unsigned int*>(zb + ring_floats);
            e = cudaMemsetAsync(ctr, 0, n_ctr * 4, s);
            if (e != cudaSuccess) { set_error("fused backward: memset: %s", cudaGetErrorString(e)); return -4; }
            tcp::RingCtl ring{ns, nrc, ctr, ctr + (size_t)(L - 1) * 2 * nrc, ctr + (size_t)(L - 1) * 2 * nrc + 2 * (size_t)ns};
            const int nblk8 = (int)((n + tcp::kDwPts - 1) / tcp::kDwPts);
            const long long stash_ls = (long long)nblk8 * tcp::kBlkFloats;
            const int n_dw = 2 * (total_clusters - nrc);
            static int n1_env = -1;
            if (n1_env < 0) { const char* ev = getenv("GPE_DW_N1"); n1_env = ev ? atoi(ev) : 0; }
            {
                TcScope t(3, s);
                fused_kern<<<(unsigned)(2 * total_clusters), tcp::kFusedThreads, smem_f, s>>>(
                    Pr, x, n, ub, gb, hb, resident, zb, dWacc, small_acc, stash_ls, nblk8, (int)pairs, rev_pf, n1_env > 0 ? n1_env : n_dw / (L - 1), ring);
            }
            e = cudaGetLastError();
            if (e != cudaSuccess) { set_error("fused backward launch: %s", cudaGetErrorString(e)); return -4; }
            if (abort_flag) *abort_flag = ring.abort;
            return 0;
        }
    }
    for (long long c0 = 0; c0 < n; c0 += chunk) {
        const long long nc = n - c0 < chunk ? n - c0 : chunk;
        const long long pairs = (nc + 2 * tcp::kTcPts - 1) / (2 * tcp::kTcPts);
        long long clusters = sms / 2;
        if (pairs < clusters) clusters = pairs;
        const float* xc = x + c0 * D;
        const int nblk8 = (int)((nc + tcp::kDwPts - 1) / tcp::kDwPts);   // image blocks (8 points) per layer of this chunk
        // resident stash (written by the forward of the same inputs, all n points): skip the recomputation and read the
        // chunk's blocks in place; otherwise recompute this chunk's stash into the workspace
        const long long stash_ls = resident ? (long long)((n + 7) / 8) * tcp::kBlkFloats : (long long)nblk8 * tcp::kBlkFloats;
        const float* st = resident ? resident + (size_t)(c0 / tcp::kDwPts) * tcp::kBlkFloats : stash;
        if (!resident) {
            TcScope t(0, s);
            fwd_kern<<<(unsigned)(2 * clusters), tcp::kTcThreads, smem, s>>>(Pf, xc, nc, nullptr, nullptr, nullptr, stash, stash_ls, nblk8, (int)pairs);
        }
        {
            TcScope t(1, s);
            rev_kern<<<(unsigned)(2 * clusters), tcp::kTcThreads, smem, s>>>(
                Pr, xc, nc, ub ? ub + c0 : nullptr, gb ? gb + c0 * D : nullptr, hb ? hb + c0 : nullptr, st, zb, small_acc, stash_ls, nblk8, (int)pairs, rev_pf);
        }
        const long long blocks = nblk8;
        long long want = blocks * (L - 1);                 // every CTA owns one hidden matrix (dw_class)
        if (want > sms) want = sms;
        const int grid = (int)(want < L - 1 ? L - 1 : want);
        static int n1_env = -1;   // GPE_DW_N1: CTAs of class l = 1 (tuning knob); default = an equal share
        if (n1_env < 0) { const char* ev = getenv("GPE_DW_N1"); n1_env = ev ? atoi(ev) : 0; }
        const int n0 = n1_env > 0 ? n1_env : grid / (L - 1);
        static int pf_env = -1;   // GPE_DW_PF: L2 prefetch distance of the dW operand blocks (blocks ahead)
        if (pf_env < 0) { const char* ev = getenv("GPE_DW_PF"); pf_env = ev ? atoi(ev) : tcp::kDwPf; }
        {
            TcScope t(2, s);
            tcp::dw_tc_kernel<D, ACT><<<grid, tcp::kDwThreads, smem_dw, s>>>(L, Pf.act, Pf.small, xc, nc, stash_ls, (long long)nblk8 * tcp::kBlkFloats, st, zb, dWacc, small_acc, nblk8, n0, pf_env);
        }
        e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("tensor-path backward launch: %s", cudaGetErrorString(e)); return -4; }
    }
    return 0;
}

template <int D>
static int bwd_tc(const tcp::TcParams& Pf, const tcp::TcParams& Pr, const float* x, long long n, const float* ub, const float* gb,
                  const float* hb, float* stash, const float* resident, float* zb, double* dWacc, float* small_acc, long long chunk, cudaStream_t s,
                  const unsigned int** abort_flag) {
    return (Pf.act == ACT_TANH && !tc_act_generic()) ? bwd_tc_act<D, ACT_TANH>(Pf, Pr, x, n, ub, gb, hb, stash, resident, zb, dWacc, small_acc, chunk, s, abort_flag)
                              : bwd_tc_act<D, -1>(Pf, Pr, x, n, ub, gb, hb, stash, resident, zb, dWacc, small_acc, chunk, s, abort_flag);
}

// chunk of points processed per (stash forward, reverse sweep, dW contraction) round
long long tc_bwd_chunk(long long n) {
    static long long cap = 0;   // GPE_TC_CHUNK (points, multiple of 64; read once): zbar + recompute-stash workspace = 12 KB per point of it
    if (cap == 0) {
        const char* ev = getenv("GPE_TC_CHUNK");
        long long v = ev ? atoll(ev) : 0;
        cap = v >= 64 ? (v + 63) / 64 * 64 : 1048576;   // measured at 4M points: 22.6 ms (262144) / 22.1 (1M) / 22.05 (4M)
    }
    long long c = (n + 63) / 64 * 64;
    return c < cap ? (c > 0 ? c : 64) : cap;
}
// image blocks: [layer][chunk / 8][128 features x 8 points x 4 streams]; st_l and zbar_l of the hidden matrices l = 1..L-1
size_t tc_stash_bytes(int L, long long n) { return (size_t)(L - 1) * (size_t)tc_bwd_chunk(n) * 512 * sizeof(float); }
size_t tc_zb_bytes(int L, long long n) { return (size_t)(L - 1) * (size_t)tc_bwd_chunk(n) * 512 * sizeof(float); }

int launch_bwd_tc(int d, int L, int act, const float* wimg_fwd, const float* wimg_rev, const float* small, const float* x,
                  long long n, const float* ub, const float* gb, const float* hb, float* stash, const float* resident, float* zb,
                  double* dWacc, float* small_acc, cudaStream_t s, const unsigned int** abort_flag) {
    tcp::TcParams Pf{L, act, wimg_fwd, small, {1.f, 1.f}, {0.f, 0.f}, tc_debias(), tc_interleave_fwd()},
        Pr{L, act, wimg_rev, small, {1.f, 1.f}, {0.f, 0.f}, 1.f, tc_interleave()};
    const long long chunk = tc_bwd_chunk(n);
    return d == 1 ? bwd_tc<1>(Pf, Pr, x, n, ub, gb, hb, stash, resident, zb, dWacc, small_acc, chunk, s, abort_flag)
                  : bwd_tc<2>(Pf, Pr, x, n, ub, gb, hb, stash, resident, zb, dWacc, small_acc, chunk, s, abort_flag);
}

// bytes of a resident stash of ALL n points: st_l image blocks of the hidden matrices l = 1..L-1
size_t tc_resident_stash_bytes(int L, long long n) { return (size_t)(L - 1) * (size_t)((n + 7) / 8) * tcp::kBlkFloats * sizeof(float); }

int launch_fwd_tc(int d, int L, int act, const float* wimg, const float* small, const float* x, long long n, float* u,
                  float* g, float* h, float* resident, cudaStream_t s, const float* xa, const float* xb) {
    // xa / xb (value-only calls: g = h = nullptr, no stash): the affine input map of the mirrored points
    tcp::TcParams P{L, act, wimg, small, {xa ? xa[0] : 1.f, xa ? xa[1] : 1.f}, {xb ? xb[0] : 0.f, xb ? xb[1] : 0.f}, tc_debias(), tc_interleave_fwd()};
    if (resident) return d == 1 ? fwd_tc<1, true>(P, x, n, u, g, h, resident, s) : fwd_tc<2, true>(P, x, n, u, g, h, resident, s);
    return d == 1 ? fwd_tc<1, false>(P, x, n, u, g, h, nullptr, s) : fwd_tc<2, false>(P, x, n, u, g, h, nullptr, s);
}

}  // namespace gpe

#ifdef GPE_TC_TRACE
extern "C" int gpe_debug_trace(long long* out, int n) {
    return (int)cudaMemcpyFromSymbol(out, gpe::tcp::g_tc_trace, sizeof(long long) * (n < 4096 ? n : 4096));
}
#endif
