// Per-width instantiation + launch of the jet kernels.  Compiled once per GPE_WP.
#include "jet_host.h"
#include "jet_kernels.cuh"

#ifndef GPE_WP
#error "compile with -DGPE_WP=<padded width>"
#endif
#define GPE_CAT2(a, b) a##b
#define GPE_CAT(a, b) GPE_CAT2(a, b)

namespace gpe {
namespace {

constexpr int WP = GPE_WP;

int num_sms() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 0;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    return sms;
}

int launch_status(const char* what, const void* fn, int grid, size_t smem) {
    const cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) return 0;
    cudaFuncAttributes a{};
    cudaFuncGetAttributes(&a, fn);
    (void)cudaGetLastError();
    set_error("%s launch: %s (grid %d, block %d, dyn smem %zu, max dyn smem %d, regs %d, local %zu)", what,
              cudaGetErrorString(e), grid, kThreads, smem, a.maxDynamicSharedSizeBytes, a.numRegs, a.localSizeBytes);
    return -4;
}

template <int D, int JET, bool RES = false>
struct Unit {
    using C = TileCfg<D, JET, WP>;
    static int plan(int L, JetPlan* pl) {
        pl->pts_per_tile = C::PTS;
        pl->rows_per_tile = C::M;
        pl->smem_fwd = C::smem_fwd(L, RES);
        pl->smem_bwd = C::smem_bwd(L, RES);
        pl->stash_floats_per_cta = C::stash_floats_per_cta(L);
        const int sms = num_sms();
        if (sms <= 0) { set_error("no CUDA device"); return -2; }
        cudaError_t e;
        e = cudaFuncSetAttribute(jet_fwd_kernel<D, JET, WP, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_fwd(kMaxLayers, RES));
        if (e != cudaSuccess) { set_error("fwd smem attr: %s", cudaGetErrorString(e)); return -3; }
        e = cudaFuncSetAttribute(jet_bwd_kernel<D, JET, WP, RES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::smem_bwd(kMaxLayers, RES));
        if (e != cudaSuccess) { set_error("bwd smem attr: %s", cudaGetErrorString(e)); return -3; }
        int of = 0, ob = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&of, jet_fwd_kernel<D, JET, WP, RES>, kThreads, pl->smem_fwd);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ob, jet_bwd_kernel<D, JET, WP, RES>, kThreads, pl->smem_bwd);
        if (of < 1 || ob < 1) { set_error("kernel does not fit on an SM (occ %d/%d)", of, ob); return -3; }
        pl->max_ctas_fwd = of * sms;
        pl->max_ctas_bwd = ob * sms;
        return 0;
    }
    static int fwd(const MlpParams& P, const float* x, long long n, const XForm& xf, float* u, float* g, float* h,
                   int grid, int num_tiles, cudaStream_t s) {
        jet_fwd_kernel<D, JET, WP, RES><<<grid, kThreads, C::smem_fwd(P.n_hidden, RES), s>>>(P, x, n, xf, u, g, h, num_tiles);
        return launch_status("jet_fwd_kernel", (const void*)jet_fwd_kernel<D, JET, WP, RES>, grid, C::smem_fwd(P.n_hidden, RES));
    }
    static int bwd(const MlpParams& P, const float* x, long long n, const XForm& xf, const float* ub, const float* gb,
                   const float* hb, float* stash, float* dWacc, float* small_acc, int grid, int num_tiles,
                   cudaStream_t s) {
        jet_bwd_kernel<D, JET, WP, RES><<<grid, kThreads, C::smem_bwd(P.n_hidden, RES), s>>>(P, x, n, xf, ub, gb, hb, stash, dWacc,
                                                                                  small_acc, num_tiles);
        return launch_status("jet_bwd_kernel", (const void*)jet_bwd_kernel<D, JET, WP, RES>, grid, C::smem_bwd(P.n_hidden, RES));
    }
};

// residual-block nets (R/box_to_gaussian_pinn_simulation.py) are 1-D; widths up to 128 (the skip image fits beside
// the two activation tiles)
#if GPE_WP <= 128
#define GPE_DISPATCH_RES(CALL)                                                     \
    if (res && d == 1 && jet == JET_VALUE) return Unit<1, JET_VALUE, true>::CALL;  \
    if (res && d == 1) return Unit<1, JET_DIAG, true>::CALL;                       \
    if (res) { set_error("residual-block networks: d = 1 only (got d=%d)", d); return -1; }
#else
#define GPE_DISPATCH_RES(CALL)                                                     \
    if (res) { set_error("residual-block networks: hidden width <= 128"); return -1; }
#endif

#define GPE_DISPATCH(CALL)                                                         \
    GPE_DISPATCH_RES(CALL)                                                         \
    if (d == 1 && jet == JET_VALUE) return Unit<1, JET_VALUE>::CALL;               \
    if (d == 1 && (jet == JET_DIAG || jet == JET_FULL || jet == JET_LAPLACIAN))    \
        return Unit<1, JET_DIAG>::CALL;                                            \
    if (d == 2 && jet == JET_VALUE) return Unit<2, JET_VALUE>::CALL;               \
    if (d == 2 && jet == JET_DIAG) return Unit<2, JET_DIAG>::CALL;                 \
    if (d == 2 && jet == JET_FULL) return Unit<2, JET_FULL>::CALL;                 \
    if (d == 2 && jet == JET_LAPLACIAN) return Unit<2, JET_LAPLACIAN>::CALL;       \
    set_error("unsupported (d=%d, jet=%d)", d, jet);                               \
    return -1;

}  // namespace

int GPE_CAT(plan_wp, GPE_WP)(int d, int jet, int L, bool res, JetPlan* plan) { GPE_DISPATCH(plan(L, plan)) }

int GPE_CAT(launch_fwd_wp, GPE_WP)(int d, int jet, const MlpParams& P, const float* x, long long n, const XForm& xf,
                                   float* u, float* g, float* h, int grid, int num_tiles, cudaStream_t s) {
    const bool res = P.act == ACT_RESBLOCK_TANH;
    GPE_DISPATCH(fwd(P, x, n, xf, u, g, h, grid, num_tiles, s))
}

int GPE_CAT(launch_bwd_wp, GPE_WP)(int d, int jet, const MlpParams& P, const float* x, long long n, const XForm& xf,
                                   const float* ub, const float* gb, const float* hb, float* stash, float* dWacc,
                                   float* small_acc, int grid, int num_tiles, cudaStream_t s) {
    const bool res = P.act == ACT_RESBLOCK_TANH;
    GPE_DISPATCH(bwd(P, x, n, xf, ub, gb, hb, stash, dWacc, small_acc, grid, num_tiles, s))
}

}  // namespace gpe
