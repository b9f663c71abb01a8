// Host-side declarations shared by the per-width launch units and the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include "jet_common.cuh"

namespace gpe {

struct PtrArray {
    const float* p[kMaxLayers + 1];
};
struct MutPtrArray {
    float* p[kMaxLayers + 1];
};

struct JetPlan {          // resolved launch geometry for (d, jet, WP, L)
    int pts_per_tile;
    int rows_per_tile;
    size_t smem_fwd, smem_bwd;
    size_t stash_floats_per_cta;
    int max_ctas_fwd, max_ctas_bwd;   // persistent grid bound (SMs x occupancy)
};

// implemented once per padded width (jet_launch.cu compiled with -DGPE_WP=...)
#define GPE_DECL_WP(WP)                                                                                          \
    int plan_wp##WP(int d, int jet, int L, bool res, JetPlan* plan);                                                     \
    int launch_fwd_wp##WP(int d, int jet, const MlpParams& P, const float* x, long long n, const XForm& xf,      \
                          float* u, float* g, float* h, int grid, int num_tiles, cudaStream_t s);                \
    int launch_bwd_wp##WP(int d, int jet, const MlpParams& P, const float* x, long long n, const XForm& xf,      \
                          const float* ub, const float* gb, const float* hb, float* stash, float* dWacc,         \
                          float* small_acc, int grid, int num_tiles, cudaStream_t s);
GPE_DECL_WP(64)
GPE_DECL_WP(128)
GPE_DECL_WP(256)
#undef GPE_DECL_WP

void set_error(const char* fmt, ...);

// NVTX ranges around the C-ABI calls (GPE_NVTX=1): fwd-jet / moments / adjoints / reverse show up as named ranges in an
// nsys or ncu timeline.  Header-only NVTX3; one getenv per process.
struct NvtxRange {
    bool on;
    explicit NvtxRange(const char* name) : on(enabled()) { if (on) nvtxRangePushA(name); }
    ~NvtxRange() { if (on) nvtxRangePop(); }
    static bool enabled() {
        static int v = -1;
        if (v < 0) { const char* e = getenv("GPE_NVTX"); v = (e && e[0] == '1') ? 1 : 0; }
        return v == 1;
    }
};


// tensor-core (tcgen05) path, jet_launch_tc.cu
bool tc_eligible(int d, int jet, int width, int L);
size_t tc_wimg_bytes(int L);
int tc_pack_wimg(const PtrArray& W, int L, int Wd, int transpose, float* wimg, cudaStream_t s);
void tc_profile_enable(bool on);
int tc_profile_read(double* ms, int n);
size_t tc_stash_bytes(int L, long long n);
size_t tc_zb_bytes(int L, long long n);
int launch_bwd_tc(int d, int L, int act, const float* wimg_fwd, const float* wimg_rev, const float* small, const float* x,
                  long long n, const float* ub, const float* gb, const float* hb, float* stash, const float* resident, float* zb,
                  double* dWacc, float* small_acc, cudaStream_t s, const unsigned int** abort_flag);
int launch_fwd_tc(int d, int L, int act, const float* wimg, const float* small, const float* x, long long n, float* u,
                  float* g, float* h, float* resident, cudaStream_t s, const float* xa = nullptr, const float* xb = nullptr);
size_t tc_resident_stash_bytes(int L, long long n);

}  // namespace gpe
