// Standalone B200 probe of the tcgen05 building blocks the tensor-path kernels rely on.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo tc_probe.cu -o tc_probe
// T1: fragment layouts of tcgen05.ld/st 16x256b / 16x128b against 32x32b
// T2: kind::tf32 MMA correctness (SS / TS, K-major / MN-major B, cta_group 1 / 2, 1 or 3 products)
// T3: MMA issue throughput (cycles per instruction)
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cstring>
#include "../../gross-pitaevskii-eigenvalue-problem_b200/csrc/tc_tmem.cuh"

using namespace tc;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

// ------------------------------------------------------------------------------------------ T1
__global__ void __launch_bounds__(128) layout_probe(uint32_t* out) {
    __shared__ uint32_t tbase_s;
    const int tid = threadIdx.x, w = tid / 32, l = tid % 32;
    if (w == 0) tmem_alloc<1>(&tbase_s, 32);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tbase_s;
    uint32_t v[8];
    for (int c = 0; c < 8; ++c) v[c] = ((32 * w + l) << 8) | c;
    tmem_st_32x32b_x8(tb + ((uint32_t)(32 * w) << 16), v);
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t a[4], b[4], c2[2], d2[2];
    tmem_ld_16x256b_x1(tb + ((uint32_t)(32 * w) << 16), a);
    tmem_ld_16x256b_x1(tb + ((uint32_t)(32 * w + 16) << 16), b);
    tmem_ld_16x128b_x1(tb + ((uint32_t)(32 * w) << 16), c2);
    tmem_ld_16x128b_x1(tb + ((uint32_t)(32 * w) << 16) + 4, d2);
    tmem_wait_ld();
    for (int i = 0; i < 4; ++i) out[tid * 12 + i] = a[i], out[tid * 12 + 4 + i] = b[i];
    out[tid * 12 + 8] = c2[0], out[tid * 12 + 9] = c2[1], out[tid * 12 + 10] = d2[0], out[tid * 12 + 11] = d2[1];
    // write-back test: store with 16x256b what we loaded, shifted by 8 columns, then re-read via 32x32b
    tmem_st_16x256b_x1(tb + ((uint32_t)(32 * w) << 16) + 8, a);
    tmem_st_16x256b_x1(tb + ((uint32_t)(32 * w + 16) << 16) + 8, b);
    tmem_wait_st();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t r[8];
    tmem_ld_32x32b_x8(tb + ((uint32_t)(32 * w) << 16) + 8, r);
    tmem_wait_ld();
    for (int c = 0; c < 8; ++c) out[128 * 12 + tid * 8 + c] = r[c];
    tc_fence_before();
    __syncthreads();
    if (w == 0) tmem_dealloc<1>(tb, 32);
}

// ------------------------------------------------------------------------------------------ T2
// image of a [rows x cols] fp32 matrix in the SWIZZLE_NONE canonical layout: 8x(16 B) core matrices,
// cores contiguous along the column direction (LBO = 128 B), row groups at SBO = (cols/4)*128 B.
__device__ __forceinline__ uint32_t img_off(int r, int c, int cols) {
    return (uint32_t)((r >> 3) * (cols / 4) * 128 + (c >> 2) * 128 + (r & 7) * 16 + (c & 3) * 4);
}
__device__ void load_img(unsigned char* dst, const float* src, int rows, int cols, int ld, int tid) {
    for (int i = tid; i < rows * cols; i += 128) {
        const int r = i / cols, c = i % cols;
        *reinterpret_cast<float*>(dst + img_off(r, c, cols)) = src[(size_t)r * ld + c];
    }
}

template <int CG, bool A_TMEM, int B_MN, int NPROD, int N>
__global__ void __launch_bounds__(128) gemm_probe(const float* A0, const float* A1, const float* B0, const float* B1, float* D,
                                                  long long* cycles) {
    constexpr int K = 128, NB = N / CG;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tbase_s;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, w = tid / 32, l = tid % 32;
    const uint32_t cta = CG == 2 ? cluster_ctarank() : 0u;
    unsigned char* sA0 = smem;
    unsigned char* sA1 = sA0 + (A_TMEM ? 0 : 128 * K * 4);
    unsigned char* sB0 = sA1 + ((A_TMEM || NPROD == 1) ? 0 : 128 * K * 4);
    unsigned char* sB1 = sB0 + NB * K * 4;
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    if (w == 0) tmem_alloc<CG>(&tbase_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = tbase_s;
    {
        uint32_t dyn;
        asm volatile("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn));
        for (uint32_t i = tid; i < dyn / 4; i += 128) reinterpret_cast<float*>(smem)[i] = 1000.f;
        __syncthreads();
    }
    const float* a0 = A0 + (size_t)cta * 128 * K;
    const float* a1 = A1 + (size_t)cta * 128 * K;
    if (A_TMEM) {
        for (int p = 0; p < (NPROD == 3 ? 2 : 1); ++p) {
            const float* a = (p ? a1 : a0) + (size_t)tid * K;
            for (int c0 = 0; c0 < K; c0 += 8) {
                uint32_t v[8];
                for (int c = 0; c < 8; ++c) v[c] = __float_as_uint(a[c0 + c]);
                tmem_st_32x32b_x8(tb + ((uint32_t)(32 * w) << 16) + 256 + 128 * p + c0, v);
            }
        }
        tmem_wait_st();
    } else {
        load_img(sA0, a0, 128, K, K, tid);
        if (NPROD == 3) load_img(sA1, a1, 128, K, K, tid);
    }
    if (!B_MN) {   // B given as [N][K]; this CTA holds rows cta*NB ..
        load_img(sB0, B0 + (size_t)cta * NB * K, NB, K, K, tid);
        if (NPROD == 3) load_img(sB1, B1 + (size_t)cta * NB * K, NB, K, K, tid);
    } else {       // B given as [K][N] (k rows, n contiguous); this CTA holds columns cta*NB ..
        load_img(sB0, B0 + cta * NB, K, NB, N, tid);
        if (NPROD == 3) load_img(sB1, B1 + cta * NB, K, NB, N, tid);
    }
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    tc_fence_after();
    long long t0 = 0, t1 = 0;
    if (tid == 0 && cta == 0) {
        constexpr uint32_t idesc = instr_desc(2, 128 * CG, N, 0, B_MN ? 1 : 0);
        t0 = clock64();
        for (int j = 0; j < K / 8; ++j) {
            for (int p = 0; p < NPROD; ++p) {
                unsigned char* sa = (p == 1) ? sA1 : sA0;
                unsigned char* sb = (p == 2) ? sB1 : sB0;
                uint64_t bdesc;
                if (!B_MN) bdesc = smem_desc(smem_u32(sb) + j * 256, 128, (K / 4) * 128);
                else if (B_MN == 1) bdesc = smem_desc(smem_u32(sb) + j * (NB / 4) * 128, (NB / 4) * 128, 128);
                else bdesc = smem_desc(smem_u32(sb) + j * (NB / 4) * 128, 128, (NB / 4) * 128);
                const uint32_t acc = (j | p) ? 1u : 0u;
                if (A_TMEM) mma_ts<0, CG>(tb, tb + 256 + (p == 1 ? 128 : 0) + 8 * j, bdesc, idesc, acc);
                else mma_ss<0, CG>(tb, smem_desc(smem_u32(sa) + j * 256, 128, (K / 4) * 128), bdesc, idesc, acc);
            }
        }
        if (CG == 2) mma_commit_pair(&bar, 3); else mma_commit(&bar);
    }
    mbar_wait(&bar, 0);
    tc_fence_after();
    if (tid == 0 && cta == 0) { t1 = clock64(); if (cycles) *cycles = t1 - t0; }
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tb + ((uint32_t)(32 * w) << 16) + c0, r);
        tmem_wait_ld();
        for (int c = 0; c < 32; ++c) D[((size_t)cta * 128 + tid) * N + c0 + c] = __uint_as_float(r[c]);
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    if (w == 0) tmem_dealloc<CG>(tb, 512);
}

// ------------------------------------------------------------------------------------------ T5
// One launch, clusters of 2: cluster 0 runs a cta_group::2 MMA (M = 256), the two CTAs of cluster 1 each run their own
// cta_group::1 MMA (M = 128).  Question: may one kernel hold both cta_group values (role-specialised CTAs)?
template <int CG>
__device__ void mix_body(const float* A, const float* B, float* D, uint32_t cta, unsigned char* smem, uint32_t* tbase_s, uint64_t* bar) {
    constexpr int K = 128, N = 128, NB = N / CG;
    const int tid = threadIdx.x, w = tid / 32;
    if (tid == 0) { mbar_init(bar, 1); mbar_fence_init(); }
    if (w == 0) tmem_alloc<CG>(tbase_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tb = *tbase_s;
    const float* a = A + (size_t)cta * 128 * K + (size_t)tid * K;
    for (int c0 = 0; c0 < K; c0 += 8) {
        uint32_t v[8];
        for (int c = 0; c < 8; ++c) v[c] = __float_as_uint(a[c0 + c]);
        tmem_st_32x32b_x8(tb + ((uint32_t)(32 * w) << 16) + 256 + c0, v);
    }
    tmem_wait_st();
    load_img(smem, B + (size_t)cta * NB * K, NB, K, K, tid);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    tc_fence_after();
    if (tid == 0 && cta == 0) {
        constexpr uint32_t idesc = instr_desc(2, 128 * CG, N, 0, 0);
        for (int j = 0; j < K / 8; ++j)
            mma_ts<0, CG>(tb, tb + 256 + 8 * j, smem_desc(smem_u32(smem) + j * 256, 128, (K / 4) * 128), idesc, j ? 1u : 0u);
        if (CG == 2) mma_commit_pair(bar, 3); else mma_commit(bar);
    }
    mbar_wait(bar, 0);
    tc_fence_after();
    for (int c0 = 0; c0 < N; c0 += 32) {
        uint32_t r[32];
        tmem_ld_32x32b_x32(tb + ((uint32_t)(32 * w) << 16) + c0, r);
        tmem_wait_ld();
        for (int c = 0; c < 32; ++c) D[((size_t)cta * 128 + tid) * N + c0 + c] = __uint_as_float(r[c]);
    }
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    if (w == 0) tmem_dealloc<CG>(tb, 512);
}
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128) mix_probe(const float* A, const float* B, float* D) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tbase_s;
    __shared__ __align__(8) uint64_t bar;
    const uint32_t cl = blockIdx.x >> 1, rank = cluster_ctarank();
    if (cl == 0) mix_body<2>(A, B, D, rank, smem, &tbase_s, &bar);                                  // rows 0..255 of D
    else mix_body<1>(A + (size_t)(2 + rank) * 128 * 128 - 0, B, D + (size_t)(2 + rank) * 128 * 128, 0u, smem, &tbase_s, &bar);   // rows 256.. / 384..
}

// ------------------------------------------------------------------------------------------ T3
template <int CG, bool A_TMEM, int N, int KIND>
__global__ void __launch_bounds__(128) rate_probe(long long* cycles, int reps) {
    constexpr int NB = N / CG;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint32_t tbase_s;
    __shared__ __align__(8) uint64_t bar;
    const int tid = threadIdx.x, w = tid / 32;
    const uint32_t cta = CG == 2 ? cluster_ctarank() : 0u;
    for (int i = tid; i < (128 * 128 + NB * 128); i += 128) reinterpret_cast<float*>(smem)[i] = 0.f;
    if (tid == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    if (w == 0) tmem_alloc<CG>(&tbase_s, 512);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    if (CG == 2) cluster_sync();
    tc_fence_after();
    const uint32_t tb = tbase_s;
    if (tid == 0 && cta == 0) {
        constexpr uint32_t idesc = instr_desc(KIND == 0 ? 2 : 1, 128 * CG, N, 0, 0);
        unsigned char* sA = smem;
        unsigned char* sB = smem + 128 * 128 * 4;
        const long long t0 = clock64();
        for (int r = 0; r < reps; ++r)
            for (int j = 0; j < 16; ++j) {
                const uint64_t bdesc = smem_desc(smem_u32(sB) + j * 256, 128, 4096);
                if (A_TMEM) mma_ts<KIND, CG>(tb, tb + 256 + 8 * j, bdesc, idesc, 1u);
                else mma_ss<KIND, CG>(tb, smem_desc(smem_u32(sA) + j * 256, 128, 4096), bdesc, idesc, 1u);
            }
        if (CG == 2) mma_commit_pair(&bar, 3); else mma_commit(&bar);
        mbar_wait(&bar, 0);
        *cycles = clock64() - t0;
    } else {
        mbar_wait(&bar, 0);
    }
    tc_fence_after();
    __syncthreads();
    if (CG == 2) cluster_sync();
    if (w == 0) tmem_dealloc<CG>(tb, 512);
}

// ------------------------------------------------------------------------------------------ host
static float tf32_trunc(float x) { uint32_t u; memcpy(&u, &x, 4); u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }
static float tf32_rn(float x) { uint32_t u; memcpy(&u, &x, 4); u += 0x1000u; u &= 0xFFFFE000u; memcpy(&x, &u, 4); return x; }
static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }

template <typename Kern>
static void launch(Kern k, int cg, size_t smem, const float* A0, const float* A1, const float* B0, const float* B1, float* D, long long* cyc) {
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cg); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = 0;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cg; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    CK(cudaLaunchKernelEx(&cfg, k, A0, A1, B0, B1, D, cyc));
    CK(cudaDeviceSynchronize());
}

template <int CG, bool A_TMEM, int B_MN, int NPROD, int N>
static void run_gemm(const char* name) {
    const int M = 128 * CG, K = 128;
    std::vector<float> a(M * K), b(N * K), a0(M * K), a1(M * K), b0(N * K), b1(N * K);
    for (auto& v : a) v = frand();
    for (auto& v : b) v = frand();
    for (int i = 0; i < M * K; ++i) { a0[i] = NPROD == 3 ? tf32_rn(a[i]) : a[i]; a1[i] = a[i] - a0[i]; }
    for (int i = 0; i < N * K; ++i) { b0[i] = NPROD == 3 ? tf32_rn(b[i]) : b[i]; b1[i] = b[i] - b0[i]; }
    // b is logically [n][k]; for B_MN the device receives it transposed, [k][n]
    auto tr = [&](const std::vector<float>& s) { std::vector<float> t(N * K); for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) t[k * N + n] = s[n * K + k]; return t; };
    std::vector<float> b0d = B_MN ? tr(b0) : b0, b1d = B_MN ? tr(b1) : b1;
    float *dA0, *dA1, *dB0, *dB1, *dD; long long* dC;
    CK(cudaMalloc(&dA0, M * K * 4)); CK(cudaMalloc(&dA1, M * K * 4)); CK(cudaMalloc(&dB0, N * K * 4)); CK(cudaMalloc(&dB1, N * K * 4));
    CK(cudaMalloc(&dD, M * N * 4)); CK(cudaMalloc(&dC, 8));
    CK(cudaMemcpy(dA0, a0.data(), M * K * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dA1, a1.data(), M * K * 4, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dB0, b0d.data(), N * K * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB1, b1d.data(), N * K * 4, cudaMemcpyHostToDevice));
    CK(cudaMemset(dD, 0xFF, M * N * 4));
    size_t smem = (A_TMEM ? 0 : (NPROD == 3 ? 2 : 1) * 128 * K * 4) + 2 * (N / CG) * K * 4;
    launch(gemm_probe<CG, A_TMEM, B_MN, NPROD, N>, CG, smem, dA0, dA1, dB0, dB1, dD, dC);
    std::vector<float> d(M * N); long long cyc;
    CK(cudaMemcpy(d.data(), dD, M * N * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
    double e_exact = 0, e_trunc = 0, e_rn = 0, dmax = 0;
    for (int m = 0; m < M; ++m) for (int n = 0; n < N; ++n) {
        double se = 0, st = 0, sr = 0;
        for (int k = 0; k < K; ++k) {
            se += (double)a[m * K + k] * b[n * K + k];
            st += (double)tf32_trunc(a[m * K + k]) * tf32_trunc(b[n * K + k]);
            sr += (double)tf32_rn(a[m * K + k]) * tf32_rn(b[n * K + k]);
        }
        const double g = d[m * N + n];
        e_exact = fmax(e_exact, fabs(g - se)); e_trunc = fmax(e_trunc, fabs(g - st)); e_rn = fmax(e_rn, fabs(g - sr)); dmax = fmax(dmax, fabs(se));
    }
    printf("T2 %-34s max|D|=%.3f  err vs exact %.3e  vs tf32-trunc %.3e  vs tf32-rn %.3e  cycles(%d mma)=%lld\n", name, dmax, e_exact,
           e_trunc, e_rn, 16 * NPROD, cyc);
    cudaFree(dA0); cudaFree(dA1); cudaFree(dB0); cudaFree(dB1); cudaFree(dD); cudaFree(dC);
}

template <int CG, bool A_TMEM, int N, int KIND>
static void run_rate(const char* name) {
    long long* dC; CK(cudaMalloc(&dC, 8));
    auto k = rate_probe<CG, A_TMEM, N, KIND>;
    const size_t smem = (128 * 128 + (N / CG) * 128) * 4;
    CK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    for (int reps : {4, 64}) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(CG); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CG; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        CK(cudaLaunchKernelEx(&cfg, k, dC, reps));
        CK(cudaDeviceSynchronize());
        long long cyc; CK(cudaMemcpy(&cyc, dC, 8, cudaMemcpyDeviceToHost));
        printf("T3 %-34s reps=%3d  %lld cycles  = %.1f cycles/mma\n", name, reps, cyc, (double)cyc / (16.0 * reps));
    }
    cudaFree(dC);
}


template <int V>
static void run_diag() {
    const int M = 128, K = 128, N = 128;
    std::vector<float> a(M * K, 0.f), bn(N * K), bk(N * K), z(N * K, 0.f);
    for (int m = 0; m < M; ++m) a[m * K + m] = 1.f;
    for (int k = 0; k < K; ++k) for (int n = 0; n < N; ++n) { bn[k * N + n] = (float)n; bk[k * N + n] = (float)k; }   // device layout [k][n]
    float *dA, *dB, *dZ, *dD; long long* dC;
    CK(cudaMalloc(&dA, M * K * 4)); CK(cudaMalloc(&dB, N * K * 4)); CK(cudaMalloc(&dZ, N * K * 4)); CK(cudaMalloc(&dD, M * N * 4)); CK(cudaMalloc(&dC, 8));
    CK(cudaMemcpy(dA, a.data(), M * K * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dZ, z.data(), N * K * 4, cudaMemcpyHostToDevice));
    std::vector<float> dn(M * N), dk(M * N);
    for (int pass = 0; pass < 2; ++pass) {
        CK(cudaMemcpy(dB, (pass ? bk : bn).data(), N * K * 4, cudaMemcpyHostToDevice));
        launch(gemm_probe<1, true, V, 1, 128>, 1, 200 * 1024, dA, dZ, dB, dZ, dD, dC);
        CK(cudaMemcpy((pass ? dk : dn).data(), dD, M * N * 4, cudaMemcpyDeviceToHost));
    }
    printf("T4 MN-major variant %d: D[m][n] fetched B(n',k') -- expected (n,m)\n", V);
    for (int m : {0, 1, 2, 7, 8, 9, 16, 64}) {
        printf("  m=%3d:", m);
        for (int n : {0, 1, 2, 3, 4, 5, 8, 12, 16, 32, 64, 127}) printf(" n%d->(%g,%g)", n, dn[m * N + n], dk[m * N + n]);
        printf("\n");
    }
}

int main(int argc, char** argv) {
    const int which = argc > 1 ? atoi(argv[1]) : 0;
    const int sub = argc > 2 ? atoi(argv[2]) : -1;
    int idx = 0;
    if (which == 0 || which == 1) {
        uint32_t* dO; CK(cudaMalloc(&dO, (128 * 12 + 128 * 8) * 4));
        layout_probe<<<1, 128>>>(dO);
        CK(cudaDeviceSynchronize());
        std::vector<uint32_t> o(128 * 12 + 128 * 8);
        CK(cudaMemcpy(o.data(), dO, o.size() * 4, cudaMemcpyDeviceToHost));
        printf("T1 16x256b.x1 / 16x128b.x1 fragments (row,col) per thread, warp 1 shown:\n");
        for (int t = 32; t < 64; ++t) {
            printf("  t%2d 256b lo:", t - 32);
            for (int i = 0; i < 4; ++i) printf(" (%3u,%u)", o[t * 12 + i] >> 8, o[t * 12 + i] & 255);
            printf("  hi:");
            for (int i = 0; i < 4; ++i) printf(" (%3u,%u)", o[t * 12 + 4 + i] >> 8, o[t * 12 + 4 + i] & 255);
            printf("  128b@0:");
            for (int i = 0; i < 2; ++i) printf(" (%3u,%u)", o[t * 12 + 8 + i] >> 8, o[t * 12 + 8 + i] & 255);
            printf("  128b@4:");
            for (int i = 0; i < 2; ++i) printf(" (%3u,%u)", o[t * 12 + 10 + i] >> 8, o[t * 12 + 10 + i] & 255);
            printf("\n");
        }
        int bad = 0;
        for (int t = 0; t < 128; ++t) for (int c = 0; c < 8; ++c) bad += o[128 * 12 + t * 8 + c] != (uint32_t)((t << 8) | c);
        printf("T1 16x256b store round trip mismatches: %d\n", bad);
    }
    if (which == 0 || which == 2) {
        if (sub < 0 || sub == idx) run_gemm<1, false, 0, 1, 128>("cg1 SS  Kmajor  1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<1, true, 0, 1, 128>("cg1 TS  Kmajor  1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<1, true, 1, 1, 128>("cg1 TS  MNmajorB(v1) 1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<1, true, 2, 1, 128>("cg1 TS  MNmajorB(v2) 1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<1, true, 0, 3, 128>("cg1 TS  Kmajor  3xTF32 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<2, true, 0, 1, 128>("cg2 TS  Kmajor  1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<2, true, 2, 1, 128>("cg2 TS  MNmajorB(v2) 1prod N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<2, true, 0, 3, 128>("cg2 TS  Kmajor  3xTF32 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<2, true, 2, 3, 128>("cg2 TS  MNmajorB(v2) 3xTF32 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_gemm<2, false, 0, 1, 128>("cg2 SS  Kmajor  1prod N128");
        ++idx;
    }
    if (which == 0 || which == 3) {
        idx = 0;
        if (sub < 0 || sub == idx) run_rate<1, false, 128, 0>("cg1 SS tf32 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<1, true, 128, 0>("cg1 TS tf32 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<2, true, 128, 0>("cg2 TS tf32 M256 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<2, false, 128, 0>("cg2 SS tf32 M256 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<2, true, 256, 0>("cg2 TS tf32 M256 N256");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<1, true, 128, 1>("cg1 TS bf16 N128");
        ++idx;
        if (sub < 0 || sub == idx) run_rate<2, true, 128, 1>("cg2 TS bf16 M256 N128");
        ++idx;
    }
    if (which == 4) { if (sub == 1) run_diag<1>(); else run_diag<2>(); }
    if (which == 5) {
        const int M = 512, K = 128, N = 128;
        std::vector<float> a(M * K), b(N * K);
        for (auto& v : a) v = tf32_rn(frand());
        for (auto& v : b) v = tf32_rn(frand());
        float *dA, *dB, *dD;
        CK(cudaMalloc(&dA, M * K * 4)); CK(cudaMalloc(&dB, N * K * 4)); CK(cudaMalloc(&dD, M * N * 4));
        CK(cudaMemcpy(dA, a.data(), M * K * 4, cudaMemcpyHostToDevice)); CK(cudaMemcpy(dB, b.data(), N * K * 4, cudaMemcpyHostToDevice));
        CK(cudaMemset(dD, 0xFF, M * N * 4));
        CK(cudaFuncSetAttribute(mix_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 128 * 4));
        mix_probe<<<4, 128, 128 * 128 * 4>>>(dA, dB, dD);
        CK(cudaDeviceSynchronize());
        std::vector<float> d(M * N);
        CK(cudaMemcpy(d.data(), dD, M * N * 4, cudaMemcpyDeviceToHost));
        for (int blk = 0; blk < 4; ++blk) {
            double emax = 0;
            for (int m = blk * 128; m < blk * 128 + 128; ++m) for (int n = 0; n < N; ++n) {
                double se = 0;
                for (int k = 0; k < K; ++k) se += (double)a[m * K + k] * b[n * K + k];
                emax = fmax(emax, fabs(d[m * N + n] - se));
            }
            printf("T5 mixed cta_group kernel: rows %3d..%3d (%s) max err %.3e\n", blk * 128, blk * 128 + 127, blk < 2 ? "cta_group::2 pair" : "cta_group::1 CTA", emax);
        }
    }
    return 0;
}
