// Launch unit of the tensor-core (tcgen05) path.  Hidden width padded to 128, up to 3 resident hidden matrices.
#include "jet_host.h"
#include "jet_tc_kernels.cuh"

namespace gpe {

bool tc_eligible(int d, int jet, int width, int L) {
    const bool lap = jet == JET_LAPLACIAN || (d == 1 && (jet == JET_DIAG || jet == JET_FULL));
    return lap && width > 64 && width <= 128 && L >= 2 && L - 1 <= tcp::kTcMaxMat;
}

size_t tc_wimg_bytes(int L) { return (size_t)(L - 1) * 2 * 2 * tcp::kImgFloats * sizeof(float); }

int tc_pack_wimg(const PtrArray& W, int L, int Wd, int transpose, float* wimg, cudaStream_t s) {
    tcp::pack_wimg_kernel<<<(L - 1) * 64, 256, 0, s>>>(W, L, Wd, transpose, wimg);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("pack_wimg_kernel: %s", cudaGetErrorString(e)); return -4; }
    return 0;
}

template <int D>
static int fwd_tc(const tcp::TcParams& P, const float* x, long long n, float* u, float* g, float* h, cudaStream_t s) {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    const size_t smem = tcp::tc_smem_bytes(D, P.n_hidden);
    cudaError_t e = cudaFuncSetAttribute(tcp::jet_fwd_tc_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_error("tc fwd smem attr (%zu B): %s", smem, cudaGetErrorString(e)); return -3; }
    const long long pairs = (n + 2 * tcp::kTcPts - 1) / (2 * tcp::kTcPts);
    long long clusters = sms / 2;
    if (pairs < clusters) clusters = pairs;
    tcp::jet_fwd_tc_kernel<D><<<(unsigned)(2 * clusters), tcp::kTcThreads, smem, s>>>(P, x, n, u, g, h, (int)pairs);
    e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("jet_fwd_tc_kernel launch: %s (grid %lld, smem %zu)", cudaGetErrorString(e), 2 * clusters, smem); return -4; }
    return 0;
}

int launch_fwd_tc(int d, int L, int act, const float* wimg, const float* small, const float* x, long long n, float* u,
                  float* g, float* h, cudaStream_t s) {
    tcp::TcParams P{L, act, wimg, small};
    return d == 1 ? fwd_tc<1>(P, x, n, u, g, h, s) : fwd_tc<2>(P, x, n, u, g, h, s);
}

}  // namespace gpe
