// estep_long.cu -- instantiations and launch of the long-document E-step (estep_long.cuh).
#include <algorithm>

#include "estep_launch.h"
#include "estep_long.cuh"

namespace oni {
namespace {

template <int T, int NSUM, int NW>
cudaError_t launch_nw(const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks)
{
    auto kernel = estep_long_kernel<T, NSUM, NW>;
    int dev = 0, optin = 0;
    cudaError_t e;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev)) != cudaSuccess) return e;
    const int need = (int)LongLayout<T, NW>::bytes;
    if (need > optin) return cudaErrorInvalidValue;
    // exclusive: ask for all the shared memory a block may have, so nothing shares the SM
    const int dyn = exclusive_blocks > 0 ? optin : need;
    if ((e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn)) != cudaSuccess) return e;
    const int want = (p.D + NW - 1) / NW;
    const int grid = exclusive_blocks > 0 ? exclusive_blocks : std::max(1, std::min(want, sm_count));
    kernel<<<grid, NW * 32, dyn, s>>>(p);
    return cudaGetLastError();
}

template <int T, int NSUM>
cudaError_t launch_long(const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks)
{
    return launch_nw<T, NSUM, LONG_NW>(p, sm_count, s, exclusive_blocks);
}

}  // namespace

cudaError_t launch_estep_long(int K, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks)
{
    if (K < 1 || K > 128 || !p.psi_tab || !p.exk_tab) return cudaErrorInvalidValue;
    if (K == 20) return launch_long<1, 20>(p, sm_count, s, exclusive_blocks);  // what the pipeline uses
    if (K <= 32) return launch_long<1, 32>(p, sm_count, s, exclusive_blocks);
    if (K <= 64) return launch_long<2, 32>(p, sm_count, s, exclusive_blocks);
    if (K <= 96) return launch_long<3, 32>(p, sm_count, s, exclusive_blocks);
    return launch_long<4, 32>(p, sm_count, s, exclusive_blocks);
}

}  // namespace oni
