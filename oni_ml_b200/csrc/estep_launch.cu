// estep_launch.cu -- (G, T) dispatch for the E-step kernel.
#include "estep_launch.h"

namespace oni {

cudaError_t launch_estep(int G, int T, bool shift, const EstepParams &p, int sm_count,
                         cudaStream_t stream, bool longdoc, int exclusive_blocks)
{
    if (longdoc && G == 32 && T <= 4) return launch_estep_g32_long(T, shift, p, sm_count, stream, exclusive_blocks);
    switch (G) {
        case 1: return launch_estep_g1(T, shift, p, sm_count, stream, exclusive_blocks);
        case 2: return launch_estep_g2(T, shift, p, sm_count, stream, exclusive_blocks);
        case 4: return launch_estep_g4(T, shift, p, sm_count, stream, exclusive_blocks);
        case 8: return launch_estep_g8(T, shift, p, sm_count, stream, exclusive_blocks);
        case 16: return launch_estep_g16(T, shift, p, sm_count, stream, exclusive_blocks);
        case 32: return launch_estep_g32(T, shift, p, sm_count, stream, exclusive_blocks);
        default: return cudaErrorInvalidValue;
    }
}

}  // namespace oni
