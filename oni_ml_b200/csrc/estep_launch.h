// estep_launch.h -- dispatch to the E-step instantiation for (G lanes per
// document, T topics per lane).  Each G lives in its own translation unit
// (estep_g*.cu) so the build runs them in parallel.
#pragma once
#include <cuda_runtime.h>

#include "estep.cuh"

namespace oni {

// Instantiated: G in {1,8,16,32} with T = 1..8; G in {2,4} with T = 5..8.
// Returns cudaErrorInvalidValue for anything else.
// longdoc = the launch holds only long documents (>= 1024 pairs): selects the latency-tuned
// build of the G = 32 kernel (T <= 4), otherwise ignored.
// exclusive_blocks > 0: launch that many blocks, each asking for all the shared memory of an SM,
// so that no other block shares the SM with them: one warp per scheduler, and the serial chain
// of word visits of a long document runs at the speed of a warp that has its scheduler to itself
// (main build: ~0.5 us per visit instead of ~1.3 us at full occupancy, at 3/4 of the throughput).
cudaError_t launch_estep(int G, int T, bool shift, const EstepParams &p, int sm_count,
                         cudaStream_t stream, bool longdoc = false, int exclusive_blocks = 0);

// The short-chain long-document kernel (estep_long.cuh): one document per warp, K <= 128, no
// max-shift.  exclusive_blocks as above; otherwise one block of 4 documents per SM at most.
cudaError_t launch_estep_long(int K, const EstepParams &p, int sm_count, cudaStream_t stream,
                              int exclusive_blocks = 0);

cudaError_t launch_estep_g1(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g2(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g4(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g8(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g16(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g32(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s, int exclusive_blocks);
cudaError_t launch_estep_g32_long(int T, bool shift, const EstepParams &p, int sm_count, cudaStream_t s,
                                  int exclusive_blocks);

}  // namespace oni
