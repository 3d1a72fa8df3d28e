// common.h -- shared host-side helpers of the engine (error reporting, RAII-free
// CUDA checks).  Internal; the public surface is include/oni_lda.h.
#pragma once
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/oni_lda.h"

#ifdef __cplusplus
extern "C" {
#endif
// Sets the thread-local message returned by oni_last_error(); returns `code`.
int oni_set_error(int code, const char *fmt, ...);
#ifdef __cplusplus
}
#endif

#ifdef __cplusplus
#include <vector>
// Internal (not exported through the header): shared by engine.cu and cli.cpp.
typedef void (*oni_iter_callback)(void *user, int iter, double likelihood, double converged);
// random_initialize_ss + lda_mle(model, ss, 0): logbeta [K][V]
void oni_internal_random_model(int K, int V, uint32_t seed, std::vector<double> &logbeta);
// run_em from a given initial model; on_iter is called after every EM iteration
int oni_internal_estimate(const oni_lda_corpus *corpus, int32_t K, double initial_alpha,
                          const oni_lda_settings *st, const double *logbeta0, int32_t ngpus,
                          oni_lda_result *res, oni_iter_callback on_iter, void *user);
#endif

#define ONI_CUDA(call)                                                                  \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess)                                                          \
            return oni_set_error(ONI_ERR_CUDA, "%s:%d: %s failed: %s", __FILE__, __LINE__, \
                                 #call, cudaGetErrorString(e_));                        \
    } while (0)

#define ONI_TRY(call)              \
    do {                           \
        int rc_ = (call);          \
        if (rc_ != ONI_OK) return rc_; \
    } while (0)
