// suffstats.cuh -- deterministic sufficient statistics, M-step and alpha Newton.
//
// Stands in for the scatter-add of doc_e_step (oni-lda-c/lda-estimate.c:48-55),
// the per-rank accumulation of likelihood / alpha_suffstats / num_docs
// (lda-estimate.c:201-208), lda_mle (lda-model.c:27-54) and opt_alpha
// (lda-alpha.c:42-68).
//
// The reference adds c_n*phi_nk into class_word[k][w] in document order.  Here
// the (document, word) pairs were sorted by word id ONCE per corpus; every EM
// iteration the rows c*phi[pair][:] of one word are summed by a fixed-shape
// reduction (no floating-point atomics), so two runs give the same bits:
//
//   level 1  the word's run of pairs is cut into chunks of <= SS_CHUNK pairs;
//            one warp sums a chunk: 32/G row slots, slot s takes rows
//            s, s+32/G, ... in order, then the slots are combined by a
//            butterfly whose shape depends only on G.
//   level 2  a word with more than one chunk has its chunk partials summed the
//            same way by one warp (partials are contiguous rows).
//
// class_total[k] is the fixed-order sum over words of class_word (the reference
// accumulates it alongside, lda-estimate.c:53); likelihood, alpha_suffstats and
// the word-visit count are fixed-order sums of the per-document values the
// E-step left behind.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "estep.cuh"

namespace oni {

constexpr int SS_CHUNK = 256;

struct SsItem {
    int64_t start;  // first position (level 1: in the word-sorted pair list; level 2: partial row)
    int32_t len;    // rows to add
    int32_t dst;    // >= 0: class_word row (word id); < 0: partial row ~dst
};

// Sum `len` rows into one.  LEVEL 1 gathers phi rows through perm and weights
// them by the pair's count; LEVEL 2 adds contiguous partial rows.
template <int G, int T, int LEVEL>
__global__ void __launch_bounds__(128) ss_rows_kernel(const SsItem *items, int32_t n_items,
                                                      const int32_t *perm, const int2 *tok,
                                                      const double *src, double *class_word,
                                                      double *partial, int32_t K, int32_t KP)
{
    const int lane = threadIdx.x & 31;
    const int item = (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5);
    if (item >= n_items) return;  // whole warps leave together
    constexpr int SLOTS = 32 / G;
    const int j = lane & (G - 1);
    const int slot = lane / G;
    const SsItem it = items[item];

    double acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = 0.0;

    for (int32_t i = slot; i < it.len; i += SLOTS) {
        size_t row;
        double wgt;
        if (LEVEL == 1) {
            const int32_t pr = __ldg(perm + it.start + i);
            row = (size_t)pr;
            wgt = (double)__ldg(&tok[pr].y);
        } else {
            row = (size_t)(it.start + i);
            wgt = 1.0;
        }
        const double *r = src + row * KP;
#pragma unroll
        for (int t = 0; t < T; t++) {
            const int k = j + G * t;
            if (k < K) {
                const double v = r[k];
                // lda-estimate.c:52  class_word += counts[n]*phi[n][k]
                acc[t] = (LEVEL == 1) ? (acc[t] + wgt * v) : (acc[t] + v);
            }
        }
    }
    // combine the row slots: lanes with the same j, butterfly over the slot bits
#pragma unroll
    for (int o = G; o < 32; o <<= 1) {
#pragma unroll
        for (int t = 0; t < T; t++) acc[t] += __shfl_xor_sync(0xffffffffu, acc[t], o);
    }
    if (slot == 0) {
        double *out = (it.dst >= 0) ? class_word + (size_t)it.dst * KP
                                    : partial + (size_t)(~it.dst) * KP;
#pragma unroll
        for (int t = 0; t < T; t++) {
            const int k = j + G * t;
            if (k < KP) out[k] = (k < K) ? acc[t] : 0.0;
        }
    }
}

// class_total[k] = sum_w class_word[w][k], one block, fixed order.
// Thread (slot, k): slot takes rows slot, slot+S, ...; then a shared-memory tree.
__global__ void __launch_bounds__(1024) class_total_kernel(const double *class_word, int32_t V,
                                                           int32_t K, int32_t KP,
                                                           double *class_total)
{
    extern __shared__ double sm[];  // [S][KP]
    const int S = blockDim.x / KP;  // row slots
    const int slot = threadIdx.x / KP, k = threadIdx.x % KP;
    double acc = 0.0;
    if (slot < S) {
        for (int w = slot; w < V; w += S) acc += class_word[(size_t)w * KP + k];
        sm[slot * KP + k] = acc;
    }
    __syncthreads();
    // S is a power of two by construction (host picks it)
    for (int h = S >> 1; h > 0; h >>= 1) {
        if (slot < h) sm[slot * KP + k] += sm[(slot + h) * KP + k];
        __syncthreads();
    }
    if (slot == 0) class_total[k] = (k < K) ? sm[k] : 0.0;
}

// Fixed-order sums over documents: likelihood, alpha_suffstats, word visits.
// Block b owns a fixed slice of documents; the last block to finish adds the
// block partials in index order.  tail = [alpha_ss, likelihood, num_docs, visits].
constexpr int DOC_RED_BLOCKS = 128;
constexpr int DOC_RED_THREADS = 256;

__global__ void __launch_bounds__(DOC_RED_THREADS)
doc_reduce_kernel(const double *doc_lik, const double *doc_ass, const int32_t *doc_iters,
                  const int64_t *doc_ptr, int32_t D, double *block_partial /* [B][3] */,
                  unsigned int *done_counter, double *tail)
{
    __shared__ double sm[3][DOC_RED_THREADS];
    __shared__ bool is_last;
    const int per = (D + gridDim.x - 1) / gridDim.x;
    const int lo = blockIdx.x * per, hi = min(D, lo + per);
    double a = 0.0, l = 0.0, v = 0.0;
    for (int d = lo + threadIdx.x; d < hi; d += blockDim.x) {
        l += doc_lik[d];
        a += doc_ass[d];
        v += (double)(doc_ptr[d + 1] - doc_ptr[d]) * (double)doc_iters[d];
    }
    sm[0][threadIdx.x] = a; sm[1][threadIdx.x] = l; sm[2][threadIdx.x] = v;
    __syncthreads();
    for (int h = blockDim.x >> 1; h > 0; h >>= 1) {
        if (threadIdx.x < h)
            for (int q = 0; q < 3; q++) sm[q][threadIdx.x] += sm[q][threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        for (int q = 0; q < 3; q++) block_partial[blockIdx.x * 3 + q] = sm[q][0];
        __threadfence();
        const unsigned prev = atomicAdd(done_counter, 1u);
        is_last = (prev == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last && threadIdx.x == 0) {
        __threadfence();
        double s[3] = {0.0, 0.0, 0.0};
        for (unsigned b = 0; b < gridDim.x; b++)
            for (int q = 0; q < 3; q++) s[q] += ((volatile double *)block_partial)[b * 3 + q];
        tail[0] = s[0];       // alpha_suffstats
        tail[1] = s[1];       // likelihood
        tail[2] = (double)D;  // num_docs
        tail[3] = s[2];       // word visits
        *done_counter = 0;    // ready for the next iteration
    }
}

// lda_mle, beta part (lda-model.c:31-44), written straight into the word-major
// table the E-step gathers from.
__global__ void __launch_bounds__(256) mle_beta_kernel(const double *class_word,
                                                       const double *class_total, int32_t V,
                                                       int32_t K, int32_t KP, double *logbeta_t)
{
    const size_t n = (size_t)V * KP;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n;
         i += (size_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % KP);
        double out = 0.0;
        if (k < K) {
            const double cw = class_word[i];
            out = (cw > 0.0) ? log(cw) - log(class_total[k]) : -100.0;
        }
        logbeta_t[i] = out;
    }
}

// Layout change between the reference's topic-major model [K][V] (lda.h:40-47, final.beta) and the
// word-major table the E-step gathers from [V][KP] (padding columns 0).  32x32 tiles through shared
// memory so that both sides are read and written in whole lines.
__global__ void __launch_bounds__(256) model_to_word_major_kernel(const double *kv, int32_t K, int32_t V,
                                                                  int32_t KP, double *vk)
{
    __shared__ double tile[32][33];
    const int w0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int k = k0 + r, w = w0 + threadIdx.x;
        tile[r][threadIdx.x] = (k < K && w < V) ? kv[(size_t)k * V + w] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int w = w0 + r, k = k0 + threadIdx.x;
        if (w < V && k < KP) vk[(size_t)w * KP + k] = tile[threadIdx.x][r];
    }
}
__global__ void __launch_bounds__(256) model_to_topic_major_kernel(const double *vk, int32_t K, int32_t V,
                                                                   int32_t KP, double *kv)
{
    __shared__ double tile[32][33];
    const int w0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int w = w0 + r, k = k0 + threadIdx.x;
        tile[r][threadIdx.x] = (w < V && k < K) ? vk[(size_t)w * KP + k] : 0.0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        const int k = k0 + r, w = w0 + threadIdx.x;
        if (k < K && w < V) kv[(size_t)k * V + w] = tile[threadIdx.x][r];
    }
}

// opt_alpha (lda-alpha.c:42-68): Newton on log(alpha) from 100, one thread.
// tail = [alpha_ss, likelihood, num_docs, visits].
__global__ void alpha_newton_kernel(const double *tail, int32_t K, double *alpha)
{
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const double ss = tail[0];
    const double D = (double)(int)tail[2];
    const double Kd = (double)K;
    double a, log_a, init_a = 100.0, df, d2f;
    int iter = 0;
    log_a = log(init_a);
    do {
        iter++;
        a = exp(log_a);
        if (a != a) {  // lda-alpha.c:53-59
            init_a = init_a * 10.0;
            a = init_a;
            log_a = log(a);
        }
        // lda-alpha.c:30-34; the reference's digamma/trigamma, literal form
        df = D * (Kd * dev_digamma(Kd * a) - Kd * dev_digamma(a)) + ss;
        d2f = D * (Kd * Kd * dev_trigamma(Kd * a) - Kd * dev_trigamma(a));
        log_a = log_a - df / (d2f * a + df);
    } while ((fabs(df) > 1e-5) && (iter < 1000));
    *alpha = exp(log_a);
}

}  // namespace oni
