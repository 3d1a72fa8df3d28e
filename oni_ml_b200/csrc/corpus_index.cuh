// corpus_index.cuh -- one-time, on-device indexing of a corpus shard.
//
// The reference re-parses model.dat into one malloc pair per document on every
// rank (oni-lda-c/lda-data.c:22-59); here the CSR arrays go to the GPU once and
// everything the EM iterations need is derived there:
//   tok        (word, count) packed per pair, word ids range-checked
//   doc_total  sum of counts per document (an int in the reference, lda.h:27)
//   order      documents by decreasing length (stable), the E-step's schedule
//   perm       pair ids sorted by word id (stable), the reduction's gather list
//   word_ptr   first position of every word in perm
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace oni {

__global__ void pack_tok_kernel(const int32_t *word, const int32_t *count, long long n, int V,
                                int2 *tok, unsigned long long *key, unsigned long long *idx,
                                int32_t *bad)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
        const int32_t w = word[i];
        if (w < 0 || w >= V) atomicExch(bad, 1);
        tok[i] = make_int2(w, count[i]);
        key[i] = (unsigned long long)(unsigned)w;  // sort key of the pair
        idx[i] = (unsigned long long)i;
    }
}

__global__ void doc_meta_kernel(const int64_t *ptr, const int32_t *count, int D, double *total,
                                unsigned long long *key, unsigned long long *idx)
{
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < D; d += gridDim.x * blockDim.x) {
        const int64_t a = ptr[d], b = ptr[d + 1];
        int32_t t = 0;  // int arithmetic as in lda-data.c:47
        for (int64_t i = a; i < b; i++) t += count[i];
        total[d] = (double)t;
        key[d] = 0xFFFFFFFFull - (unsigned long long)(b - a);  // ascending key = longest first
        idx[d] = (unsigned long long)d;
    }
}

__global__ void narrow_kernel(const unsigned long long *src, long long n, int32_t *dst)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x)
        dst[i] = (int32_t)src[i];
}

// word_ptr[w] = first position in the sorted key list whose word id is >= w
__global__ void word_ptr_kernel(const unsigned long long *sorted_word, long long n, int V,
                                int64_t *word_ptr)
{
    for (int w = blockIdx.x * blockDim.x + threadIdx.x; w <= V; w += gridDim.x * blockDim.x) {
        long long lo = 0, hi = n;
        while (lo < hi) {
            const long long mid = (lo + hi) >> 1;
            if (sorted_word[mid] < (unsigned long long)w) lo = mid + 1; else hi = mid;
        }
        word_ptr[w] = lo;
    }
}

}  // namespace oni
