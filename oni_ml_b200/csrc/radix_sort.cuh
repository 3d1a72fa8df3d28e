// radix_sort.cuh -- stable LSD radix sort of (key, index) pairs, 8-bit digits.
//
// Used by the event scorer (rank survivors by score, ties by event index) and by
// the engine's one-time corpus indexing (documents by length, pairs by word id).
// Passes 0..7 take the digits of the 64-bit index (minor), passes 8..15 those of
// the 64-bit key (major); a caller whose input is already in index order starts
// at pass 8 and gets a stable sort by key alone.  Digit positions on which every
// element agrees are skipped.  No floating point, no atomics on the data path
// other than integer histogram counters: the result is deterministic.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <vector>

#include "common.h"

namespace oni {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ROUNDS = 16;                          // rounds of 32 keys per warp
constexpr int RS_TILE = RS_THREADS * RS_ROUNDS;        // keys per block
constexpr int RS_MAX_PASSES = 16;                      // 8 index digits + 8 key digits

__device__ __forceinline__ unsigned digit_of(const unsigned long long *key,
                                             const unsigned long long *idx, long long i, int pass)
{
    // passes 0..7 take the event index (minor), passes 8..15 the key (major)
    return (pass < 8) ? (unsigned)((idx[i] >> (8 * pass)) & 255u)
                      : (unsigned)((key[i] >> (8 * (pass - 8))) & 255u);
}

// totals[pass][digit] over all survivors, every pass at once.
static __global__ void __launch_bounds__(RS_THREADS)
radix_totals_kernel(const unsigned long long *key, const unsigned long long *idx, long long n,
                    unsigned long long *totals)
{
    __shared__ unsigned int sh[RS_MAX_PASSES][256];
    for (int i = threadIdx.x; i < RS_MAX_PASSES * 256; i += blockDim.x) (&sh[0][0])[i] = 0;
    __syncthreads();
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
         i += (long long)gridDim.x * blockDim.x) {
#pragma unroll
        for (int p = 0; p < RS_MAX_PASSES; p++) atomicAdd(&sh[p][digit_of(key, idx, i, p)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < RS_MAX_PASSES * 256; i += blockDim.x) {
        const unsigned v = (&sh[0][0])[i];
        if (v) atomicAdd(&totals[i], (unsigned long long)v);
    }
}

// counts[digit][tile] for one pass.
static __global__ void __launch_bounds__(RS_THREADS)
radix_count_kernel(const unsigned long long *key, const unsigned long long *idx, long long n,
                   int pass, int ntiles, unsigned long long *counts)
{
    __shared__ unsigned int sh[256];
    sh[threadIdx.x] = 0;
    __syncthreads();
    const long long lo = (long long)blockIdx.x * RS_TILE;
    const long long hi = min(n, lo + RS_TILE);
    for (long long i = lo + threadIdx.x; i < hi; i += RS_THREADS)
        atomicAdd(&sh[digit_of(key, idx, i, pass)], 1u);
    __syncthreads();
    counts[(size_t)threadIdx.x * ntiles + blockIdx.x] = sh[threadIdx.x];
}

// Exclusive scan of counts (digit-major [256][ntiles]) in place: block d scans row d,
// starting from the number of elements with a smaller digit (from the digit totals of
// this pass).  256 blocks.
static __global__ void __launch_bounds__(256)
radix_scan_kernel(unsigned long long *counts, int ntiles, const unsigned long long *pass_totals)
{
    __shared__ unsigned long long part[256];
    __shared__ unsigned long long base_sh;
    const int d = blockIdx.x;
    unsigned long long *row = counts + (size_t)d * ntiles;
    // digits below d
    unsigned long long b = (threadIdx.x < d) ? pass_totals[threadIdx.x] : 0ull;
    part[threadIdx.x] = b;
    __syncthreads();
    for (int h = 128; h > 0; h >>= 1) {
        if (threadIdx.x < h) part[threadIdx.x] += part[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) base_sh = part[0];
    __syncthreads();
    const int per = (ntiles + 255) / 256;
    const int lo = min(ntiles, (int)threadIdx.x * per), hi = min(ntiles, lo + per);
    unsigned long long s = 0;
    for (int i = lo; i < hi; i++) s += row[i];
    part[threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = base_sh;
        for (int i = 0; i < 256; i++) { const unsigned long long t = part[i]; part[i] = run; run += t; }
    }
    __syncthreads();
    unsigned long long run = part[threadIdx.x];
    for (int i = lo; i < hi; i++) { const unsigned long long t = row[i]; row[i] = run; run += t; }
}

// Stable scatter of one pass.  A warp owns RS_ROUNDS*32 consecutive keys and
// walks them in order; ranks inside a round come from match_any.
static __global__ void __launch_bounds__(RS_THREADS)
radix_scatter_kernel(const unsigned long long *kin, const unsigned long long *vin,
                     unsigned long long *kout, unsigned long long *vout, long long n, int pass,
                     int ntiles, const unsigned long long *offsets)
{
    __shared__ unsigned long long base[RS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&base[0][0])[i] = 0;
    __syncthreads();
    const long long wlo = (long long)blockIdx.x * RS_TILE + (long long)warp * (RS_ROUNDS * 32);
    // phase 1: digit counts of this warp's keys
    for (int r = 0; r < RS_ROUNDS; r++) {
        const long long i = wlo + r * 32 + lane;
        const bool ok = i < n;
        const unsigned d = ok ? digit_of(kin, vin, i, pass) : (0x10000u + lane);
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        if (ok && lane == __ffs(peers) - 1) base[warp][d] += __popc(peers);
        __syncwarp();
    }
    __syncthreads();
    // phase 2: first output slot of every (warp, digit)
    {
        const int d = threadIdx.x;  // 256 threads, 256 digits
        unsigned long long run = offsets[(size_t)d * ntiles + blockIdx.x];
        for (int w = 0; w < RS_WARPS; w++) {
            const unsigned long long t = base[w][d];
            base[w][d] = run;
            run += t;
        }
    }
    __syncthreads();
    // phase 3: same walk, now placing
    for (int r = 0; r < RS_ROUNDS; r++) {
        const long long i = wlo + r * 32 + lane;
        const bool ok = i < n;
        const unsigned d = ok ? digit_of(kin, vin, i, pass) : (0x10000u + lane);
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        unsigned long long pos = 0;
        if (ok) pos = base[warp][d] + __popc(peers & ((1u << lane) - 1u));
        __syncwarp();
        if (ok && lane == __ffs(peers) - 1) base[warp][d] += __popc(peers);
        __syncwarp();
        if (ok) {
            kout[pos] = kin[i];
            vout[pos] = vin[i];
        }
    }
}

// Sort n pairs.  key[0]/idx[0] hold the input; key[1]/idx[1] are same-sized scratch;
// counts needs 256*ceil(n/RS_TILE) words, totals RS_MAX_PASSES*256 words.
// On return *cur says which buffer holds the sorted pairs.  Synchronises the stream
// once (the digit totals come back to the host to pick the passes).
static inline int radix_sort_pairs(unsigned long long *const key[2], unsigned long long *const idx[2],
                                   unsigned long long *counts, unsigned long long *totals, long long n,
                                   int first_pass, int sm_count, cudaStream_t s, int *cur, int64_t *launches)
{
    *cur = 0;
    if (n <= 1) return ONI_OK;
    ONI_CUDA(cudaMemsetAsync(totals, 0, sizeof(unsigned long long) * RS_MAX_PASSES * 256, s));
    const int tblocks = (int)std::min<long long>((n + RS_THREADS - 1) / RS_THREADS, (long long)sm_count * 8);
    radix_totals_kernel<<<tblocks, RS_THREADS, 0, s>>>(key[0], idx[0], n, totals);
    std::vector<unsigned long long> h((size_t)RS_MAX_PASSES * 256);
    ONI_CUDA(cudaMemcpyAsync(h.data(), totals, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost, s));
    ONI_CUDA(cudaStreamSynchronize(s));
    if (launches) *launches += 1;
    const int ntiles = (int)((n + RS_TILE - 1) / RS_TILE);
    for (int pass = first_pass; pass < RS_MAX_PASSES; pass++) {
        bool trivial = false;
        for (int d = 0; d < 256; d++)
            if (h[(size_t)pass * 256 + d] == (unsigned long long)n) { trivial = true; break; }
        if (trivial) continue;  // every element has the same digit here
        const int in = *cur, out = *cur ^ 1;
        radix_count_kernel<<<ntiles, RS_THREADS, 0, s>>>(key[in], idx[in], n, pass, ntiles, counts);
        radix_scan_kernel<<<256, 256, 0, s>>>(counts, ntiles, totals + (size_t)pass * 256);
        radix_scatter_kernel<<<ntiles, RS_THREADS, 0, s>>>(key[in], idx[in], key[out], idx[out], n, pass, ntiles, counts);
        if (launches) *launches += 3;
        *cur = out;
    }
    ONI_CUDA(cudaGetLastError());
    return ONI_OK;
}

}  // namespace oni
