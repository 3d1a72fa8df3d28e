// scoring.cu -- post-LDA event scoring, threshold filter and ranking on sm_100a.
//
// Stands in for the scoring loops of flow_post_lda.scala:227-245 and
// dns_post_lda.scala:312-326, and for the table shaping of lda_post.py.
//
//   score kernel   one thread per event: theta row . p(w|z) row with one rounded
//                  multiply and one rounded add per topic, i = 0..K-1 in order
//                  (the JVM evaluates `s += a(i)*b(i)` without fusing), flow key
//                  = min(src, dst); events with key < tol are appended, warp-
//                  aggregated, to the survivor list (order irrelevant: see sort)
//   radix sort     stable LSD radix sort, 8-bit digits, over the composite
//                  (key bits, event index): index digits first, then key digits,
//                  so equal scores rank by ascending event index and the result
//                  does not depend on the order survivors were appended in.
//                  Digit positions where every survivor agrees are skipped.
//
// Scratch: survivor keys/indices (two buffers each, ping-pong), per-tile digit
// counts and their scan.  All of it is global memory owned by the scoring
// context and sized to the survivor count.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "common.h"
#include "radix_sort.cuh"

using namespace oni;

namespace {

// ------------------------------------------------------------------- score math
// sum_{i<K} a[i]*b[i], strictly sequential, no contraction.
__device__ __forceinline__ double dot_rows(const double *__restrict__ a,
                                           const double *__restrict__ b, int K, bool vec2)
{
    double acc = 0.0;
    if (vec2) {
        const double2 *a2 = reinterpret_cast<const double2 *>(a);
        const double2 *b2 = reinterpret_cast<const double2 *>(b);
        for (int i = 0; i < K / 2; i++) {
            const double2 x = __ldg(a2 + i), y = __ldg(b2 + i);
            acc = __dadd_rn(acc, __dmul_rn(x.x, y.x));
            acc = __dadd_rn(acc, __dmul_rn(x.y, y.y));
        }
    } else {
        for (int i = 0; i < K; i++) acc = __dadd_rn(acc, __dmul_rn(__ldg(a + i), __ldg(b + i)));
    }
    return acc;
}

// One side of an event.  A missing document or word scores with the constant
// vector (flow_post_lda.scala:228-231, dns_post_lda.scala:313-314).
__device__ __forceinline__ double side_score(const double *theta, const double *pwz, int K,
                                             int32_t doc, int32_t word, double dflt, bool vec2)
{
    if (doc >= 0 && word >= 0)
        return dot_rows(theta + (size_t)doc * K, pwz + (size_t)word * K, K, vec2);
    double acc = 0.0;
    const double *row = (doc >= 0) ? theta + (size_t)doc * K : (word >= 0 ? pwz + (size_t)word * K : nullptr);
    for (int i = 0; i < K; i++) {
        const double other = row ? __ldg(row + i) : dflt;
        // keep the operand order theta*p of the reference
        const double prod = (doc >= 0) ? __dmul_rn(other, dflt) : __dmul_rn(dflt, other);
        acc = __dadd_rn(acc, prod);
    }
    return acc;
}

__device__ __forceinline__ unsigned long long key_bits(double v)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(v);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);  // total order of doubles as unsigned
}
__host__ __device__ inline double key_value(unsigned long long u)
{
    u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)u);
#else
    double d;
    memcpy(&d, &u, sizeof d);
    return d;
#endif
}

template <bool FLOW>
__global__ void __launch_bounds__(256)
score_filter_kernel(int K, int D, int V, const double *__restrict__ theta, const double *__restrict__ pwz,
                    long long n, const int32_t *__restrict__ a_doc, const int32_t *__restrict__ b_doc,
                    const int32_t *__restrict__ a_word, const int32_t *__restrict__ b_word,
                    double dflt, double tol, long long idx_base, unsigned long long *surv_key,
                    unsigned long long *surv_idx, unsigned long long cap,
                    unsigned long long *counter)
{
    const bool vec2 = (K % 2) == 0;
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    // whole warps iterate together so the ballot below is over a full warp
    for (long long base = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride) {
        const long long e = base + lane;
        bool keep = false;
        double key = 0.0;
        if (e < n) {
            // an index beyond its table is the caller's error: flagged (counter[1]) and read as a miss
            int32_t ad = a_doc[e], aw = a_word[e], bd = FLOW ? b_doc[e] : -1, bw = FLOW ? b_word[e] : -1;
            if (ad >= D || aw >= V || bd >= D || bw >= V) {
                counter[1] = 1;
                ad = ad >= D ? -1 : ad; aw = aw >= V ? -1 : aw; bd = bd >= D ? -1 : bd; bw = bw >= V ? -1 : bw;
            }
            const double sa = side_score(theta, pwz, K, ad, aw, dflt, vec2);
            if (FLOW) {
                const double sb = side_score(theta, pwz, K, bd, bw, dflt, vec2);
                key = (sb < sa) ? sb : sa;  // flow_post_lda.scala:238
            } else {
                key = sa;
            }
            keep = key < tol;  // strict, flow_post_lda.scala:245
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m) {
            unsigned long long pos = 0;
            if (lane == __ffs(m) - 1) pos = atomicAdd(counter, (unsigned long long)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, __ffs(m) - 1);
            if (keep) {
                pos += __popc(m & ((1u << lane) - 1u));
                if (pos < cap) {
                    surv_key[pos] = key_bits(key);
                    surv_idx[pos] = (unsigned long long)(idx_base + e);
                }
            }
        }
    }
}

// ---- paired variant (K = 2*CH, CH 16-byte pieces per row) -------------------------------
// The kernel above is bound by the L1 data pipe, not by HBM: each lane of a 16-byte load
// touches a different row, so a warp load is 32 half-used 32-byte sectors (ncu: lsu
// wavefronts 96 % of peak, lts 78 %, dram 31 %).  Here lanes 2m and 2m+1 share their two
// events: for every row, the even lane loads the even 16-byte pieces and the odd lane the
// odd ones, so the pair reads whole 32-byte sectors, and one shuffle per piece hands the
// pieces to the lane that owns the event.  Half the sectors per load, every sector requested
// once; the sum itself is unchanged (one lane, topics in order, unfused).
template <int CH>
__device__ __forceinline__ double dot_paired(const double *theta, const double *pwz, int32_t my_doc,
                                             int32_t my_word, double dflt, int lane)
{
    constexpr int HALF = (CH + 1) / 2;  // pieces per lane and row (the odd lane may hold one less)
    const int q = lane & 1;
    const int32_t pal_doc = __shfl_xor_sync(0xffffffffu, my_doc, 1);
    const int32_t pal_word = __shfl_xor_sync(0xffffffffu, my_word, 1);
    // rows of the even lane's event ("r0") and of the odd lane's event ("r1")
    const int32_t d0 = q ? pal_doc : my_doc, w0 = q ? pal_word : my_word;
    const int32_t d1 = q ? my_doc : pal_doc, w1 = q ? my_word : pal_word;
    const double2 fill = make_double2(dflt, dflt);
    double2 t0[HALF], p0[HALF], t1[HALF], p1[HALF];
#pragma unroll
    for (int i = 0; i < HALF; i++) {
        const int piece = 2 * i + q;  // this lane's pieces: q, q+2, ...
        const bool in = piece < CH;
        t0[i] = (in && d0 >= 0) ? __ldg(reinterpret_cast<const double2 *>(theta + (size_t)d0 * (2 * CH)) + piece) : fill;
        p0[i] = (in && w0 >= 0) ? __ldg(reinterpret_cast<const double2 *>(pwz + (size_t)w0 * (2 * CH)) + piece) : fill;
        t1[i] = (in && d1 >= 0) ? __ldg(reinterpret_cast<const double2 *>(theta + (size_t)d1 * (2 * CH)) + piece) : fill;
        p1[i] = (in && w1 >= 0) ? __ldg(reinterpret_cast<const double2 *>(pwz + (size_t)w1 * (2 * CH)) + piece) : fill;
    }
    // hand over and accumulate, piece pair by piece pair.  The lane that loaded a piece also
    // forms its two products (__dmul_rn: the same rounding wherever it is done), so only the
    // products travel: the even lane gives away piece 2i of the odd lane's event and receives
    // piece 2i+1 of its own; the odd lane the other way round.  The owner then adds the
    // products of pieces 2i and 2i+1 in topic order, unfused.
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < HALF; i++) {
        const double2 ta = q ? t0[i] : t1[i], pa = q ? p0[i] : p1[i];  // the partner's piece
        const double2 tb = q ? t1[i] : t0[i], pb = q ? p1[i] : p0[i];  // my own piece
        const double gx = __dmul_rn(ta.x, pa.x), gy = __dmul_rn(ta.y, pa.y);
        const double mx = __dmul_rn(tb.x, pb.x), my = __dmul_rn(tb.y, pb.y);
        const double rx = __shfl_xor_sync(0xffffffffu, gx, 1);
        const double ry = __shfl_xor_sync(0xffffffffu, gy, 1);
        // piece 2i then piece 2i+1: (own, received) on the even lane, (received, own) on the odd
        acc = __dadd_rn(acc, q ? rx : mx);
        acc = __dadd_rn(acc, q ? ry : my);
        if (2 * i + 1 < CH) {
            acc = __dadd_rn(acc, q ? mx : rx);
            acc = __dadd_rn(acc, q ? my : ry);
        }
    }
    return acc;
}

#ifndef ONI_SCORE_MINB
#define ONI_SCORE_MINB 4
#endif
template <bool FLOW, int CH>
__global__ void __launch_bounds__(128, ONI_SCORE_MINB)
score_filter_paired_kernel(int D, int V, const double *__restrict__ theta, const double *__restrict__ pwz, long long n,
                           const int32_t *__restrict__ a_doc, const int32_t *__restrict__ b_doc,
                           const int32_t *__restrict__ a_word, const int32_t *__restrict__ b_word,
                           double dflt, double tol, long long idx_base, unsigned long long *surv_key,
                           unsigned long long *surv_idx, unsigned long long cap,
                           unsigned long long *counter)
{
    const int lane = threadIdx.x & 31;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long base = (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); base < n; base += stride) {
        const long long e = base + lane;
        const bool in = e < n;
        // lanes past the end take part in the exchange with "missing" rows; an index beyond its
        // table is the caller's error: flagged (counter[1]) and read as a miss
        int32_t ad = in ? a_doc[e] : -1, aw = in ? a_word[e] : -1;
        int32_t bd = (FLOW && in) ? b_doc[e] : -1, bw = (FLOW && in) ? b_word[e] : -1;
        if (ad >= D || aw >= V || bd >= D || bw >= V) {
            counter[1] = 1;
            ad = ad >= D ? -1 : ad; aw = aw >= V ? -1 : aw; bd = bd >= D ? -1 : bd; bw = bw >= V ? -1 : bw;
        }
        const double sa = dot_paired<CH>(theta, pwz, ad, aw, dflt, lane);
        double key = sa;
        if (FLOW) {
            // (the second side's 2*CH loads must stay below the first side's sum: all 4*CH in
            //  flight at once need 192 registers and leave 8 warps per SM; the compiler hoists
            //  them over any barrier, so their addresses are made to depend on sa)
            const int order = (sa != sa) ? 1 : 0;  // always 0: makes the addresses below depend on sa
            const double sb = dot_paired<CH>(theta, pwz, bd >= 0 ? bd + order : -1, bw >= 0 ? bw + order : -1, dflt, lane);
            key = (sb < sa) ? sb : sa;  // flow_post_lda.scala:238
        }
        const bool keep = in && (key < tol);  // strict, flow_post_lda.scala:245
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        if (m) {
            unsigned long long pos = 0;
            if (lane == __ffs(m) - 1) pos = atomicAdd(counter, (unsigned long long)__popc(m));
            pos = __shfl_sync(0xffffffffu, pos, __ffs(m) - 1);
            if (keep) {
                pos += __popc(m & ((1u << lane) - 1u));
                if (pos < cap) {
                    surv_key[pos] = key_bits(key);
                    surv_idx[pos] = (unsigned long long)(idx_base + e);
                }
            }
        }
    }
}

// Two staged variants of this kernel were measured and dropped (profiles/r1_score.txt): rows
// gathered into shared memory with cp.async (LDGSTS, consecutive lanes on consecutive 16-byte
// pieces, double-buffered) ran at 37-40 ms per 250 M events, and with per-lane 160-byte TMA
// bulk copies (UBLKCP + mbarrier) at 49 ms, against 28 ms for the direct gather above: the
// first is short of warps at 45 KB of shared memory each, the second is bound by the TMA
// request rate (one 160-byte request per ~14 cycles per SM).
// src/dst score of the ranked survivors (what the Scala stage appends to a row).
template <bool FLOW>
__global__ void rescored_kernel(int K, const double *theta, const double *pwz, long long m,
                                const unsigned long long *idx, long long idx_base,
                                const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                                const int32_t *b_word, double dflt, double *out_a, double *out_b)
{
    const bool vec2 = (K % 2) == 0;
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < m;
         r += (long long)gridDim.x * blockDim.x) {
        const long long e = idx ? (long long)idx[r] - idx_base : r;  // (no list: the events ARE the ranked ones)
        out_a[r] = side_score(theta, pwz, K, a_doc[e], a_word[e], dflt, vec2);
        if (FLOW) out_b[r] = side_score(theta, pwz, K, b_doc[e], b_word[e], dflt, vec2);
    }
}

// ---------------------------------------------------------- lda_post.py on device
// theta = gamma / sum(gamma), python sum(): left to right (lda_post.py:44-54)
__global__ void theta_kernel(int D, int K, const double *gamma, double *theta)
{
    for (int d = blockIdx.x * blockDim.x + threadIdx.x; d < D; d += gridDim.x * blockDim.x) {
        const double *g = gamma + (size_t)d * K;
        double total = 0.0;
        for (int k = 0; k < K; k++) total = __dadd_rn(total, g[k]);
        for (int k = 0; k < K; k++) theta[(size_t)d * K + k] = (total > 0.0) ? g[k] / total : 0.0;
    }
}
// pwz[w][k] = exp(beta[k][w]) / sum_w exp(beta[k][w]), sum left to right (lda_post.py:88-96).
// One block per topic; thread 0 forms the sequential total from the staged exps.
__global__ void pwz_kernel(int K, int V, const double *beta, double *raw, double *pwz)
{
    const int k = blockIdx.x;
    __shared__ double total_sh;
    for (int w = threadIdx.x; w < V; w += blockDim.x) raw[(size_t)k * V + w] = exp(beta[(size_t)k * V + w]);
    __syncthreads();
    if (threadIdx.x == 0) {
        double total = 0.0;
        for (int w = 0; w < V; w++) total = __dadd_rn(total, raw[(size_t)k * V + w]);
        total_sh = total;
    }
    __syncthreads();
    const double total = total_sh;
    for (int w = threadIdx.x; w < V; w += blockDim.x) pwz[(size_t)w * K + k] = raw[(size_t)k * V + w] / total;
}

struct DevBuf {
    void *p = nullptr;
    size_t bytes = 0;
    int reserve(size_t need)
    {
        if (need <= bytes) return ONI_OK;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        ONI_CUDA(cudaMalloc(&p, need ? need : 1));
        bytes = need;
        return ONI_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};

}  // namespace

// ------------------------------------------------------------- scoring context
struct oni_score_ctx {
    int device = 0, K = 0, D = 0, V = 0, sm_count = 0;
    cudaStream_t stream = nullptr;
    double *d_theta = nullptr, *d_pwz = nullptr;
    unsigned long long *d_counter = nullptr, *d_totals = nullptr;
    DevBuf key[2], idx[2], counts, ev[4], out_a, out_b;
    // events from host arrays: three sets of chunk buffers, uploads on a stream of their own so
    // that chunk i+1 travels while chunk i is scored
    cudaStream_t copy_stream = nullptr;
    DevBuf up[2][4];
    cudaEvent_t up_done[2] = {};
    unsigned long long cap = 0;     // survivor capacity (entries)
    unsigned long long n_surv = 0;  // survivors of the last run
    int cur = 0;                    // which ping-pong buffer holds the ranked list
    int64_t launches = 0;
    float ms_score = 0, ms_sort = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
};

extern "C" void oni_score_destroy(oni_score_ctx *c);

extern "C" int oni_score_create(int32_t device, int32_t K, int32_t D, int32_t V, const double *theta,
                                const double *pwz, oni_score_ctx **out)
{
    if (!out || !theta || !pwz || K < 1 || D < 0 || V < 0)
        return oni_set_error(ONI_ERR_ARG, "score_create: bad argument");
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    if (device < 0 || device >= ndev) return oni_set_error(ONI_ERR_ARG, "device %d of %d", device, ndev);
    oni_score_ctx *c = new oni_score_ctx();
    c->device = device; c->K = K; c->D = D; c->V = V;
    auto build = [&]() -> int {
        ONI_CUDA(cudaSetDevice(device));
        cudaDeviceProp prop;
        ONI_CUDA(cudaGetDeviceProperties(&prop, device));
        c->sm_count = prop.multiProcessorCount;
        ONI_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        ONI_CUDA(cudaEventCreate(&c->e0)); ONI_CUDA(cudaEventCreate(&c->e1)); ONI_CUDA(cudaEventCreate(&c->e2));
        ONI_CUDA(cudaMalloc((void **)&c->d_theta, sizeof(double) * std::max<size_t>(1, (size_t)D * K)));
        ONI_CUDA(cudaMalloc((void **)&c->d_pwz, sizeof(double) * std::max<size_t>(1, (size_t)V * K)));
        ONI_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        for (int b = 0; b < 2; b++) ONI_CUDA(cudaEventCreateWithFlags(&c->up_done[b], cudaEventDisableTiming));
        ONI_CUDA(cudaMalloc((void **)&c->d_counter, 2 * sizeof(unsigned long long)));  // [0] survivors, [1] bad-index flag
        ONI_CUDA(cudaMalloc((void **)&c->d_totals, sizeof(unsigned long long) * RS_MAX_PASSES * 256));
        ONI_CUDA(cudaMemsetAsync(c->d_counter, 0, 2 * sizeof(unsigned long long), c->stream));
        ONI_CUDA(cudaMemcpyAsync(c->d_theta, theta, sizeof(double) * (size_t)D * K, cudaMemcpyHostToDevice, c->stream));
        ONI_CUDA(cudaMemcpyAsync(c->d_pwz, pwz, sizeof(double) * (size_t)V * K, cudaMemcpyHostToDevice, c->stream));
        ONI_CUDA(cudaStreamSynchronize(c->stream));  // (the caller's arrays may go away)
        return ONI_OK;
    };
    const int rc = build();
    if (rc != ONI_OK) { oni_score_destroy(c); return rc; }  // nothing half-built survives a failure
    *out = c;
    return ONI_OK;
}

extern "C" void oni_score_destroy(oni_score_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaFree(c->d_theta); cudaFree(c->d_pwz); cudaFree(c->d_counter); cudaFree(c->d_totals);
    for (auto &b : c->key) b.release();
    for (auto &b : c->idx) b.release();
    for (auto &b : c->ev) b.release();
    c->counts.release(); c->out_a.release(); c->out_b.release();
    for (auto &set : c->up) for (auto &b : set) b.release();
    for (auto &e : c->up_done) if (e) cudaEventDestroy(e);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->e0) cudaEventDestroy(c->e0);
    if (c->e1) cudaEventDestroy(c->e1);
    if (c->e2) cudaEventDestroy(c->e2);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

static int reserve_survivors(oni_score_ctx *c, unsigned long long cap)
{
    if (cap <= c->cap) return ONI_OK;
    // growing keeps what is already there (chunked runs append across calls)
    for (int b = 0; b < 2; b++) {
        DevBuf nk, ni;
        ONI_TRY(nk.reserve(cap * 8));
        { const int rc = ni.reserve(cap * 8); if (rc != ONI_OK) { nk.release(); return rc; } }
        if (c->cap && b == c->cur) {
            ONI_CUDA(cudaMemcpyAsync(nk.p, c->key[b].p, c->cap * 8, cudaMemcpyDeviceToDevice, c->stream));
            ONI_CUDA(cudaMemcpyAsync(ni.p, c->idx[b].p, c->cap * 8, cudaMemcpyDeviceToDevice, c->stream));
            ONI_CUDA(cudaStreamSynchronize(c->stream));
        }
        c->key[b].release(); c->idx[b].release();
        c->key[b] = nk; c->idx[b] = ni;
    }
    c->cap = cap;
    return ONI_OK;
}

// Start a scoring run (clears the survivor list).
extern "C" int oni_score_begin(oni_score_ctx *c)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "score_begin: null context");
    ONI_CUDA(cudaSetDevice(c->device));
    ONI_CUDA(cudaMemsetAsync(c->d_counter, 0, 2 * sizeof(unsigned long long), c->stream));
    c->n_surv = 0; c->cur = 0; c->ms_score = c->ms_sort = 0;
    return ONI_OK;
}

// Score n events whose indices are on the device; survivors accumulate.
// idx_base = global index of the chunk's first event.
extern "C" int oni_score_events_device(oni_score_ctx *c, int32_t is_flow, int64_t n, int64_t idx_base,
                                       const int32_t *a_doc, const int32_t *b_doc,
                                       const int32_t *a_word, const int32_t *b_word, double tol)
{
    if (!c || n < 0 || (n > 0 && (!a_doc || !a_word || (is_flow && (!b_doc || !b_word)))))
        return oni_set_error(ONI_ERR_ARG, "score_events: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    if (n == 0) return ONI_OK;
    const double dflt = is_flow ? 0.05 : 0.1;
    const int blocks = (int)std::min<long long>((n + 255) / 256, (long long)c->sm_count * 16);
    if (c->cap == 0) ONI_TRY(reserve_survivors(c, std::max<unsigned long long>(1u << 20, (unsigned long long)n / 64)));
    for (int attempt = 0; attempt < 2; attempt++) {
        ONI_CUDA(cudaEventRecord(c->e0, c->stream));
        unsigned long long *sk = (unsigned long long *)c->key[0].p, *si = (unsigned long long *)c->idx[0].p;
        // K = 20 is what lda_post.py and both Scala stages hard-code (lda_post.py:115,
        // flow_post_lda.scala:234): that case gets the lane-paired gather
        const bool paired = (c->K == 20) && !getenv("ONI_SCORE_SIMPLE");
        const int pblocks = (int)std::min<long long>((n + 127) / 128, (long long)c->sm_count * 32);
        if (paired && is_flow)
            score_filter_paired_kernel<true, 10><<<pblocks, 128, 0, c->stream>>>(
                c->D, c->V, c->d_theta, c->d_pwz, n, a_doc, b_doc, a_word, b_word, dflt, tol, idx_base, sk, si, c->cap, c->d_counter);
        else if (paired)
            score_filter_paired_kernel<false, 10><<<pblocks, 128, 0, c->stream>>>(
                c->D, c->V, c->d_theta, c->d_pwz, n, a_doc, nullptr, a_word, nullptr, dflt, tol, idx_base, sk, si, c->cap, c->d_counter);
        else if (is_flow)
            score_filter_kernel<true><<<blocks, 256, 0, c->stream>>>(
                c->K, c->D, c->V, c->d_theta, c->d_pwz, n, a_doc, b_doc, a_word, b_word, dflt, tol, idx_base, sk, si, c->cap, c->d_counter);
        else
            score_filter_kernel<false><<<blocks, 256, 0, c->stream>>>(
                c->K, c->D, c->V, c->d_theta, c->d_pwz, n, a_doc, nullptr, a_word, nullptr, dflt, tol, idx_base, sk, si, c->cap, c->d_counter);
        c->launches++;
        ONI_CUDA(cudaEventRecord(c->e1, c->stream));
        unsigned long long two[2] = {0, 0};
        ONI_CUDA(cudaMemcpyAsync(two, c->d_counter, sizeof two, cudaMemcpyDeviceToHost, c->stream));
        ONI_CUDA(cudaStreamSynchronize(c->stream));
        const unsigned long long total = two[0];
        if (two[1])
            return oni_set_error(ONI_ERR_ARG, "score: an event's document index is >= %d or its word index >= %d "
                                 "(events %lld..%lld)", c->D, c->V, (long long)idx_base, (long long)(idx_base + n - 1));
        float ms = 0;
        cudaEventElapsedTime(&ms, c->e0, c->e1);
        if (total <= c->cap) { c->n_surv = total; c->ms_score += ms; return ONI_OK; }
        // the list overflowed: grow to what was counted, rewind this chunk, run it again
        const unsigned long long before = c->n_surv;
        ONI_TRY(reserve_survivors(c, total + total / 8));
        ONI_CUDA(cudaMemcpyAsync(c->d_counter, &before, sizeof before, cudaMemcpyHostToDevice, c->stream));
        ONI_CUDA(cudaStreamSynchronize(c->stream));
    }
    return oni_set_error(ONI_ERR_CUDA, "score_events: survivor list overflowed twice");
}

// Rank the accumulated survivors (ascending key, ties by event index).
extern "C" int oni_score_rank(oni_score_ctx *c, int64_t *n_survivors)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "score_rank: null context");
    ONI_CUDA(cudaSetDevice(c->device));
    const long long n = (long long)c->n_surv;
    if (n_survivors) *n_survivors = n;
    c->cur = 0;
    if (n <= 1) return ONI_OK;
    cudaStream_t s = c->stream;
    ONI_CUDA(cudaEventRecord(c->e1, s));
    const int ntiles = (int)((n + RS_TILE - 1) / RS_TILE);
    ONI_TRY(c->counts.reserve(sizeof(unsigned long long) * 256 * (size_t)ntiles));
    unsigned long long *key[2] = {(unsigned long long *)c->key[0].p, (unsigned long long *)c->key[1].p};
    unsigned long long *idx[2] = {(unsigned long long *)c->idx[0].p, (unsigned long long *)c->idx[1].p};
    ONI_TRY(radix_sort_pairs(key, idx, (unsigned long long *)c->counts.p, c->d_totals, n, 0, c->sm_count, s,
                             &c->cur, &c->launches));
    ONI_CUDA(cudaEventRecord(c->e2, s));
    ONI_CUDA(cudaStreamSynchronize(s));
    ONI_CUDA(cudaGetLastError());
    float ms = 0;
    cudaEventElapsedTime(&ms, c->e1, c->e2);
    c->ms_sort = ms;
    return ONI_OK;
}

// Copy the first `cap` ranked survivors to the host.
extern "C" int oni_score_fetch(oni_score_ctx *c, int64_t cap, int64_t *out_idx, double *out_key)
{
    if (!c || cap < 0) return oni_set_error(ONI_ERR_ARG, "score_fetch: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    const size_t m = (size_t)std::min<unsigned long long>((unsigned long long)cap, c->n_surv);
    if (m == 0) return ONI_OK;
    std::vector<unsigned long long> k(m);
    if (out_idx)
        ONI_CUDA(cudaMemcpy(out_idx, c->idx[c->cur].p, sizeof(int64_t) * m, cudaMemcpyDeviceToHost));
    if (out_key) {
        ONI_CUDA(cudaMemcpy(k.data(), c->key[c->cur].p, sizeof(unsigned long long) * m, cudaMemcpyDeviceToHost));
        for (size_t i = 0; i < m; i++) out_key[i] = key_value(k[i]);
    }
    return ONI_OK;
}

// src/dst scores of the first `cap` ranked survivors; the event index arrays are
// the device arrays of the chunk that starts at idx_base and holds them all.
extern "C" int oni_score_fetch_sides(oni_score_ctx *c, int32_t is_flow, int64_t cap, int64_t idx_base,
                                     const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                                     const int32_t *b_word, double *out_a, double *out_b)
{
    if (!c || cap < 0) return oni_set_error(ONI_ERR_ARG, "score_fetch_sides: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    const long long m = (long long)std::min<unsigned long long>((unsigned long long)cap, c->n_surv);
    if (m == 0) return ONI_OK;
    ONI_TRY(c->out_a.reserve(sizeof(double) * (size_t)m));
    ONI_TRY(c->out_b.reserve(sizeof(double) * (size_t)m));
    const double dflt = is_flow ? 0.05 : 0.1;
    const int blocks = (int)std::min<long long>((m + 255) / 256, 4096);
    if (is_flow)
        rescored_kernel<true><<<blocks, 256, 0, c->stream>>>(c->K, c->d_theta, c->d_pwz, m,
            (unsigned long long *)c->idx[c->cur].p, idx_base, a_doc, b_doc, a_word, b_word, dflt,
            (double *)c->out_a.p, (double *)c->out_b.p);
    else
        rescored_kernel<false><<<blocks, 256, 0, c->stream>>>(c->K, c->d_theta, c->d_pwz, m,
            (unsigned long long *)c->idx[c->cur].p, idx_base, a_doc, nullptr, a_word, nullptr, dflt,
            (double *)c->out_a.p, (double *)c->out_b.p);
    c->launches++;
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    ONI_CUDA(cudaGetLastError());
    if (out_a) ONI_CUDA(cudaMemcpy(out_a, c->out_a.p, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost));
    if (out_b && is_flow) ONI_CUDA(cudaMemcpy(out_b, c->out_b.p, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost));
    return ONI_OK;
}

extern "C" int oni_score_get_stats(oni_score_ctx *c, double *ms_score, double *ms_sort, int64_t *launches)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "score_get_stats: null context");
    if (ms_score) *ms_score = c->ms_score;
    if (ms_sort) *ms_sort = c->ms_sort;
    if (launches) *launches = c->launches;
    return ONI_OK;
}

// ------------------------------------------------------ host-buffer entry points
// Score n events whose indices are in HOST arrays; survivors accumulate as with
// oni_score_events_device.  The events travel in chunks (2^24 by default, ONI_SCORE_CHUNK) through
// two sets of device buffers on a copy stream of their own, so that chunk i+1 is on the bus while
// chunk i is scored: with pinned host arrays a day of events costs its PCIe time (16 bytes per
// flow event) and little more.  Pageable arrays work, at the speed of staged copies.
extern "C" int oni_score_events_host(oni_score_ctx *c, int32_t is_flow, int64_t n, int64_t idx_base,
                                     const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                                     const int32_t *b_word, double tol)
{
    if (!c || n < 0 || (n > 0 && (!a_doc || !a_word || (is_flow && (!b_doc || !b_word)))))
        return oni_set_error(ONI_ERR_ARG, "score_events_host: bad argument");
    if (n == 0) return ONI_OK;
    ONI_CUDA(cudaSetDevice(c->device));
    const int nsrc = is_flow ? 4 : 2;
    const int32_t *src[4] = {a_doc, a_word, b_doc, b_word};
    const char *env = getenv("ONI_SCORE_CHUNK");
    int64_t chunk = env ? atoll(env) : ((int64_t)1 << 24);
    if (chunk < 1) chunk = 1;
    chunk = std::min(chunk, n);
    const int64_t nchunks = (n + chunk - 1) / chunk;
    for (int b = 0; b < (nchunks > 1 ? 2 : 1); b++)
        for (int q = 0; q < nsrc; q++) ONI_TRY(c->up[b][q].reserve(sizeof(int32_t) * (size_t)chunk));
    // (the buffers may still be read by a kernel of an earlier call on this context)
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    auto upload = [&](int64_t i) -> int {
        const int b = (int)(i & 1);
        const int64_t lo = i * chunk, m = std::min(chunk, n - lo);
        for (int q = 0; q < nsrc; q++)
            ONI_CUDA(cudaMemcpyAsync(c->up[b][q].p, src[q] + lo, sizeof(int32_t) * (size_t)m, cudaMemcpyHostToDevice,
                                     c->copy_stream));
        ONI_CUDA(cudaEventRecord(c->up_done[b], c->copy_stream));
        return ONI_OK;
    };
    ONI_TRY(upload(0));
    for (int64_t i = 0; i < nchunks; i++) {
        // chunk i+1 goes into the buffers of chunk i-1, whose kernel has finished (every chunk ends
        // with a read-back of the survivor count)
        if (i + 1 < nchunks) ONI_TRY(upload(i + 1));
        const int b = (int)(i & 1);
        const int64_t lo = i * chunk, m = std::min(chunk, n - lo);
        ONI_CUDA(cudaStreamWaitEvent(c->stream, c->up_done[b], 0));
        const int rc = oni_score_events_device(c, is_flow, m, idx_base + lo, (int32_t *)c->up[b][0].p,
                                               (int32_t *)c->up[b][2].p, (int32_t *)c->up[b][1].p,
                                               (int32_t *)c->up[b][3].p, tol);
        if (rc != ONI_OK) { cudaStreamSynchronize(c->copy_stream); return rc; }
    }
    return ONI_OK;
}

// src/dst scores of the first `cap` ranked survivors of events that live in HOST arrays: their
// index quadruples are gathered on the host (cap entries, however many chunks the events took),
// uploaded, and scored once more in ranked order.
extern "C" int oni_score_fetch_sides_host(oni_score_ctx *c, int32_t is_flow, int64_t cap, int64_t idx_base,
                                          const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                                          const int32_t *b_word, double *out_a, double *out_b)
{
    if (!c || cap < 0 || !a_doc || !a_word || (is_flow && (!b_doc || !b_word)))
        return oni_set_error(ONI_ERR_ARG, "score_fetch_sides_host: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    const size_t m = (size_t)std::min<unsigned long long>((unsigned long long)cap, c->n_surv);
    if (m == 0) return ONI_OK;
    std::vector<int64_t> idx(m);
    ONI_CUDA(cudaMemcpy(idx.data(), c->idx[c->cur].p, sizeof(int64_t) * m, cudaMemcpyDeviceToHost));
    const int nsrc = is_flow ? 4 : 2;
    const int32_t *src[4] = {a_doc, a_word, b_doc, b_word};
    std::vector<int32_t> g[4];
    for (int q = 0; q < nsrc; q++) {
        g[q].resize(m);
        for (size_t r = 0; r < m; r++) g[q][r] = src[q][idx[r] - idx_base];
        ONI_TRY(c->ev[q].reserve(sizeof(int32_t) * m));
        ONI_CUDA(cudaMemcpyAsync(c->ev[q].p, g[q].data(), sizeof(int32_t) * m, cudaMemcpyHostToDevice, c->stream));
    }
    ONI_TRY(c->out_a.reserve(sizeof(double) * m));
    ONI_TRY(c->out_b.reserve(sizeof(double) * m));
    const double dflt = is_flow ? 0.05 : 0.1;
    const int blocks = (int)std::min<long long>(((long long)m + 255) / 256, 4096);
    if (is_flow)
        rescored_kernel<true><<<blocks, 256, 0, c->stream>>>(c->K, c->d_theta, c->d_pwz, (long long)m, nullptr, 0,
            (int32_t *)c->ev[0].p, (int32_t *)c->ev[2].p, (int32_t *)c->ev[1].p, (int32_t *)c->ev[3].p, dflt,
            (double *)c->out_a.p, (double *)c->out_b.p);
    else
        rescored_kernel<false><<<blocks, 256, 0, c->stream>>>(c->K, c->d_theta, c->d_pwz, (long long)m, nullptr, 0,
            (int32_t *)c->ev[0].p, nullptr, (int32_t *)c->ev[1].p, nullptr, dflt,
            (double *)c->out_a.p, (double *)c->out_b.p);
    c->launches++;
    ONI_CUDA(cudaStreamSynchronize(c->stream));  // (the gathered vectors die at return)
    ONI_CUDA(cudaGetLastError());
    if (out_a) ONI_CUDA(cudaMemcpy(out_a, c->out_a.p, sizeof(double) * m, cudaMemcpyDeviceToHost));
    if (out_b && is_flow) ONI_CUDA(cudaMemcpy(out_b, c->out_b.p, sizeof(double) * m, cudaMemcpyDeviceToHost));
    return ONI_OK;
}

static int score_host(int32_t device, int is_flow, int32_t K, int32_t D, int32_t V, const double *theta,
                      const double *pwz, int64_t n, const int32_t *a_doc, const int32_t *b_doc,
                      const int32_t *a_word, const int32_t *b_word, double tol, int64_t out_cap,
                      int64_t *out_idx, double *out_key, double *out_a, double *out_b, int64_t *out_n,
                      int64_t base = 0 /* global index of this call's first event */)
{
    if (n < 0 || out_cap < 0 || !theta || !pwz) return oni_set_error(ONI_ERR_ARG, "score: bad argument");
    oni_score_ctx *c = nullptr;
    ONI_TRY(oni_score_create(device, K, D, V, theta, pwz, &c));
    int rc = oni_score_begin(c);
    // (indices are range-checked by the scoring kernel itself; -1 = lookup miss)
    if (rc == ONI_OK) rc = oni_score_events_host(c, is_flow, n, base, a_doc, b_doc, a_word, b_word, tol);
    int64_t total = 0;
    if (rc == ONI_OK) rc = oni_score_rank(c, &total);
    if (rc == ONI_OK) rc = oni_score_fetch(c, out_cap, out_idx, out_key);
    if (rc == ONI_OK && (out_a || out_b) && n > 0)
        rc = oni_score_fetch_sides_host(c, is_flow, out_cap, base, a_doc, b_doc, a_word, b_word, out_a, out_b);
    if (out_n) *out_n = total;
    oni_score_destroy(c);
    return rc;
}

extern "C" int oni_score_flow(int32_t device, int32_t K, int32_t D, int32_t V, const double *theta,
                              const double *pwz, int64_t n, const int32_t *src_doc, const int32_t *dst_doc,
                              const int32_t *src_word, const int32_t *dst_word, double tol, int64_t out_cap,
                              int64_t *out_idx, double *out_key, double *out_src, double *out_dst,
                              int64_t *out_n)
{
    if (n > 0 && (!src_doc || !dst_doc || !src_word || !dst_word))
        return oni_set_error(ONI_ERR_ARG, "score_flow: null event array");
    return score_host(device, 1, K, D, V, theta, pwz, n, src_doc, dst_doc, src_word, dst_word, tol, out_cap,
                      out_idx, out_key, out_src, out_dst, out_n);
}

extern "C" int oni_score_dns(int32_t device, int32_t K, int32_t D, int32_t V, const double *theta,
                             const double *pwz, int64_t n, const int32_t *doc, const int32_t *word, double tol,
                             int64_t out_cap, int64_t *out_idx, double *out_key, int64_t *out_n)
{
    if (n > 0 && (!doc || !word)) return oni_set_error(ONI_ERR_ARG, "score_dns: null event array");
    return score_host(device, 0, K, D, V, theta, pwz, n, doc, nullptr, word, nullptr, tol, out_cap, out_idx,
                      out_key, nullptr, nullptr, out_n);
}

// Events split evenly over the GPUs of this process: tables replicated, every GPU
// filters and ranks its own contiguous share, the host merges the ranked shares on (key, global
// event index).  No collective: the shares meet only in the merge.
static int score_multi(int32_t ngpus, int is_flow, int32_t K, int32_t D, int32_t V, const double *theta,
                       const double *pwz, int64_t n, const int32_t *a_doc, const int32_t *b_doc,
                       const int32_t *a_word, const int32_t *b_word, double tol, int64_t out_cap,
                       int64_t *out_idx, double *out_key, double *out_a, double *out_b, int64_t *out_n)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    if (ngpus < 1 || ngpus > ndev) return oni_set_error(ONI_ERR_ARG, "score: %d GPUs asked, %d visible", ngpus, ndev);
    if (n < 0 || out_cap < 0) return oni_set_error(ONI_ERR_ARG, "score: bad argument");
    struct Share {
        int64_t lo = 0, n = 0, total = 0;
        std::vector<int64_t> idx;
        std::vector<double> key, a, b;
        int rc = ONI_OK;
        std::string msg;
    };
    std::vector<Share> sh(ngpus);
    std::vector<std::thread> th;
    const bool sides = out_a || out_b;
    for (int g = 0; g < ngpus; g++) {
        Share &s = sh[g];
        s.lo = n * g / ngpus;
        s.n = n * (g + 1) / ngpus - s.lo;
        const size_t cap = (size_t)std::max<int64_t>(1, std::min(out_cap, s.n));
        s.idx.resize(cap); s.key.resize(cap);
        if (sides) { s.a.resize(cap); s.b.resize(cap); }
        th.emplace_back([&, g] {
            Share &me = sh[g];
            me.rc = score_host(g, is_flow, K, D, V, theta, pwz, me.n, a_doc + me.lo, b_doc ? b_doc + me.lo : nullptr,
                               a_word + me.lo, b_word ? b_word + me.lo : nullptr, tol, std::min(out_cap, me.n),
                               me.idx.data(), me.key.data(), sides ? me.a.data() : nullptr,
                               (sides && is_flow) ? me.b.data() : nullptr, &me.total, me.lo);
            if (me.rc != ONI_OK) me.msg = oni_last_error();  // (the message is per thread)
        });
    }
    for (auto &t : th) t.join();
    int64_t total = 0;
    for (int g = 0; g < ngpus; g++) {
        if (sh[g].rc != ONI_OK) return oni_set_error(sh[g].rc, "GPU %d: %s", g, sh[g].msg.c_str());
        total += sh[g].total;
    }
    // merge: ascending key, ties by global event index (each share is already in that order)
    std::vector<int64_t> head(ngpus, 0), have(ngpus);
    for (int g = 0; g < ngpus; g++) have[g] = std::min<int64_t>(sh[g].total, std::min(out_cap, sh[g].n));
    const int64_t emit = std::min(out_cap, total);
    for (int64_t r = 0; r < emit; r++) {
        int best = -1;
        for (int g = 0; g < ngpus; g++) {
            if (head[g] >= have[g]) continue;
            if (best < 0) { best = g; continue; }
            const double kb = sh[best].key[head[best]], kg = sh[g].key[head[g]];
            if (kg < kb || (kg == kb && sh[g].idx[head[g]] < sh[best].idx[head[best]])) best = g;
        }
        if (best < 0) break;
        const int64_t h = head[best]++;
        if (out_idx) out_idx[r] = sh[best].idx[h];
        if (out_key) out_key[r] = sh[best].key[h];
        if (out_a) out_a[r] = sh[best].a[h];
        if (out_b && is_flow) out_b[r] = sh[best].b[h];
    }
    if (out_n) *out_n = total;
    return ONI_OK;
}

extern "C" int oni_score_flow_multi(int32_t ngpus, int32_t K, int32_t D, int32_t V, const double *theta,
                                    const double *pwz, int64_t n, const int32_t *src_doc, const int32_t *dst_doc,
                                    const int32_t *src_word, const int32_t *dst_word, double tol, int64_t out_cap,
                                    int64_t *out_idx, double *out_key, double *out_src, double *out_dst,
                                    int64_t *out_n)
{
    if (!theta || !pwz || (n > 0 && (!src_doc || !dst_doc || !src_word || !dst_word)))
        return oni_set_error(ONI_ERR_ARG, "score_flow_multi: null array");
    return score_multi(ngpus, 1, K, D, V, theta, pwz, n, src_doc, dst_doc, src_word, dst_word, tol, out_cap, out_idx,
                       out_key, out_src, out_dst, out_n);
}

extern "C" int oni_score_dns_multi(int32_t ngpus, int32_t K, int32_t D, int32_t V, const double *theta,
                                   const double *pwz, int64_t n, const int32_t *doc, const int32_t *word, double tol,
                                   int64_t out_cap, int64_t *out_idx, double *out_key, int64_t *out_n)
{
    if (!theta || !pwz || (n > 0 && (!doc || !word))) return oni_set_error(ONI_ERR_ARG, "score_dns_multi: null array");
    return score_multi(ngpus, 0, K, D, V, theta, pwz, n, doc, nullptr, word, nullptr, tol, out_cap, out_idx, out_key,
                       nullptr, nullptr, out_n);
}

// ------------------------------------------------------------------ lda_post.py
extern "C" int oni_lda_post(int32_t device, int32_t D, int32_t K, int32_t V, const double *gamma,
                            const double *logbeta, int32_t text_digits, double *theta, double *pwz)
{
    if (D < 0 || K < 1 || V < 0 || (text_digits != 0 && (text_digits < 1 || text_digits > 17)))
        return oni_set_error(ONI_ERR_ARG, "lda_post: bad argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    ONI_CUDA(cudaSetDevice(device));
    char fmt[16];
    snprintf(fmt, sizeof fmt, "%%.%dg", text_digits);
    auto roundtrip = [&](double *a, size_t n) {
        if (!text_digits) return;
        char buf[64];
        for (size_t i = 0; i < n; i++) {  // Python-2 str(float): 12 significant digits
            snprintf(buf, sizeof buf, fmt, a[i]);
            a[i] = strtod(buf, nullptr);
        }
    };
    DevBuf buf[3];  // released on every way out
    auto run = [&]() -> int {
        if (gamma && theta && D > 0) {
            const size_t bytes = sizeof(double) * (size_t)D * K;
            ONI_TRY(buf[0].reserve(bytes)); ONI_TRY(buf[1].reserve(bytes));
            ONI_CUDA(cudaMemcpy(buf[0].p, gamma, bytes, cudaMemcpyHostToDevice));
            theta_kernel<<<std::min(4096, (D + 127) / 128), 128>>>(D, K, (double *)buf[0].p, (double *)buf[1].p);
            ONI_CUDA(cudaMemcpy(theta, buf[1].p, bytes, cudaMemcpyDeviceToHost));
            roundtrip(theta, (size_t)D * K);
        }
        if (logbeta && pwz && V > 0) {
            const size_t bytes = sizeof(double) * (size_t)V * K;
            ONI_TRY(buf[0].reserve(bytes)); ONI_TRY(buf[1].reserve(bytes)); ONI_TRY(buf[2].reserve(bytes));
            ONI_CUDA(cudaMemcpy(buf[0].p, logbeta, bytes, cudaMemcpyHostToDevice));
            pwz_kernel<<<K, 256>>>(K, V, (double *)buf[0].p, (double *)buf[1].p, (double *)buf[2].p);
            ONI_CUDA(cudaMemcpy(pwz, buf[2].p, bytes, cudaMemcpyDeviceToHost));
            roundtrip(pwz, (size_t)V * K);
        }
        ONI_CUDA(cudaGetLastError());
        return ONI_OK;
    };
    const int rc = run();
    for (auto &b : buf) b.release();
    return rc;
}

// ============================================================================
// Flow scoring + ranking without any allocation.
//
// Who this is for: hosts that are not allowed to grow device memory while they
// run.  A Spark executor or serving process that was handed a fixed slice of
// the GPU by a pool allocator (everything reserved up front, nothing left for
// the driver to give out), or a process with latency-critical streams next to
// this one: cudaMalloc/cudaFree may fail late when the pool owns the memory and
// they stall every stream of the device while they run.  Such a caller brings
// its own device buffers for the inputs and outputs; the working memory of the
// filter and the radix sort lives in the fixed array below instead of in
// buffers sized and allocated per call.
//
// The price is a hard ceiling: the array has room for ONI_NOALLOC_MAX_SURVIVORS
// events under the threshold (2^18).  It is part of the library image, so every
// GPU that loads liboni_lda.so carries it (about 8.2 MB) until the process ends.
//
// All calls on a device share the array and it keeps whatever the previous call
// left behind; each call clears the two pieces that must start at zero (the
// survivor counter and the digit totals) and overwrites the rest before reading
// it.  Two calls that overlap on the same GPU (different streams or threads)
// would trample each other's state -- run them one after the other.
// ============================================================================
namespace {

constexpr unsigned long long NA_MAX_SURVIVORS = 1ull << 18;
constexpr unsigned long long NA_MAX_TILES = NA_MAX_SURVIVORS / RS_TILE;  // 64
// layout, in 8-byte words
constexpr size_t NA_OFF_KEY0 = 0;
constexpr size_t NA_OFF_IDX0 = NA_OFF_KEY0 + NA_MAX_SURVIVORS;
constexpr size_t NA_OFF_KEY1 = NA_OFF_IDX0 + NA_MAX_SURVIVORS;
constexpr size_t NA_OFF_IDX1 = NA_OFF_KEY1 + NA_MAX_SURVIVORS;
constexpr size_t NA_OFF_COUNTS = NA_OFF_IDX1 + NA_MAX_SURVIVORS;       // [256][tiles]
constexpr size_t NA_OFF_TOTALS = NA_OFF_COUNTS + 256 * NA_MAX_TILES;   // [passes][256]
constexpr size_t NA_OFF_COUNTER = NA_OFF_TOTALS + RS_MAX_PASSES * 256;  // [1] (+ padding)
constexpr size_t NA_WORDS = NA_OFF_COUNTER + 32;

static __device__ __align__(256) unsigned long long g_noalloc_scratch[NA_WORDS];

// In host code `g_noalloc_scratch` is only a handle for the symbol; the address
// kernels can use comes from cudaGetSymbolAddress, once per device.
constexpr int NA_MAX_DEVICES = 64;
unsigned long long *g_noalloc_addr[NA_MAX_DEVICES] = {};

__global__ void emit_ranked_kernel(const unsigned long long *key, const unsigned long long *idx,
                                   long long m, long long *out_idx, double *out_key)
{
    for (long long r = blockIdx.x * (long long)blockDim.x + threadIdx.x; r < m;
         r += (long long)gridDim.x * blockDim.x) {
        if (out_idx) out_idx[r] = (long long)idx[r];
        if (out_key) out_key[r] = key_value(key[r]);
    }
}

}  // namespace

extern "C" int oni_score_flow_noalloc(int32_t device, int32_t K, const double *d_theta,
                                      const double *d_pwz, int64_t n, const int32_t *d_src_doc,
                                      const int32_t *d_dst_doc, const int32_t *d_src_word,
                                      const int32_t *d_dst_word, double tol, int64_t out_cap,
                                      int64_t *d_out_idx, double *d_out_key, int64_t *out_n,
                                      void *cuda_stream)
{
    if (K < 1 || n < 0 || out_cap < 0 || !d_theta || !d_pwz || !out_n ||
        (n > 0 && (!d_src_doc || !d_dst_doc || !d_src_word || !d_dst_word)))
        return oni_set_error(ONI_ERR_ARG, "score_flow_noalloc: bad argument");
    if (device < 0 || device >= NA_MAX_DEVICES)
        return oni_set_error(ONI_ERR_ARG, "score_flow_noalloc: device %d outside 0..%d", device, NA_MAX_DEVICES - 1);
    if ((unsigned long long)out_cap > NA_MAX_SURVIVORS)
        return oni_set_error(ONI_ERR_ARG, "score_flow_noalloc: out_cap %lld exceeds the fixed scratch (%llu ranked events)",
                             (long long)out_cap, NA_MAX_SURVIVORS);
    ONI_CUDA(cudaSetDevice(device));
    // set-up, first call on this device: resolve the array's device address
    if (!g_noalloc_addr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return oni_set_error(ONI_ERR_CUDA, "score_flow_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        g_noalloc_addr[device] = (unsigned long long *)p;
    }
    unsigned long long *base = g_noalloc_addr[device];
    unsigned long long *key[2] = {base + NA_OFF_KEY0, base + NA_OFF_KEY1};
    unsigned long long *idx[2] = {base + NA_OFF_IDX0, base + NA_OFF_IDX1};
    unsigned long long *counts = base + NA_OFF_COUNTS, *totals = base + NA_OFF_TOTALS;
    unsigned long long *counter = base + NA_OFF_COUNTER;
    cudaStream_t s = (cudaStream_t)cuda_stream;
    *out_n = 0;

    // the previous call's leftovers: only these two must start from zero
    ONI_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned long long), s));
    ONI_CUDA(cudaMemsetAsync(totals, 0, sizeof(unsigned long long) * RS_MAX_PASSES * 256, s));
    if (n == 0) { ONI_CUDA(cudaStreamSynchronize(s)); return ONI_OK; }

    int sms = 0;
    ONI_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    const int blocks = (int)std::min<long long>((n + 255) / 256, (long long)sms * 16);
    // (this entry point is not told the table sizes: no index is "beyond" them, as before)
    score_filter_kernel<true><<<blocks, 256, 0, s>>>(K, 0x7fffffff, 0x7fffffff, d_theta, d_pwz, n, d_src_doc, d_dst_doc, d_src_word,
                                                     d_dst_word, 0.05, tol, 0, key[0], idx[0],
                                                     NA_MAX_SURVIVORS, counter);
    unsigned long long total = 0;
    ONI_CUDA(cudaMemcpyAsync(&total, counter, sizeof total, cudaMemcpyDeviceToHost, s));
    ONI_CUDA(cudaStreamSynchronize(s));
    ONI_CUDA(cudaGetLastError());
    if (total > NA_MAX_SURVIVORS)
        return oni_set_error(ONI_ERR_ARG,
                             "score_flow_noalloc: %llu events fall below the threshold, the fixed scratch holds %llu; "
                             "lower the threshold, split the events, or use oni_score_flow",
                             total, NA_MAX_SURVIVORS);
    *out_n = (int64_t)total;
    const long long m = (long long)total;
    int cur = 0;
    if (m > 1) {
        const int tblocks = (int)std::min<long long>((m + RS_THREADS - 1) / RS_THREADS, (long long)sms * 8);
        radix_totals_kernel<<<tblocks, RS_THREADS, 0, s>>>(key[0], idx[0], m, totals);
        unsigned long long h_totals[RS_MAX_PASSES * 256];
        ONI_CUDA(cudaMemcpyAsync(h_totals, totals, sizeof h_totals, cudaMemcpyDeviceToHost, s));
        ONI_CUDA(cudaStreamSynchronize(s));
        const int ntiles = (int)((m + RS_TILE - 1) / RS_TILE);
        for (int pass = 0; pass < RS_MAX_PASSES; pass++) {
            bool trivial = false;
            for (int d = 0; d < 256; d++)
                if (h_totals[pass * 256 + d] == (unsigned long long)m) { trivial = true; break; }
            if (trivial) continue;
            radix_count_kernel<<<ntiles, RS_THREADS, 0, s>>>(key[cur], idx[cur], m, pass, ntiles, counts);
            radix_scan_kernel<<<256, 256, 0, s>>>(counts, ntiles, totals + (size_t)pass * 256);
            radix_scatter_kernel<<<ntiles, RS_THREADS, 0, s>>>(key[cur], idx[cur], key[cur ^ 1], idx[cur ^ 1], m,
                                                               pass, ntiles, counts);
            cur ^= 1;
        }
    }
    const long long emit = std::min<long long>(m, out_cap);
    if (emit > 0 && (d_out_idx || d_out_key)) {
        const int eblocks = (int)std::min<long long>((emit + 255) / 256, 1024);
        emit_ranked_kernel<<<eblocks, 256, 0, s>>>(key[cur], idx[cur], emit, (long long *)d_out_idx, d_out_key);
    }
    ONI_CUDA(cudaStreamSynchronize(s));
    ONI_CUDA(cudaGetLastError());
    return ONI_OK;
}
