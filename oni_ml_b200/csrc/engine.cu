// engine.cu -- the step-wise LDA engine behind include/oni_lda.h: one context per
// GPU holding a shard of the documents, the replicated model and the packed
// sufficient statistics.  Replaces run_em (oni-lda-c/lda-estimate.c:112-321).
//
// HBM layout of a context (D docs, NNZ pairs in the shard, V terms, K topics,
// KP = K rounded up to 4 so every row starts on a 32-byte boundary):
//   tok        int2   [NNZ]      (word, count) of each pair, document order
//   doc_ptr    int64  [D+1]
//   order      int32  [D]        documents, longest first (schedule only)
//   doc_total  double [D]        sum of counts
//   perm       int32  [NNZ]      pair ids sorted by word id (stable)
//   items/items2                 chunk lists of the deterministic reduction
//   logbeta_t  double [V][KP]    word-major log p(w|k): one row per word visit
//   phi        double [NNZ][KP]  variational phi; read+written every sweep
//   gamma      double [D][K]
//   doc_lik/doc_ass double [D], doc_iters int32 [D]
//   ss         double [V*KP + KP + 4]  class_word | class_total | tail
//   partial    double [n_partial][KP]  scratch of the reduction's level 1
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <numeric>
#include <thread>
#include <vector>

#include "common.h"
#include "corpus_index.cuh"
#include "estep_launch.h"
#include "mt19937.h"
#include "psi_table.h"
#include "radix_sort.cuh"
#include "suffstats.cuh"

using namespace oni;

static const int TIER_LEN_CAP = 65535;  // document lengths above it share one bucket of docs_ge
static const MathTables g_host_tables = ONI_MATH_TABLES_INIT;  // host copy (digamma_ref on the host)

// ----------------------------------------------------------------- NCCL, lazily
// The library must load on hosts without NCCL (single-GPU use), so the few
// entry points are resolved with dlopen on first use.
namespace {
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[ONI_NCCL_ID_BYTES]; } ncclUniqueId;
enum { kNcclSum = 0, kNcclFloat64 = 8 };
struct NcclApi {
    void *handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};
NcclApi g_nccl;

int nccl_load()
{
    if (g_nccl.handle) return ONI_OK;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    void *h = nullptr;
    for (const char *n : names) {
        h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (h) break;
    }
    if (!h) return oni_set_error(ONI_ERR_NCCL, "cannot load libnccl: %s", dlerror());
#define ONI_SYM(field, sym)                                                         \
    *(void **)(&g_nccl.field) = dlsym(h, sym);                                      \
    if (!g_nccl.field) return oni_set_error(ONI_ERR_NCCL, "libnccl lacks %s", sym);
    ONI_SYM(GetUniqueId, "ncclGetUniqueId")
    ONI_SYM(CommInitRank, "ncclCommInitRank")
    ONI_SYM(CommInitAll, "ncclCommInitAll")
    ONI_SYM(CommDestroy, "ncclCommDestroy")
    ONI_SYM(AllReduce, "ncclAllReduce")
    ONI_SYM(GroupStart, "ncclGroupStart")
    ONI_SYM(GroupEnd, "ncclGroupEnd")
    ONI_SYM(GetErrorString, "ncclGetErrorString")
#undef ONI_SYM
    g_nccl.handle = h;
    return ONI_OK;
}
#define ONI_NCCL(call)                                                              \
    do {                                                                            \
        int r_ = (call);                                                            \
        if (r_ != 0)                                                                \
            return oni_set_error(ONI_ERR_NCCL, "%s failed: %s", #call,              \
                                 g_nccl.GetErrorString(r_));                        \
    } while (0)
}  // namespace

// ------------------------------------------------------------------ the context
struct oni_lda_ctx {
    int device = 0;
    int K = 0, KP = 0, V = 0, G = 0, T = 0;
    int G_long = 32;       // long documents get wider groups: G_long lanes, T_long topics per lane
    int T_long = 0;
    int long_min = 0;      // documents with at least this many pairs are "long" (0 = none)
    bool long_min_auto = false;  // ... chosen per corpus from its size (no ONI_LDA_LONG_DOC)
    int32_t n_long = 0;    // they are the first n_long entries of `order`
    int32_t n_excl = 0;    // the first n_excl of those get an SM to themselves (4 documents per SM)
    std::vector<int32_t> docs_ge;  // docs_ge[L] = documents of >= L pairs (L <= TIER_LEN_CAP); empty = tiers fixed
    int32_t n_small = 0;   // documents of at most one pair: the LAST n_small entries of `order`
    oni_lda_settings st{};
    int var_max_iter = 0;  // doubles when the bound decreases (lda-estimate.c:261)
    cudaStream_t stream = nullptr, stream_long = nullptr, stream_excl = nullptr;
    cudaEvent_t ev[6] = {}, ev_fork = nullptr, ev_join = nullptr, ev_join_excl = nullptr,
                ev_t_long = nullptr, ev_t_excl = nullptr, ev_t_main = nullptr;  // ONI_LDA_TIMING: ends of the three launches
    int sm_count = 0;

    // corpus shard.  The buffers grow and never shrink while the context lives: a caller that
    // runs job after job on one handle (bench.py's e2e, a daily pipeline) pays cudaMalloc / cudaFree
    // of ~10 GB once -- those calls were seen to take anything between 30 and 300 ms.
    int D = 0;
    int64_t nnz = 0;
    size_t cap_docs = 0, cap_pairs = 0, cap_items = 0, cap_items2 = 0, cap_partial = 0;
    int2 *d_tok = nullptr;
    int64_t *d_doc_ptr = nullptr;
    int32_t *d_order = nullptr;
    double *d_doc_total = nullptr;
    int32_t *d_perm = nullptr;
    SsItem *d_items = nullptr, *d_items2 = nullptr;
    int32_t n_items = 0, n_items2 = 0, n_partial = 0;
    std::vector<int32_t> level_off;  // items2[level_off[l] .. level_off[l+1]) = level l+2 of the reduction

    // model + variational state
    double *d_logbeta_t = nullptr, *d_alpha = nullptr;
    double *d_phi = nullptr, *d_gamma = nullptr, *d_doc_lik = nullptr, *d_doc_ass = nullptr;
    int32_t *d_doc_iters = nullptr;
    double *d_ss = nullptr, *d_partial = nullptr, *d_block_partial = nullptr;
    double *d_long_tabs = nullptr;  // psi_table.h: [PSI_N*6] psi coefficients, then [EXK_N] 2^(i/2048)
    bool long_kernel = false;       // long documents go to estep_long_kernel (K <= 128)
    // What the unshifted kernels need to know about a model that came from outside (set_model,
    // `lda inf`, a model-root start): the smallest entry, and the smallest per-word maximum over
    // topics.  Models written by the M-step never go below the -100 floor (lda-model.c:42).
    double model_min = -100.0, model_min_rowmax = -100.0;
    int64_t max_doc_pairs = 0;
    unsigned int *d_ticket = nullptr, *d_done = nullptr;
    int32_t *d_nan = nullptr;
    double *h_pinned = nullptr;  // [8] tail + alpha + nan flag readback

    // EM bookkeeping (lda-estimate.c:174-176)
    int em_iter = 0;
    double alpha = 0.0, lik_old = 0.0, converged = 1.0;
    bool have_model = false, estep_done = false;

    ncclComm_t comm = nullptr;
    int nranks = 1;
    int64_t launches = 0;
    oni_lda_stats stats{};

    size_t ss_count() const { return (size_t)V * KP + KP + 4; }
    double *ss_class_total() const { return d_ss + (size_t)V * KP; }
    double *ss_tail() const { return d_ss + (size_t)V * KP + KP; }
};

namespace {

// lanes per document (G) and topics per lane (T): fewest padded slots with
// T <= 8, ties to the smaller G (fewer shuffle steps).
bool pick_config(int K, int *G, int *T)
{
    // instantiated: G = 1 with T = 1..8, and G in {2,4,8,16,32} with T = 5..8
    int best_g = 0, best_t = 0, best_slots = 1 << 30;
    for (int g = 1; g <= 32; g <<= 1) {
        const int t = (K + g - 1) / g;
        if (t > 8 || (g > 1 && t < 5)) continue;
        if (g * t < best_slots) { best_slots = g * t; best_g = g; best_t = t; }
    }
    if (!best_g) return false;
    *G = best_g; *T = best_t;
    return true;
}

template <typename Tp>
int dev_alloc(Tp **p, size_t n)
{
    ONI_CUDA(cudaMalloc((void **)p, sizeof(Tp) * (n ? n : 1)));
    return ONI_OK;
}

// the buffers sized by the number of documents / pairs
void free_corpus_buffers(oni_lda_ctx *c)
{
    cudaFree(c->d_tok); cudaFree(c->d_doc_ptr); cudaFree(c->d_order); cudaFree(c->d_doc_total);
    cudaFree(c->d_perm); cudaFree(c->d_phi);
    cudaFree(c->d_gamma); cudaFree(c->d_doc_lik); cudaFree(c->d_doc_ass); cudaFree(c->d_doc_iters);
    c->d_tok = nullptr; c->d_doc_ptr = nullptr; c->d_order = nullptr; c->d_doc_total = nullptr;
    c->d_perm = nullptr; c->d_phi = nullptr;
    c->d_gamma = c->d_doc_lik = c->d_doc_ass = nullptr; c->d_doc_iters = nullptr;
    c->cap_docs = c->cap_pairs = 0;
}

void free_corpus(oni_lda_ctx *c)
{
    free_corpus_buffers(c);
    cudaFree(c->d_items); cudaFree(c->d_items2); cudaFree(c->d_partial);
    c->d_items = c->d_items2 = nullptr; c->d_partial = nullptr;
    c->cap_items = c->cap_items2 = c->cap_partial = 0;
    c->D = 0; c->nnz = 0; c->n_long = 0; c->n_excl = 0; c->n_small = 0;
}

double now_seconds();

struct StageTimer {
    bool on = getenv("ONI_LDA_TIMING") != nullptr;
    double t = now_seconds();
    void lap(const char *what)
    {
        if (!on) return;
        const double n = now_seconds();
        fprintf(stderr, "[oni timing] %-28s %8.1f ms\n", what, (n - t) * 1e3);
        t = n;
    }
};

// Once per corpus: everything is derived on the device from the CSR arrays
// (corpus_index.cuh); only the chunk lists of the reduction (one entry per word
// plus one per 256 pairs) are formed on the host from the word histogram.
// d_ptr/d_word/d_count: device arrays; h_ptr: host copy of doc_ptr.
int build_indices(oni_lda_ctx *c, int32_t D, int64_t nnz, const int64_t *h_ptr, const int64_t *d_ptr,
                  const int32_t *d_word, const int32_t *d_count)
{
    StageTimer tm;
    if (nnz > 0x7fffffffLL)
        return oni_set_error(ONI_ERR_ARG, "shard has %lld pairs; at most 2^31-1 per GPU", (long long)nnz);
    const int V = c->V;
    for (int32_t d = 0; d < D; d++)
        if (h_ptr[d + 1] < h_ptr[d]) return oni_set_error(ONI_ERR_ARG, "doc_ptr not monotone at %d", d);
    if (h_ptr[0] != 0 || h_ptr[D] != nnz) return oni_set_error(ONI_ERR_ARG, "doc_ptr does not span the pairs");
    // Three tiers, by the length of a document's serial chain of word visits against the time
    // the shard's bulk takes (~1.6 ns per pair on one B200 at K = 20):
    //   main      12 warps per SM, 8 documents per warp: ~1.3 us per visit of one document
    //   long      estep_long_kernel, one document per warp, its blocks of four warps sharing SMs
    //             with the main launch's: ~0.35 us per visit
    //   longest   the same kernel alone on its SMs (one warp per scheduler): 0.2 us per visit
    // A document moves up a tier when its chain at the lower tier's pace would last ~0.8 of the
    // bulk time: from nnz/15500 pairs to the long tier (flow day on one GPU: capped at 1024; an
    // eighth of it: 440 -- profiles/tier_sweep.sh: 360 and 576 both end the main launch ~1 ms later,
    // 256 and 768 ~4 ms later), from nnz/6000 to the top one.  (Measured and dropped for the middle
    // tier: the main build with one block per SM -- a lone warp of it runs at 1.7 us per visit,
    // slower than at full occupancy -- and twelve long-kernel warps per SM, see estep_long.cuh.)
    // ONI_LDA_LONG_DOC / ONI_LDA_EXCL_DOC override the two thresholds (0 = tier off).
    int long_min = c->long_min;
    if (c->long_min_auto && c->long_kernel) long_min = (int)std::min<int64_t>(1024, std::max<int64_t>(128, nnz / 15500));
    int32_t n_long = 0;
    int64_t pairs_long = 0, max_pairs = 0;
    int32_t n_small = 0;
    const bool retier = c->long_min_auto && c->long_kernel && !getenv("ONI_LDA_EXCL_DOC") && c->G_long == 32 &&
                        c->T_long <= 4 && !(getenv("ONI_LDA_RETIER") && atoi(getenv("ONI_LDA_RETIER")) == 0);
    c->docs_ge.clear();
    if (retier) c->docs_ge.assign(TIER_LEN_CAP + 2, 0);
    for (int32_t d = 0; d < D; d++) {
        const int64_t n = h_ptr[d + 1] - h_ptr[d];
        max_pairs = std::max(max_pairs, n);
        n_small += n <= 1;
        if (long_min > 0 && n >= long_min) { n_long++; pairs_long += n; }
        if (retier) c->docs_ge[std::min<int64_t>(n, TIER_LEN_CAP)]++;
    }
    if (retier)
        for (int L = TIER_LEN_CAP - 1; L >= 0; L--) c->docs_ge[L] += c->docs_ge[L + 1];
    int32_t n_excl = 0;
    int64_t pairs_excl = 0;
    if (n_long > 0 && c->G_long == 32 && c->T_long <= 4) {
        const char *env = getenv("ONI_LDA_EXCL_DOC");
        const int64_t excl_min = env ? atoll(env) : std::max<int64_t>(long_min, nnz / (c->long_kernel ? 6000 : 20000));
        if (excl_min > 0)
            for (int32_t d = 0; d < D; d++)
                if (h_ptr[d + 1] - h_ptr[d] >= excl_min) { n_excl++; pairs_excl += h_ptr[d + 1] - h_ptr[d]; }
        const int32_t cap = std::min<int32_t>(n_long, 4 * std::max(1, c->sm_count / 8));  // at most 1/8 of the SMs
        if (n_excl > cap) { n_excl = cap; pairs_excl = 0; /* (only used for the middle tier's share below) */ }
    }
    tm.lap("validate doc_ptr");

    c->D = D; c->nnz = nnz; c->n_long = n_long; c->n_excl = n_excl; c->max_doc_pairs = max_pairs;
    c->n_small = std::min<int32_t>(n_small, D - n_long);
    cudaStream_t s = c->stream;
    if (!c->d_doc_ptr || (size_t)D > c->cap_docs || (size_t)nnz > c->cap_pairs) {
        // (one of the two grew: everything sized by it is replaced; sizes are kept with a little headroom)
        const size_t nd = std::max<size_t>((size_t)D, c->cap_docs), np = std::max<size_t>((size_t)nnz, c->cap_pairs);
        ONI_CUDA(cudaStreamSynchronize(s));
        free_corpus_buffers(c);
        ONI_TRY(dev_alloc(&c->d_tok, np));
        ONI_TRY(dev_alloc(&c->d_doc_ptr, nd + 1));
        ONI_TRY(dev_alloc(&c->d_order, nd));
        ONI_TRY(dev_alloc(&c->d_doc_total, nd));
        ONI_TRY(dev_alloc(&c->d_perm, np));
        ONI_TRY(dev_alloc(&c->d_phi, np * c->KP));
        ONI_TRY(dev_alloc(&c->d_gamma, nd * c->K));
        ONI_TRY(dev_alloc(&c->d_doc_lik, nd));
        ONI_TRY(dev_alloc(&c->d_doc_ass, nd));
        ONI_TRY(dev_alloc(&c->d_doc_iters, nd));
        c->cap_docs = nd; c->cap_pairs = np;
    }
    // scratch of the two sorts: 4 x 8 bytes per entry, carved out of the phi buffer (unused until the
    // first E-step) when it is large enough, i.e. whenever KP >= 4 pairs' worth covers the documents
    const size_t m = (size_t)std::max<int64_t>(nnz, D);
    const size_t ntiles = (m + RS_TILE - 1) / RS_TILE;
    unsigned long long *key[2] = {nullptr, nullptr}, *idx[2] = {nullptr, nullptr}, *counts = nullptr, *totals = nullptr;
    int64_t *d_wptr = nullptr;
    const bool in_phi = 4 * m <= c->cap_pairs * (size_t)c->KP;
    int rc = ONI_OK;
    auto release = [&]() {
        if (!in_phi) { cudaFree(key[0]); cudaFree(key[1]); cudaFree(idx[0]); cudaFree(idx[1]); }
        cudaFree(counts); cudaFree(totals); cudaFree(d_wptr);
    };
#define ONI_IDX(call) do { rc = (call); if (rc != ONI_OK) { release(); return rc; } } while (0)
#define ONI_IDX_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { release(); \
        return oni_set_error(ONI_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
    if (in_phi) {
        unsigned long long *base = reinterpret_cast<unsigned long long *>(c->d_phi);
        key[0] = base; key[1] = base + m; idx[0] = base + 2 * m; idx[1] = base + 3 * m;
    } else {
        for (int b = 0; b < 2; b++) { ONI_IDX(dev_alloc(&key[b], m)); ONI_IDX(dev_alloc(&idx[b], m)); }
    }
    ONI_IDX(dev_alloc(&counts, 256 * ntiles));
    ONI_IDX(dev_alloc(&totals, (size_t)RS_MAX_PASSES * 256));
    ONI_IDX(dev_alloc(&d_wptr, (size_t)V + 1));
    tm.lap("device allocations");

    const int grid = c->sm_count * 8;
    ONI_IDX_CUDA(cudaMemcpyAsync(c->d_doc_ptr, d_ptr, sizeof(int64_t) * ((size_t)D + 1), cudaMemcpyDeviceToDevice, s));
    ONI_IDX_CUDA(cudaMemsetAsync(c->d_nan, 0, sizeof(int32_t), s));
    // documents: totals + longest-first order (stable: ties keep document order)
    int cur = 0;
    if (D > 0) {
        doc_meta_kernel<<<grid, 256, 0, s>>>(d_ptr, d_count, D, c->d_doc_total, key[0], idx[0]);
        ONI_IDX(radix_sort_pairs(key, idx, counts, totals, D, 8, c->sm_count, s, &cur, &c->launches));
        narrow_kernel<<<grid, 256, 0, s>>>(idx[cur], D, c->d_order);
        c->launches += 2;
    }
    // pairs: pack, range-check, and sort by word id (stable: document order inside a word)
    int32_t bad = 0;
    if (nnz > 0) {
        pack_tok_kernel<<<grid, 256, 0, s>>>(d_word, d_count, nnz, V, c->d_tok, key[0], idx[0], c->d_nan);
        ONI_IDX_CUDA(cudaMemcpyAsync(&bad, c->d_nan, sizeof bad, cudaMemcpyDeviceToHost, s));
        ONI_IDX(radix_sort_pairs(key, idx, counts, totals, nnz, 8, c->sm_count, s, &cur, &c->launches));
        narrow_kernel<<<grid, 256, 0, s>>>(idx[cur], nnz, c->d_perm);
        c->launches += 2;
    }
    word_ptr_kernel<<<(V + 256) / 256, 256, 0, s>>>(nnz > 0 ? key[cur] : key[0], nnz, V, d_wptr);
    c->launches++;
    std::vector<int64_t> wptr((size_t)V + 1);
    ONI_IDX_CUDA(cudaMemcpyAsync(wptr.data(), d_wptr, sizeof(int64_t) * wptr.size(), cudaMemcpyDeviceToHost, s));
    ONI_IDX_CUDA(cudaMemsetAsync(c->d_gamma, 0, sizeof(double) * (size_t)D * c->K, s));
    ONI_IDX_CUDA(cudaMemsetAsync(c->d_doc_iters, 0, sizeof(int32_t) * (size_t)D, s));
    ONI_IDX_CUDA(cudaStreamSynchronize(s));
    ONI_IDX_CUDA(cudaMemsetAsync(c->d_nan, 0, sizeof(int32_t), s));
    release();
#undef ONI_IDX
#undef ONI_IDX_CUDA
    if (bad) return oni_set_error(ONI_ERR_ARG, "a word id lies outside [0,%d)", V);
    tm.lap("device indexing");

    // chunk lists of the deterministic reduction: level 1 cuts a word's pairs into chunks
    // of <= SS_CHUNK; a word with several chunks has its partial rows reduced the same way
    // by the next level, and so on until one row per word is left (a word holding 5 % of a
    // flow-day corpus needs three levels instead of one warp adding 10^4 rows in sequence)
    std::vector<SsItem> items, items2;
    std::vector<int32_t> level_off{0};
    int32_t n_partial = 0;
    struct Pending { int w; int64_t first, n; };
    std::vector<Pending> pending, next;
    for (int w = 0; w < V; w++) {
        const int64_t a = wptr[w], n = wptr[w + 1] - a;
        if (n <= SS_CHUNK) {
            items.push_back(SsItem{a, (int32_t)n, w});
        } else {
            const int32_t first = n_partial;
            for (int64_t o = 0; o < n; o += SS_CHUNK) {
                items.push_back(SsItem{a + o, (int32_t)std::min<int64_t>(SS_CHUNK, n - o), ~n_partial});
                n_partial++;
            }
            pending.push_back(Pending{w, first, n_partial - first});
        }
    }
    while (!pending.empty()) {
        next.clear();
        for (const Pending &pd : pending) {
            if (pd.n <= SS_CHUNK) {
                items2.push_back(SsItem{pd.first, (int32_t)pd.n, pd.w});
            } else {
                const int32_t first = n_partial;
                for (int64_t o = 0; o < pd.n; o += SS_CHUNK) {
                    items2.push_back(SsItem{pd.first + o, (int32_t)std::min<int64_t>(SS_CHUNK, pd.n - o), ~n_partial});
                    n_partial++;
                }
                next.push_back(Pending{pd.w, first, n_partial - first});
            }
        }
        level_off.push_back((int32_t)items2.size());
        pending.swap(next);
    }
    c->level_off = level_off;
    c->n_items = (int32_t)items.size(); c->n_items2 = (int32_t)items2.size(); c->n_partial = n_partial;
    if (items.size() > c->cap_items) { cudaFree(c->d_items); c->d_items = nullptr; c->cap_items = 0;
        ONI_TRY(dev_alloc(&c->d_items, items.size())); c->cap_items = items.size(); }
    if (items2.size() > c->cap_items2 || !c->d_items2) { cudaFree(c->d_items2); c->d_items2 = nullptr; c->cap_items2 = 0;
        ONI_TRY(dev_alloc(&c->d_items2, items2.size())); c->cap_items2 = items2.size(); }
    if ((size_t)n_partial > c->cap_partial || !c->d_partial) { cudaFree(c->d_partial); c->d_partial = nullptr; c->cap_partial = 0;
        ONI_TRY(dev_alloc(&c->d_partial, (size_t)n_partial * c->KP)); c->cap_partial = (size_t)n_partial; }
    ONI_CUDA(cudaMemcpyAsync(c->d_items, items.data(), sizeof(SsItem) * items.size(), cudaMemcpyHostToDevice, s));
    ONI_CUDA(cudaMemcpyAsync(c->d_items2, items2.data(), sizeof(SsItem) * items2.size(), cudaMemcpyHostToDevice, s));
    ONI_CUDA(cudaStreamSynchronize(s));  // host vectors die at return
    tm.lap("chunk lists");
    return ONI_OK;
}

template <int G, int T>
void launch_ss(oni_lda_ctx *c)
{
    const int warps_per_block = 4;
    if (c->n_items > 0) {
        const int blocks = (c->n_items + warps_per_block - 1) / warps_per_block;
        ss_rows_kernel<G, T, 1><<<blocks, 128, 0, c->stream>>>(
            c->d_items, c->n_items, c->d_perm, c->d_tok, c->d_phi, c->d_ss, c->d_partial, c->K, c->KP);
        c->launches++;
    }
    for (size_t l = 0; l + 1 < c->level_off.size(); l++) {
        const int32_t lo = c->level_off[l], n = c->level_off[l + 1] - lo;
        if (n <= 0) continue;
        const int blocks = (n + warps_per_block - 1) / warps_per_block;
        ss_rows_kernel<G, T, 2><<<blocks, 128, 0, c->stream>>>(
            c->d_items2 + lo, n, nullptr, nullptr, c->d_partial, c->d_ss, c->d_partial, c->K, c->KP);
        c->launches++;
    }
}

int dispatch_ss(oni_lda_ctx *c)
{
#define ONI_SS_CASE(G_, T_) \
    if (c->G == G_ && c->T == T_) { launch_ss<G_, T_>(c); return ONI_OK; }
#define ONI_SS_T58(G_) ONI_SS_CASE(G_, 5) ONI_SS_CASE(G_, 6) ONI_SS_CASE(G_, 7) ONI_SS_CASE(G_, 8)
    ONI_SS_CASE(1, 1) ONI_SS_CASE(1, 2) ONI_SS_CASE(1, 3) ONI_SS_CASE(1, 4)
    ONI_SS_T58(1) ONI_SS_T58(2) ONI_SS_T58(4) ONI_SS_T58(8) ONI_SS_T58(16) ONI_SS_T58(32)
#undef ONI_SS_T58
#undef ONI_SS_CASE
    return oni_set_error(ONI_ERR_ARG, "no reduction kernel for G=%d T=%d", c->G, c->T);
}

double now_seconds()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

}  // namespace

// ------------------------------------------------------------------- lifecycle
static int create_impl(int32_t device, int32_t K, int32_t V, const oni_lda_settings *st, oni_lda_ctx *c);

extern "C" int oni_lda_create(int32_t device, int32_t K, int32_t V, const oni_lda_settings *st,
                              oni_lda_ctx **out)
{
    if (!out || !st) return oni_set_error(ONI_ERR_ARG, "create: null argument");
    *out = nullptr;
    if (K < 1 || V < 1) return oni_set_error(ONI_ERR_ARG, "create: num_topics=%d num_terms=%d", K, V);
    int G, T;
    if (!pick_config(K, &G, &T))
        return oni_set_error(ONI_ERR_ARG, "num_topics=%d not supported (1..256)", K);
    // (the reference runs ZERO sweeps for 0 and for anything below -1 and then reports a bound of
    //  0 -- lda-inference.c:46-47; that is a bug there, refused here)
    if (st->var_max_iter == 0 || st->var_max_iter < -1)
        return oni_set_error(ONI_ERR_ARG, "var max iter %d is not supported (a positive cap, or -1 for none)",
                             st->var_max_iter);
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    if (device < 0 || device >= ndev)
        return oni_set_error(ONI_ERR_ARG, "device %d of %d", device, ndev);
    oni_lda_ctx *c = new oni_lda_ctx();
    c->G = G; c->T = T;
    const int rc = create_impl(device, K, V, st, c);
    if (rc != ONI_OK) { oni_lda_destroy(c); return rc; }  // nothing half-built survives a failure
    *out = c;
    return ONI_OK;
}

static int create_impl(int32_t device, int32_t K, int32_t V, const oni_lda_settings *st, oni_lda_ctx *c)
{
    const int G = c->G, T = c->T;
    ONI_CUDA(cudaSetDevice(device));
    c->device = device; c->K = K; c->KP = (K + 3) & ~3; c->V = V; c->G = G; c->T = T;
    c->st = *st; c->var_max_iter = st->var_max_iter;
    cudaDeviceProp prop;
    ONI_CUDA(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    ONI_CUDA(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    ONI_CUDA(cudaStreamCreateWithFlags(&c->stream_long, cudaStreamNonBlocking));
    ONI_CUDA(cudaStreamCreateWithFlags(&c->stream_excl, cudaStreamNonBlocking));
    ONI_CUDA(cudaEventCreateWithFlags(&c->ev_join_excl, cudaEventDisableTiming));
    ONI_CUDA(cudaEventCreate(&c->ev_t_long)); ONI_CUDA(cudaEventCreate(&c->ev_t_excl)); ONI_CUDA(cudaEventCreate(&c->ev_t_main));
    for (auto &e : c->ev) ONI_CUDA(cudaEventCreate(&e));
    ONI_CUDA(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    ONI_CUDA(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
    // A document's words are visited strictly in order, so a very long document is a
    // serial chain; those get a whole warp (G = 32) to shorten the chain per visit.
    if (G < 32) {
        const char *env = getenv("ONI_LDA_LONG_DOC");
        c->long_min = env ? atoi(env) : 1024;
        c->long_min_auto = !env;
        if (c->long_min > 0 && c->long_min < 64) c->long_min = 64;  // the long-document build requests rows 16 visits ahead
        const char *g = getenv("ONI_LDA_LONG_G");  // tuning knob: 8, 16 or 32
        c->G_long = (g && (atoi(g) == 8 || atoi(g) == 16) && atoi(g) > G) ? atoi(g) : 32;
        c->T_long = (K + c->G_long - 1) / c->G_long;
        if (c->T_long > 8) { c->G_long = 32; c->T_long = (K + 31) / 32; }
    }
    if (G < 32 && c->G_long == 32 && K <= 128) {
        // the short-chain kernel for long documents and its tables (ONI_LDA_LONG_KERNEL=0: the
        // one-document-per-warp build of estep_kernel instead, for A/B runs)
        const char *env = getenv("ONI_LDA_LONG_KERNEL");
        c->long_kernel = !(env && atoi(env) == 0);
    }
    if (c->long_kernel) {
        std::vector<double> tabs((size_t)PSI_N * PSI_COEF + EXK_N);
        build_long_tables(tabs.data(), tabs.data() + (size_t)PSI_N * PSI_COEF);
        ONI_TRY(dev_alloc(&c->d_long_tabs, tabs.size()));
        ONI_CUDA(cudaMemcpy(c->d_long_tabs, tabs.data(), sizeof(double) * tabs.size(), cudaMemcpyHostToDevice));
    }
    ONI_TRY(dev_alloc(&c->d_logbeta_t, (size_t)V * c->KP));
    if (getenv("ONI_LDA_L2_PERSIST") && atoi(getenv("ONI_LDA_L2_PERSIST")) > 0) {
        // Experiment (profiles/r2_estep_experiments.txt): keep the log-beta table (8 MB at flow-day
        // size) in a persisting L2 window, so that the phi rows streaming through the L2 (267 GB per
        // launch) do not evict the rows every word visit gathers.
        const size_t bytes = sizeof(double) * (size_t)V * c->KP;
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min<size_t>(bytes, (size_t)prop.persistingL2CacheMaxSize));
        cudaStreamAttrValue av{};
        av.accessPolicyWindow.base_ptr = c->d_logbeta_t;
        av.accessPolicyWindow.num_bytes = std::min<size_t>(bytes, (size_t)prop.accessPolicyMaxWindowSize);
        av.accessPolicyWindow.hitRatio = 1.0f;
        av.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        av.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        for (cudaStream_t st2 : {c->stream, c->stream_long, c->stream_excl})
            ONI_CUDA(cudaStreamSetAttribute(st2, cudaStreamAttributeAccessPolicyWindow, &av));
    }
    ONI_TRY(dev_alloc(&c->d_alpha, 1));
    ONI_TRY(dev_alloc(&c->d_ss, c->ss_count()));
    ONI_TRY(dev_alloc(&c->d_block_partial, (size_t)DOC_RED_BLOCKS * 3));
    ONI_TRY(dev_alloc(&c->d_ticket, 4));
    ONI_TRY(dev_alloc(&c->d_done, 1));
    ONI_TRY(dev_alloc(&c->d_nan, 1));
    ONI_CUDA(cudaMemset(c->d_done, 0, sizeof(unsigned)));
    ONI_CUDA(cudaMemset(c->d_nan, 0, sizeof(int32_t)));
    ONI_CUDA(cudaMemset(c->d_ss, 0, sizeof(double) * c->ss_count()));
    ONI_CUDA(cudaMallocHost((void **)&c->h_pinned, sizeof(double) * 8));
    return ONI_OK;
}

extern "C" void oni_lda_destroy(oni_lda_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    free_corpus(c);
    cudaFree(c->d_logbeta_t); cudaFree(c->d_alpha); cudaFree(c->d_ss); cudaFree(c->d_block_partial);
    cudaFree(c->d_ticket); cudaFree(c->d_done); cudaFree(c->d_nan); cudaFree(c->d_long_tabs);
    if (c->h_pinned) cudaFreeHost(c->h_pinned);
    for (auto &e : c->ev) if (e) cudaEventDestroy(e);
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    if (c->stream_long) cudaStreamDestroy(c->stream_long);
    if (c->stream_excl) cudaStreamDestroy(c->stream_excl);
    if (c->ev_join_excl) cudaEventDestroy(c->ev_join_excl);
    if (c->ev_t_long) cudaEventDestroy(c->ev_t_long);
    if (c->ev_t_excl) cudaEventDestroy(c->ev_t_excl);
    if (c->ev_t_main) cudaEventDestroy(c->ev_t_main);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" int oni_lda_set_corpus(oni_lda_ctx *c, int32_t D, int64_t nnz, const int64_t *ptr,
                                  const int32_t *word, const int32_t *count)
{
    if (!c || !ptr || (nnz > 0 && (!word || !count)) || D < 0 || nnz < 0)
        return oni_set_error(ONI_ERR_ARG, "set_corpus: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    c->estep_done = false;
    // stage the CSR arrays on the device, index there
    StageTimer tm;
    int64_t *d_ptr = nullptr;
    int32_t *d_word = nullptr, *d_count = nullptr;
    int rc = dev_alloc(&d_ptr, (size_t)D + 1);
    if (rc == ONI_OK) rc = dev_alloc(&d_word, (size_t)nnz);
    if (rc == ONI_OK) rc = dev_alloc(&d_count, (size_t)nnz);
    if (rc == ONI_OK) {
        cudaError_t e = cudaMemcpyAsync(d_ptr, ptr, sizeof(int64_t) * ((size_t)D + 1), cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(d_word, word, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess && nnz) e = cudaMemcpyAsync(d_count, count, sizeof(int32_t) * (size_t)nnz, cudaMemcpyHostToDevice, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) rc = oni_set_error(ONI_ERR_CUDA, "set_corpus upload: %s", cudaGetErrorString(e));
    }
    tm.lap("corpus upload");
    if (rc == ONI_OK) rc = build_indices(c, D, nnz, ptr, d_ptr, d_word, d_count);
    cudaFree(d_ptr); cudaFree(d_word); cudaFree(d_count);
    return rc;
}

extern "C" int oni_lda_set_corpus_device(oni_lda_ctx *c, int32_t D, int64_t nnz,
                                         const int64_t *d_ptr, const int32_t *d_word,
                                         const int32_t *d_count)
{
    if (!c || !d_ptr || D < 0 || nnz < 0 || (nnz > 0 && (!d_word || !d_count)))
        return oni_set_error(ONI_ERR_ARG, "set_corpus_device: bad argument");
    ONI_CUDA(cudaSetDevice(c->device));
    std::vector<int64_t> ptr((size_t)D + 1);
    ONI_CUDA(cudaMemcpy(ptr.data(), d_ptr, sizeof(int64_t) * ((size_t)D + 1), cudaMemcpyDeviceToHost));
    c->estep_done = false;
    return build_indices(c, D, nnz, ptr.data(), d_ptr, d_word, d_count);
}

// Upload a topic-major host model into the word-major device table.
static int upload_model(oni_lda_ctx *c, const double *logbeta, double alpha)
{
    // the [K][V] -> [V][KP] layout change happens on the device (a strided host loop over K*V
    // doubles took 100 ms at K = 50, V = 200k)
    {
        // exp(digamma(gamma) + log beta) must be a normal number for at least one topic of every
        // word, or the unshifted E-step normalises garbage (the reference's log_sum chain,
        // lda-inference.c:53-64, has no such limit): remember how low this model goes
        const size_t V = (size_t)c->V;
        std::vector<double> rowmax(V, -INFINITY);
        double lo = INFINITY;
        for (int k = 0; k < c->K; k++) {
            const double *row = logbeta + (size_t)k * V;
            for (size_t w = 0; w < V; w++) {
                const double x = row[w];
                if (x != x) return oni_set_error(ONI_ERR_ARG, "model: log beta[%d][%zu] is NaN", k, w);
                rowmax[w] = x > rowmax[w] ? x : rowmax[w];
                lo = x < lo ? x : lo;
            }
        }
        c->model_min = lo;
        c->model_min_rowmax = *std::min_element(rowmax.begin(), rowmax.end());
    }
    double *stage = nullptr;
    ONI_TRY(dev_alloc(&stage, (size_t)c->K * c->V));
    cudaError_t e = cudaMemcpyAsync(stage, logbeta, sizeof(double) * (size_t)c->K * c->V, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) {
        const dim3 grid((c->V + 31) / 32, (c->KP + 31) / 32), block(32, 8);
        model_to_word_major_kernel<<<grid, block, 0, c->stream>>>(stage, c->K, c->V, c->KP, c->d_logbeta_t);
        c->launches++;
        e = cudaMemcpyAsync(c->d_alpha, &alpha, sizeof(double), cudaMemcpyHostToDevice, c->stream);
    }
    if (e == cudaSuccess) e = cudaMemsetAsync(c->d_nan, 0, sizeof(int32_t), c->stream);  // a new model, a clean slate
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    cudaFree(stage);
    if (e != cudaSuccess) return oni_set_error(ONI_ERR_CUDA, "model upload: %s", cudaGetErrorString(e));
    c->alpha = alpha;
    c->have_model = true;
    c->em_iter = 0; c->lik_old = 0.0; c->converged = 1.0;
    c->var_max_iter = c->st.var_max_iter;
    return ONI_OK;
}

// random_initialize_ss + lda_mle(model, ss, 0) on the host with the reference's
// stream and libm, so 000.beta is byte-identical (lda-model.c:104-117,27-44).
void oni_internal_random_model(int K, int V, uint32_t seed, std::vector<double> &logbeta)
{
    CokusMT mt(seed);
    logbeta.assign((size_t)K * V, 0.0);
    std::vector<double> lct((size_t)K);
    // the draws and the running totals are one sequential stream (k outer, n inner) ...
    for (int k = 0; k < K; k++) {
        double ct = 0.0;
        double *cw = &logbeta[(size_t)k * V];
        for (int n = 0; n < V; n++) {
            cw[n] = 1.0 / V + mt.uniform();
            ct += cw[n];
        }
        lct[k] = log(ct);
    }
    // ... the K*V logarithms after them are independent: topics over host threads (same libm)
    const int T = std::max(1, std::min<int>({K, (int)std::thread::hardware_concurrency(), 16}));
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++)
        th.emplace_back([&, t] {
            for (int k = t; k < K; k += T) {
                double *cw = &logbeta[(size_t)k * V];
                for (int n = 0; n < V; n++) cw[n] = (cw[n] > 0) ? log(cw[n]) - lct[k] : -100.0;
            }
        });
    for (auto &x : th) x.join();
}

extern "C" int oni_lda_init_random(oni_lda_ctx *c, uint32_t seed, double alpha)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "init_random: null context");
    ONI_CUDA(cudaSetDevice(c->device));
    std::vector<double> lb;
    oni_internal_random_model(c->K, c->V, seed, lb);
    return upload_model(c, lb.data(), alpha);
}

extern "C" int oni_lda_set_model(oni_lda_ctx *c, const double *logbeta, double alpha)
{
    if (!c || !logbeta) return oni_set_error(ONI_ERR_ARG, "set_model: null argument");
    ONI_CUDA(cudaSetDevice(c->device));
    return upload_model(c, logbeta, alpha);
}

// ------------------------------------------------------------------------ steps
extern "C" int oni_lda_e_step(oni_lda_ctx *c)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "e_step: null context");
    if (!c->have_model) return oni_set_error(ONI_ERR_ARG, "e_step: no model (init_random / set_model first)");
    if (!c->d_doc_ptr) return oni_set_error(ONI_ERR_ARG, "e_step: no corpus");
    ONI_CUDA(cudaSetDevice(c->device));
    cudaStream_t s = c->stream;

    EstepParams p{};
    p.tok = c->d_tok; p.doc_ptr = c->d_doc_ptr; p.order = c->d_order; p.doc_total = c->d_doc_total;
    p.logbeta_t = c->d_logbeta_t; p.phi = c->d_phi; p.gamma = c->d_gamma; p.doc_lik = c->d_doc_lik;
    p.doc_ass = c->d_doc_ass; p.doc_iters = c->d_doc_iters; p.alpha = c->d_alpha;
    p.ticket = c->d_ticket; p.nan_flag = c->d_nan;
    p.D = c->D; p.K = c->K; p.KP = c->KP;
    p.var_max_iter = c->var_max_iter;
    p.var_converged = (double)c->st.var_converged;  // float promoted, lda-inference.c:46
    p.psi_tab = c->d_long_tabs;
    p.exk_tab = c->d_long_tabs ? c->d_long_tabs + (size_t)PSI_N * PSI_COEF : nullptr;

    // exp(x) cannot underflow for every topic of a word unless digamma(alpha) is hugely
    // negative (alpha tiny) or the model itself is (a loaded model may hold a word whose every
    // topic sits below -600); then use the max-shifted kernel.  gamma >= alpha, so
    // digamma(alpha) + the smallest per-word maximum bounds the best exponent of any visit.
    const bool shift = !(c->alpha > 0.01) ||
                       !(digamma_ref(c->alpha, g_host_tables.lg) + c->model_min_rowmax > -600.0);
    // the short-chain kernel's exp takes arguments down to -7e5 and its offsets are 32-bit
    const bool use_long = c->long_kernel && !shift && c->model_min > -1.0e5 &&
                          (double)c->max_doc_pairs * c->KP < 4.0e9;

    ONI_CUDA(cudaMemsetAsync(c->d_ticket, 0, 4 * sizeof(unsigned), s));
    ONI_CUDA(cudaEventRecord(c->ev[0], s));
    if (c->n_long > 0) ONI_CUDA(cudaEventRecord(c->ev_fork, s));
    if (c->n_excl > 0) {
        // the very longest documents before anything else, on SMs nothing else will share
        ONI_CUDA(cudaStreamWaitEvent(c->stream_excl, c->ev_fork, 0));
        EstepParams px = p;
        px.D = c->n_excl;
        px.ticket = c->d_ticket + 2;
        cudaError_t e = use_long
            ? launch_estep_long(c->K, px, c->sm_count, c->stream_excl, /*exclusive_blocks=*/(c->n_excl + 3) / 4)
            : launch_estep(c->G_long, c->T_long, shift, px, c->sm_count, c->stream_excl, /*longdoc=*/true,
                           /*exclusive_blocks=*/(c->n_excl + 3) / 4);
        if (e != cudaSuccess)
            return oni_set_error(ONI_ERR_CUDA, "E-step launch (exclusive, T=%d): %s", c->T_long, cudaGetErrorString(e));
        c->launches++;
        ONI_CUDA(cudaEventRecord(c->ev_join_excl, c->stream_excl));
        ONI_CUDA(cudaEventRecord(c->ev_t_excl, c->stream_excl));
    }
    if (c->n_long > c->n_excl) {
        // the other long documents next, one per warp, on a second stream; their blocks (four
        // warps) share SMs with the main launch's, so the rest of the corpus fills in around them
        ONI_CUDA(cudaStreamWaitEvent(c->stream_long, c->ev_fork, 0));
        EstepParams pl = p;
        pl.order = c->d_order + c->n_excl;
        pl.D = c->n_long - c->n_excl;
        pl.ticket = c->d_ticket + 1;
        cudaError_t e = use_long
            ? launch_estep_long(c->K, pl, c->sm_count, c->stream_long)
            : launch_estep(c->G_long, c->T_long, shift, pl, c->sm_count, c->stream_long, /*longdoc=*/true);
        if (e != cudaSuccess)
            return oni_set_error(ONI_ERR_CUDA, "E-step launch (G=%d,T=%d): %s", c->G_long, c->T_long, cudaGetErrorString(e));
        c->launches++;
        ONI_CUDA(cudaEventRecord(c->ev_join, c->stream_long));
        ONI_CUDA(cudaEventRecord(c->ev_t_long, c->stream_long));
    }
    if (c->D - c->n_small > c->n_long) {
        EstepParams ps = p;
        ps.order = c->d_order + c->n_long;
        ps.D = c->D - c->n_long - c->n_small;
        cudaError_t e = launch_estep(c->G, c->T, shift, ps, c->sm_count, s);
        if (e != cudaSuccess)
            return oni_set_error(ONI_ERR_CUDA, "E-step launch (G=%d,T=%d): %s", c->G, c->T, cudaGetErrorString(e));
        c->launches++;
    }
    if (c->n_small > 0) {
        // documents of one pair (and empty ones) last, through the build that knows a document's only
        // row is both the one it writes and the next one it reads
        EstepParams pt = p;
        pt.order = c->d_order + (c->D - c->n_small);
        pt.D = c->n_small;
        pt.ticket = c->d_ticket + 3;
        pt.no_exact = 1;
        cudaError_t e = launch_estep(c->G, c->T, shift, pt, c->sm_count, s);
        if (e != cudaSuccess)
            return oni_set_error(ONI_ERR_CUDA, "E-step launch (one-pair documents): %s", cudaGetErrorString(e));
        c->launches++;
    }
    ONI_CUDA(cudaEventRecord(c->ev_t_main, s));
    if (c->n_long > c->n_excl) ONI_CUDA(cudaStreamWaitEvent(s, c->ev_join, 0));
    if (c->n_excl > 0) ONI_CUDA(cudaStreamWaitEvent(s, c->ev_join_excl, 0));
    ONI_CUDA(cudaEventRecord(c->ev[1], s));
    ONI_TRY(dispatch_ss(c));
    {
        int S = 1;
        while (S * 2 * c->KP <= 1024) S *= 2;
        class_total_kernel<<<1, S * c->KP, sizeof(double) * S * c->KP, s>>>(c->d_ss, c->V, c->K, c->KP,
                                                                           c->ss_class_total());
        doc_reduce_kernel<<<DOC_RED_BLOCKS, DOC_RED_THREADS, 0, s>>>(
            c->d_doc_lik, c->d_doc_ass, c->d_doc_iters, c->d_doc_ptr, c->D, c->d_block_partial, c->d_done,
            c->ss_tail());
        c->launches += 2;
    }
    ONI_CUDA(cudaEventRecord(c->ev[2], s));
    // this rank's own visit count, before an all-reduce sums the tail over the ranks
    ONI_CUDA(cudaMemcpyAsync(c->h_pinned + 6, c->ss_tail() + 3, sizeof(double), cudaMemcpyDeviceToHost, s));
    ONI_CUDA(cudaGetLastError());
    c->estep_done = true;
    return ONI_OK;
}

extern "C" int oni_lda_suffstats(oni_lda_ctx *c, double **d_buffer, size_t *count)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "suffstats: null context");
    if (d_buffer) *d_buffer = c->d_ss;
    if (count) *count = c->ss_count();
    return ONI_OK;
}

extern "C" int oni_lda_comm_unique_id(unsigned char id[ONI_NCCL_ID_BYTES])
{
    ONI_TRY(nccl_load());
    ncclUniqueId u;
    ONI_NCCL(g_nccl.GetUniqueId(&u));
    memcpy(id, u.internal, ONI_NCCL_ID_BYTES);
    return ONI_OK;
}

extern "C" int oni_lda_comm_init(oni_lda_ctx *c, const unsigned char id[ONI_NCCL_ID_BYTES],
                                 int32_t rank, int32_t nranks)
{
    if (!c || !id || nranks < 1 || rank < 0 || rank >= nranks)
        return oni_set_error(ONI_ERR_ARG, "comm_init: bad argument");
    ONI_TRY(nccl_load());
    ONI_CUDA(cudaSetDevice(c->device));
    ncclUniqueId u;
    memcpy(u.internal, id, ONI_NCCL_ID_BYTES);
    ONI_NCCL(g_nccl.CommInitRank(&c->comm, nranks, u, rank));
    c->nranks = nranks;
    return ONI_OK;
}

extern "C" int oni_lda_allreduce(oni_lda_ctx *c)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "allreduce: null context");
    if (!c->comm || c->nranks == 1) return ONI_OK;
    ONI_CUDA(cudaSetDevice(c->device));
    // one collective for what the reference does with 2K+3 (lda-estimate.c:211-252)
    ONI_NCCL(g_nccl.AllReduce(c->d_ss, c->d_ss, c->ss_count(), kNcclFloat64, kNcclSum, c->comm, c->stream));
    return ONI_OK;
}

// M-step kernels only (asynchronous); finish_m_step reads the scalars back.
static int launch_m_step(oni_lda_ctx *c)
{
    cudaStream_t s = c->stream;
    ONI_CUDA(cudaEventRecord(c->ev[3], s));
    const size_t n = (size_t)c->V * c->KP;
    const int blocks = (int)std::min<size_t>((n + 255) / 256, (size_t)c->sm_count * 8);
    mle_beta_kernel<<<blocks, 256, 0, s>>>(c->d_ss, c->ss_class_total(), c->V, c->K, c->KP, c->d_logbeta_t);
    c->launches++;
    if (c->st.estimate_alpha == 1) {  // lda-model.c:46-53
        alpha_newton_kernel<<<1, 32, 0, s>>>(c->ss_tail(), c->K, c->d_alpha);
        c->launches++;
    }
    ONI_CUDA(cudaEventRecord(c->ev[4], s));
    ONI_CUDA(cudaMemcpyAsync(c->h_pinned, c->ss_tail(), sizeof(double) * 4, cudaMemcpyDeviceToHost, s));
    ONI_CUDA(cudaMemcpyAsync(c->h_pinned + 4, c->d_alpha, sizeof(double), cudaMemcpyDeviceToHost, s));
    ONI_CUDA(cudaMemcpyAsync(c->h_pinned + 5, c->d_nan, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    return ONI_OK;
}

// After an EM iteration the number of word visits is known, and with it how long the bulk of the
// next E-step will take (~0.107 ns per visit at K = 20 with the GPU full); the tier thresholds of
// build_indices, which had to guess it from the pair count, are set again from it: a document's
// chain in the main launch (~1.26 us per visit of one document, var_max_iter sweeps at worst) and
// in the co-resident long launch (~0.39 us) should both end at ~0.9 of the bulk.  A corpus whose
// documents converge in few sweeps (DNS day) gets its long tier from ~570 pairs instead of 1024 on
// four GPUs.  The visit count is a property of the documents and the model, not of timing, so a
// run stays reproducible; the tiers are prefixes of the same length-sorted order.
static void retier_from_visits(oni_lda_ctx *c)
{
    if (c->docs_ge.empty() || c->stats.word_visits <= 0) return;
    const double visits = (double)c->stats.word_visits;  // (this rank's shard)
    const double sweeps = c->var_max_iter > 0 ? std::min(c->var_max_iter, 1000) : 20;
    const double bulk_us = visits * 0.107e-3 * (c->K / 20.0);
    const int64_t long_min = std::min<int64_t>(1024, std::max<int64_t>(128, (int64_t)(0.9 * bulk_us / (sweeps * 1.26))));
    const int64_t excl_min = std::min<int64_t>(TIER_LEN_CAP, std::max<int64_t>(long_min, (int64_t)(0.9 * bulk_us / (sweeps * 0.39))));
    const int32_t n_long = c->docs_ge[long_min];
    const int32_t n_excl = std::min(std::min(c->docs_ge[excl_min], n_long), 4 * std::max(1, c->sm_count / 8));
    c->n_long = n_long;
    c->n_excl = n_excl;
    c->n_small = std::min<int32_t>(c->n_small, c->D - n_long);
}

static int finish_m_step(oni_lda_ctx *c, double *likelihood, double *converged)
{
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    int32_t nan_flag;
    memcpy(&nan_flag, c->h_pinned + 5, sizeof nan_flag);
    const double lik = c->h_pinned[1];
    c->alpha = c->h_pinned[4];
    c->model_min = c->model_min_rowmax = -100.0;  // the table is lda_mle's now (floor: lda-model.c:42)
    c->stats.word_visits = (int64_t)c->h_pinned[6];
    float ms = 0;
    if (cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]) == cudaSuccess) c->stats.ms_estep_kernel = ms;
    {
        float a = -1, b = -1, m = -1;
        if (c->n_excl > 0) cudaEventElapsedTime(&a, c->ev[0], c->ev_t_excl);
        if (c->n_long > c->n_excl) cudaEventElapsedTime(&b, c->ev[0], c->ev_t_long);
        if (c->D > c->n_long) cudaEventElapsedTime(&m, c->ev[0], c->ev_t_main);
        c->stats.ms_end_exclusive = a; c->stats.ms_end_long = b; c->stats.ms_end_main = m;
        c->stats.docs_exclusive = c->n_excl; c->stats.docs_long = c->n_long - c->n_excl;
        c->stats.longest_doc_pairs = c->max_doc_pairs;
        if (getenv("ONI_LDA_TIMING")) fprintf(stderr, "[oni lda] E-step %.2f ms: exclusive (%d docs) done at %.2f, long (%d docs) at %.2f, rest at %.2f\n",
                c->stats.ms_estep_kernel, c->n_excl, a, c->n_long - c->n_excl, b, m);
    }
    if (cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]) == cudaSuccess) c->stats.ms_suffstats = ms;
    if (cudaEventElapsedTime(&ms, c->ev[3], c->ev[4]) == cudaSuccess) c->stats.ms_mstep = ms;
    c->stats.kernel_launches = c->launches;
    if (nan_flag == 2) {  // (ONI_RING_CHECK builds only)
        cudaMemsetAsync(c->d_nan, 0, sizeof(int32_t), c->stream);
        return oni_set_error(ONI_ERR_CUDA, "ring self-check: a word visit of the long-document kernel read something else than plain loads of the same words");
    }
    if (nan_flag || lik != lik) {
        cudaMemsetAsync(c->d_nan, 0, sizeof(int32_t), c->stream);  // reported once; the handle stays usable
        return oni_set_error(ONI_ERR_NAN, "a document bound is NaN (EM iteration %d)", c->em_iter + 1);
    }
    // lda-estimate.c:260-262
    c->em_iter++;
    c->converged = (c->lik_old - lik) / c->lik_old;
    // lda-estimate.c:261 doubles on every decrease of the bound, -1 included (-1 -> -2 = zero sweeps
    // from then on) and without an overflow guard; here "no cap" stays "no cap" and the doubling
    // stops at 2^30
    if (c->converged < 0 && c->var_max_iter > 0 && c->var_max_iter < (1 << 30)) c->var_max_iter *= 2;
    retier_from_visits(c);
    c->lik_old = lik;
    c->estep_done = false;
    if (likelihood) *likelihood = lik;
    if (converged) *converged = c->converged;
    return ONI_OK;
}

extern "C" int oni_lda_m_step(oni_lda_ctx *c, double *likelihood, double *converged)
{
    if (!c) return oni_set_error(ONI_ERR_ARG, "m_step: null context");
    if (!c->estep_done) return oni_set_error(ONI_ERR_ARG, "m_step: no E-step since the last M-step");
    ONI_CUDA(cudaSetDevice(c->device));
    ONI_TRY(launch_m_step(c));
    return finish_m_step(c, likelihood, converged);
}

extern "C" int oni_lda_em_iteration(oni_lda_ctx *c, double *likelihood, double *converged)
{
    ONI_TRY(oni_lda_e_step(c));
    ONI_TRY(oni_lda_allreduce(c));
    return oni_lda_m_step(c, likelihood, converged);
}

extern "C" int oni_lda_em_should_continue(const oni_lda_ctx *c)
{
    if (!c) return 0;
    // lda-estimate.c:186
    const double em_conv = (double)c->st.em_converged;
    return (((c->converged < 0) || (c->converged > em_conv) || (c->em_iter <= 2)) &&
            (c->em_iter <= c->st.em_max_iter)) ? 1 : 0;
}

// --------------------------------------------------------------------- readback
extern "C" int oni_lda_get_logbeta(oni_lda_ctx *c, double *logbeta)
{
    if (!c || !logbeta) return oni_set_error(ONI_ERR_ARG, "get_logbeta: null argument");
    ONI_CUDA(cudaSetDevice(c->device));
    double *stage = nullptr;
    ONI_TRY(dev_alloc(&stage, (size_t)c->K * c->V));
    const dim3 grid((c->V + 31) / 32, (c->K + 31) / 32), block(32, 8);
    model_to_topic_major_kernel<<<grid, block, 0, c->stream>>>(c->d_logbeta_t, c->K, c->V, c->KP, stage);
    c->launches++;
    cudaError_t e = cudaMemcpyAsync(logbeta, stage, sizeof(double) * (size_t)c->K * c->V, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    cudaFree(stage);
    if (e != cudaSuccess) return oni_set_error(ONI_ERR_CUDA, "model download: %s", cudaGetErrorString(e));
    return ONI_OK;
}

extern "C" int oni_lda_get_gamma(oni_lda_ctx *c, double *gamma)
{
    if (!c || !gamma) return oni_set_error(ONI_ERR_ARG, "get_gamma: null argument");
    ONI_CUDA(cudaSetDevice(c->device));
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    ONI_CUDA(cudaMemcpy(gamma, c->d_gamma, sizeof(double) * (size_t)c->D * c->K, cudaMemcpyDeviceToHost));
    return ONI_OK;
}

extern "C" int oni_lda_get_alpha(oni_lda_ctx *c, double *alpha)
{
    if (!c || !alpha) return oni_set_error(ONI_ERR_ARG, "get_alpha: null argument");
    *alpha = c->alpha;
    return ONI_OK;
}

extern "C" int oni_lda_get_doc_likelihood(oni_lda_ctx *c, double *lik)
{
    if (!c || !lik) return oni_set_error(ONI_ERR_ARG, "get_doc_likelihood: null argument");
    ONI_CUDA(cudaSetDevice(c->device));
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    ONI_CUDA(cudaMemcpy(lik, c->d_doc_lik, sizeof(double) * (size_t)c->D, cudaMemcpyDeviceToHost));
    return ONI_OK;
}

extern "C" int oni_lda_get_doc_iters(oni_lda_ctx *c, int32_t *iters)
{
    if (!c || !iters) return oni_set_error(ONI_ERR_ARG, "get_doc_iters: null argument");
    ONI_CUDA(cudaSetDevice(c->device));
    ONI_CUDA(cudaStreamSynchronize(c->stream));
    ONI_CUDA(cudaMemcpy(iters, c->d_doc_iters, sizeof(int32_t) * (size_t)c->D, cudaMemcpyDeviceToHost));
    return ONI_OK;
}

extern "C" int oni_lda_get_stats(oni_lda_ctx *c, oni_lda_stats *out)
{
    if (!c || !out) return oni_set_error(ONI_ERR_ARG, "get_stats: null argument");
    c->stats.kernel_launches = c->launches;
    *out = c->stats;
    return ONI_OK;
}

extern "C" int oni_lda_stream(oni_lda_ctx *c, void **cuda_stream)
{
    if (!c || !cuda_stream) return oni_set_error(ONI_ERR_ARG, "stream: null argument");
    *cuda_stream = (void *)c->stream;
    return ONI_OK;
}

// ---------------------------------------------------------- whole-run, in memory
// Contiguous shards balanced by pair count: cut[r]..cut[r+1].
static void shard_docs(const oni_lda_corpus *c, int n, std::vector<int32_t> &cut)
{
    cut.assign((size_t)n + 1, 0);
    cut[n] = c->num_docs;
    for (int r = 1; r < n; r++) {
        const int64_t target = c->nnz * r / n;
        const int64_t *it = std::lower_bound(c->doc_ptr, c->doc_ptr + c->num_docs + 1, target);
        cut[r] = std::max<int32_t>(cut[r - 1], (int32_t)std::min<int64_t>(it - c->doc_ptr, c->num_docs));
    }
}

int oni_internal_estimate(const oni_lda_corpus *corpus, int32_t K, double initial_alpha,
                          const oni_lda_settings *st, const double *logbeta0, int32_t ngpus,
                          oni_lda_result *res, oni_iter_callback on_iter, void *user)
{
    if (!corpus || !st || !res) return oni_set_error(ONI_ERR_ARG, "estimate: null argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return oni_set_error(ONI_ERR_CUDA, "no CUDA device: this engine has no CPU path");
    if (ngpus < 1) ngpus = 1;
    if (ngpus > ndev)
        return oni_set_error(ONI_ERR_ARG, "%d GPUs requested, %d visible", ngpus, ndev);
    const int V = corpus->num_terms, D = corpus->num_docs;

    std::vector<oni_lda_ctx *> ctx((size_t)ngpus, nullptr);
    std::vector<int32_t> cut;
    shard_docs(corpus, ngpus, cut);
    int rc = ONI_OK;
    auto cleanup = [&]() { for (auto *c : ctx) oni_lda_destroy(c); };
#define ONI_EST(call) do { rc = (call); if (rc != ONI_OK) { cleanup(); return rc; } } while (0)

    if (res->logbeta_init) memcpy(res->logbeta_init, logbeta0, sizeof(double) * (size_t)K * V);

    for (int r = 0; r < ngpus; r++) {
        ONI_EST(oni_lda_create(r, K, V, st, &ctx[r]));
        const int32_t lo = cut[r], hi = cut[r + 1];
        const int64_t a = corpus->doc_ptr[lo], b = corpus->doc_ptr[hi];
        std::vector<int64_t> ptr((size_t)(hi - lo) + 1);
        for (int32_t d = lo; d <= hi; d++) ptr[d - lo] = corpus->doc_ptr[d] - a;
        ONI_EST(oni_lda_set_corpus(ctx[r], hi - lo, b - a, ptr.data(), corpus->word + a, corpus->count + a));
        ONI_EST(upload_model(ctx[r], logbeta0, initial_alpha));
    }
    if (ngpus > 1) {
        rc = nccl_load();
        if (rc != ONI_OK) { cleanup(); return rc; }
        std::vector<ncclComm_t> comms((size_t)ngpus);
        std::vector<int> devs((size_t)ngpus);
        std::iota(devs.begin(), devs.end(), 0);
        int nr = g_nccl.CommInitAll(comms.data(), ngpus, devs.data());
        if (nr != 0) { cleanup(); return oni_set_error(ONI_ERR_NCCL, "ncclCommInitAll: %s", g_nccl.GetErrorString(nr)); }
        for (int r = 0; r < ngpus; r++) { ctx[r]->comm = comms[r]; ctx[r]->nranks = ngpus; }
    }

    const double t0 = now_seconds();
    int it = 0;
    while (oni_lda_em_should_continue(ctx[0])) {
        for (int r = 0; r < ngpus; r++) ONI_EST(oni_lda_e_step(ctx[r]));
        if (ngpus > 1) {
            g_nccl.GroupStart();
            for (int r = 0; r < ngpus; r++) {
                rc = oni_lda_allreduce(ctx[r]);
                if (rc != ONI_OK) break;
            }
            g_nccl.GroupEnd();
            if (rc != ONI_OK) { cleanup(); return rc; }
        }
        for (int r = 0; r < ngpus; r++) {
            rc = cudaSetDevice(ctx[r]->device) == cudaSuccess ? launch_m_step(ctx[r]) : ONI_ERR_CUDA;
            if (rc != ONI_OK) { cleanup(); return rc; }
        }
        double lik = 0, conv = 0;
        for (int r = 0; r < ngpus; r++) {
            cudaSetDevice(ctx[r]->device);
            ONI_EST(finish_m_step(ctx[r], &lik, &conv));
        }
        if (it < res->likelihood_cap) {
            if (res->likelihood) res->likelihood[it] = lik;
            if (res->converged) res->converged[it] = conv;
        }
        it++;
        if (on_iter) on_iter(user, it, lik, conv);
    }
    res->seconds_em = now_seconds() - t0;
    res->em_iters = it;
    res->alpha = ctx[0]->alpha;
    if (res->logbeta) ONI_EST(oni_lda_get_logbeta(ctx[0], res->logbeta));
    if (res->gamma)
        for (int r = 0; r < ngpus; r++)
            ONI_EST(oni_lda_get_gamma(ctx[r], res->gamma + (size_t)cut[r] * K));
    (void)D;
    cleanup();
#undef ONI_EST
    return ONI_OK;
}

extern "C" int oni_lda_estimate(const oni_lda_corpus *corpus, int32_t K, double initial_alpha,
                                const oni_lda_settings *st, uint32_t seed, int32_t ngpus,
                                oni_lda_result *res)
{
    if (!corpus || !st || !res) return oni_set_error(ONI_ERR_ARG, "estimate: null argument");
    if (K < 1 || corpus->num_terms < 1) return oni_set_error(ONI_ERR_ARG, "estimate: empty model");
    std::vector<double> lb0;
    oni_internal_random_model(K, corpus->num_terms, seed, lb0);  // lda-estimate.c:155-161
    return oni_internal_estimate(corpus, K, initial_alpha, st, lb0.data(), ngpus, res, nullptr, nullptr);
}
