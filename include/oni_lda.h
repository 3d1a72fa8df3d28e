/*
 * oni_lda.h -- C ABI of the B200-native variational-EM LDA engine and event
 * scorer that replaces the oni-lda-c step of rabarona/oni-ml.
 *
 * The reference has no in-process interface for this path: its "plugin API" is
 * a process command line plus files (ml_ops.sh:79-80 launches
 * `mpiexec ... ./lda est 2.5 20 settings.txt 20 model.dat random <dir>`), and
 * the scoring loop is a spark-shell script (ml_ops.sh:108).  This header is the
 * FFI a maintainer of that pipeline would bind instead (see INTEGRATION.md for
 * the ml_ops.sh edit, a JNI/JNA stub for the Scala stage and a ctypes stub).
 * Each entry point names the reference code it stands in for; paths are
 * relative to the reference repository root.
 *
 * Conventions
 *   - plain C: pointers, sizes, PODs; no C++ or framework types cross the ABI;
 *   - every function returns 0 on success and a non-zero oni_status otherwise;
 *     oni_last_error() then returns a message (thread-local).  The reference
 *     never reports failure (main always returns 0, lda-estimate.c:467), so any
 *     caller that ignored the status keeps working;
 *   - "host" pointers are ordinary process memory, "device" pointers are CUDA
 *     global memory on the context's device;
 *   - matrices are row-major doubles.  Topic-major [K][V] is the layout of the
 *     reference's log_prob_w / class_word (lda.h:40-57) and of final.beta;
 *   - there is no CPU fallback: without a CUDA device the compute entry points
 *     fail with ONI_ERR_CUDA.
 */
#ifndef ONI_LDA_H
#define ONI_LDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    ONI_OK = 0,
    ONI_ERR_ARG = 1,    /* bad argument / size                               */
    ONI_ERR_IO = 2,     /* file missing, unreadable, malformed, unwritable   */
    ONI_ERR_CUDA = 3,   /* CUDA runtime error, or no device                  */
    ONI_ERR_NCCL = 4,   /* NCCL not loadable or a collective failed          */
    ONI_ERR_NAN = 5,    /* a document bound became NaN (the reference aborts
                           on assert(!isnan), lda-inference.c:78)            */
    ONI_ERR_NOMEM = 6
} oni_status;

/* Message for the last failure on the calling thread ("" if none). */
const char *oni_last_error(void);

/* Version string of the library ("oni-lda-b200 x.y"). */
const char *oni_version(void);

/* ------------------------------------------------------------------------ */
/* settings.txt  (read_settings, oni-lda-c/lda-estimate.c:331-350)           */
/* The two thresholds are floats in the reference (lda-inference.h:10,       */
/* lda-estimate.h:21) and are compared after promotion to double.            */
typedef struct {
    int32_t var_max_iter;   /* "var max iter %d"      (-1 = no cap)           */
    float var_converged;    /* "var convergence %f"                           */
    int32_t em_max_iter;    /* "em max iter %d"                               */
    float em_converged;     /* "em convergence %f"                            */
    int32_t estimate_alpha; /* "alpha %s": 0 for the literal "fixed", else 1  */
} oni_lda_settings;

int oni_lda_read_settings(const char *path, oni_lda_settings *out);

/* ------------------------------------------------------------------------ */
/* Corpus in host memory, CSR.  Replaces document/corpus (lda.h:23-37) and    */
/* read_data (lda-data.c:22-59).  Duplicate word ids inside a document are    */
/* separate pairs.  num_terms = largest word id + 1 (lda-data.c:49).          */
typedef struct {
    int32_t num_docs;
    int32_t num_terms;
    int64_t nnz;
    const int64_t *doc_ptr; /* [num_docs+1]                                  */
    const int32_t *word;    /* [nnz]                                         */
    const int32_t *count;   /* [nnz]                                         */
} oni_lda_corpus;

/* Parse model.dat ("N w:c w:c ..." per line).  The arrays are malloc'ed by the
 * library; release them with oni_lda_free_corpus. */
int oni_lda_read_corpus(const char *model_dat_path, oni_lda_corpus *out);
void oni_lda_free_corpus(oni_lda_corpus *c);
/* The same text written from CSR arrays: one line per document, "N w:c w:c ...", as
 * lda_pre.py:94 writes it and read_data (lda-data.c:35-51) parses it. */
int oni_lda_write_corpus(const char *model_dat_path, const oni_lda_corpus *c);

/* The corpus indexing stage, lda_pre.py (ml_ops.sh:67): reads <dir>doc_wc.dat
 * ("ip,word,count" rows) and writes <dir>words.dat, <dir>doc.dat and <dir>model.dat
 * byte for byte as the script does (word ids from 0 and document ids from 1 in order
 * of first appearance, lda_pre.py:33-44,57-73; a word repeated inside a document stays
 * a separate pair, :66; model.dat line = pair count then " w:c"..., :94).  `dir` is a
 * prefix, used like the script's argv[1] ("path/"). */
int oni_lda_pre(const char *dir);

/* ------------------------------------------------------------------------ */
/* Whole-run entry points                                                    */

/* The reference's process entry (main, lda-estimate.c:403-468) with the same
 * argv:   lda est <alpha> <k> <settings> <ngpus> <model.dat> <random|seeded|root> <outdir>
 *         lda inf <settings> <model_root> <model.dat> <out_name>
 * argv[5] ("#processes", the MPI stride of lda-estimate.c:201,428) is read as
 * the number of GPUs to shard documents over.  The RNG seed is the wall clock
 * as in the reference (lda-estimate.c:414-416) unless the environment variable
 * ONI_LDA_SEED is set.  Writes <outdir>/000.beta, 000.other, likelihood.dat,
 * final.beta, final.other, final.gamma (formats of lda-model.c:193-217,
 * lda-estimate.c:88-104,266).  Returns 0 or an oni_status. */
int oni_lda_main(int argc, char **argv);

/* Result of an in-memory estimation; all arrays are caller-allocated host
 * memory (any pointer may be NULL to skip that output). */
typedef struct {
    double *logbeta_init;  /* [K][V]  model before EM (000.beta)             */
    double *logbeta;       /* [K][V]  final.beta                             */
    double *gamma;         /* [D][K]  final.gamma                            */
    double *likelihood;    /* [likelihood_cap] bound after each EM iteration */
    double *converged;     /* [likelihood_cap] relative change (first = inf) */
    int32_t likelihood_cap;
    int32_t em_iters;      /* out: EM iterations run                         */
    double alpha;          /* out: final alpha (final.other)                 */
    double seconds_em;     /* out: wall time of the EM loop                  */
} oni_lda_result;

/* run_em (lda-estimate.c:112-321) for start = "random", over host buffers:
 * uploads the corpus, initialises the model with the reference's MT19937
 * stream (cokus.c, lda-model.c:104-117), runs EM on `ngpus` GPUs of this
 * process (documents sharded, one NCCL all-reduce of the sufficient statistics
 * per iteration when ngpus > 1), downloads the results.
 * initial_alpha is used as given (the CLI rounds it through float as
 * lda-estimate.c:425 does). */
int oni_lda_estimate(const oni_lda_corpus *corpus, int32_t num_topics, double initial_alpha,
                     const oni_lda_settings *settings, uint32_t seed, int32_t ngpus,
                     oni_lda_result *result);

/* ------------------------------------------------------------------------ */
/* Step-wise engine (one context = one GPU = one shard of the documents).     */
/* This is what a host that owns the process layout (one process per GPU,     */
/* e.g. under torchrun) drives; oni_lda_estimate is built on it.              */
typedef struct oni_lda_ctx oni_lda_ctx;

/* new_lda_model + new_lda_suffstats (lda-model.c:151-168,61-80) on `device`. */
int oni_lda_create(int32_t device, int32_t num_topics, int32_t num_terms,
                   const oni_lda_settings *settings, oni_lda_ctx **out);
void oni_lda_destroy(oni_lda_ctx *ctx);

/* Shard of documents for this context, from host CSR arrays (copied) ...     */
int oni_lda_set_corpus(oni_lda_ctx *ctx, int32_t num_docs, int64_t nnz, const int64_t *doc_ptr,
                       const int32_t *word, const int32_t *count);
/* ... or from arrays already resident on the context's device (copied d2d). */
int oni_lda_set_corpus_device(oni_lda_ctx *ctx, int32_t num_docs, int64_t nnz,
                              const int64_t *d_doc_ptr, const int32_t *d_word,
                              const int32_t *d_count);

/* random_initialize_ss + lda_mle(model, ss, 0) + alpha = initial
 * (lda-estimate.c:155-161): bit-identical to the reference for the same seed. */
int oni_lda_init_random(oni_lda_ctx *ctx, uint32_t seed, double initial_alpha);
/* Start from a given model: logbeta host [K][V] (load_lda_model semantics are
 * the caller's business), alpha. */
int oni_lda_set_model(oni_lda_ctx *ctx, const double *logbeta, double alpha);

/* E-step over the shard (doc_e_step for every document, lda-estimate.c:201-208,
 * 28-60; lda_inference + compute_likelihood, lda-inference.c:27-127) followed by
 * the deterministic reduction into the packed sufficient-statistics buffer.
 * Asynchronous on the context's stream. */
int oni_lda_e_step(oni_lda_ctx *ctx);

/* The packed sufficient statistics of the last E-step, on the device:
 *   [num_terms*KP] class_word, word-major, KP = num_topics rounded up to 4
 *   [KP]           class_total
 *   [1] alpha_suffstats  [1] likelihood  [1] num_docs  [1] word visits
 * This is the operand of the one all-reduce per EM iteration that replaces the
 * 2K+3 MPI_Allreduce calls of lda-estimate.c:211-252.  *count = doubles. */
int oni_lda_suffstats(oni_lda_ctx *ctx, double **d_buffer, size_t *count);

/* NCCL plumbing for one-process-per-GPU hosts: rank 0 makes an id, the host
 * broadcasts the bytes by its own means, every rank calls comm_init. */
#define ONI_NCCL_ID_BYTES 128
int oni_lda_comm_unique_id(unsigned char id[ONI_NCCL_ID_BYTES]);
int oni_lda_comm_init(oni_lda_ctx *ctx, const unsigned char id[ONI_NCCL_ID_BYTES],
                      int32_t rank, int32_t nranks);
/* Sum the packed buffer over all ranks (no-op without a communicator). */
int oni_lda_allreduce(oni_lda_ctx *ctx);

/* M-step: lda_mle (lda-model.c:27-54) incl. opt_alpha (lda-alpha.c:42-68) when
 * settings.estimate_alpha == 1, on the device; then the convergence bookkeeping
 * of lda-estimate.c:260-262 (relative change; var_max_iter doubles when the
 * bound decreased).  Synchronises; returns the bound and its relative change. */
int oni_lda_m_step(oni_lda_ctx *ctx, double *likelihood, double *converged);

/* e_step + allreduce + m_step. */
int oni_lda_em_iteration(oni_lda_ctx *ctx, double *likelihood, double *converged);

/* 1 while the loop condition of lda-estimate.c:186 holds. */
int oni_lda_em_should_continue(const oni_lda_ctx *ctx);

/* Readback (host buffers).  gamma is [shard docs][K] in shard order. */
int oni_lda_get_logbeta(oni_lda_ctx *ctx, double *logbeta /* [K][V] */);
int oni_lda_get_gamma(oni_lda_ctx *ctx, double *gamma /* [D][K] */);
int oni_lda_get_alpha(oni_lda_ctx *ctx, double *alpha);
int oni_lda_get_doc_likelihood(oni_lda_ctx *ctx, double *lik /* [D] */);
/* Per-document variational passes of the last E-step (diagnostics). */
int oni_lda_get_doc_iters(oni_lda_ctx *ctx, int32_t *iters /* [D] */);

/* Timing / accounting of the last E-step, for the bench. */
typedef struct {
    double ms_estep_kernel;  /* CUDA-event time of the per-document kernel    */
    double ms_suffstats;     /* deterministic reduction kernels               */
    double ms_mstep;         /* M-step + alpha kernels                        */
    int64_t word_visits;     /* sum over this rank's docs of pairs x passes   */
    int64_t kernel_launches; /* kernels launched by the library so far        */
    /* where the E-step's time goes: its three launches overlap on three streams; each   */
    /* figure is the time from the start of the E-step to the END of that launch (ms;   */
    /* -1 = not launched).  A shard's longest documents are serial chains of word visits */
    /* (lda-inference.c:50-75) and pace a small shard.                                   */
    double ms_end_exclusive; /* the longest documents, alone on their SMs     */
    double ms_end_long;      /* the other long documents, one per warp        */
    double ms_end_main;      /* everything else                               */
    int32_t docs_exclusive, docs_long; /* how many documents each of the two took */
    int64_t longest_doc_pairs;         /* pairs of the shard's longest document   */
} oni_lda_stats;
int oni_lda_get_stats(oni_lda_ctx *ctx, oni_lda_stats *out);
/* The CUDA stream (cudaStream_t) every kernel of this context is launched on,
 * so a host can bracket steps with its own events. */
int oni_lda_stream(oni_lda_ctx *ctx, void **cuda_stream);

/* ------------------------------------------------------------------------ */
/* File writers (save_lda_model lda-model.c:193-217, save_gamma              */
/* lda-estimate.c:88-104).  logbeta [K][V], gamma [D][K], host.               */
int oni_lda_save_model(const char *root, int32_t num_topics, int32_t num_terms,
                       const double *logbeta, double alpha);
int oni_lda_save_gamma(const char *path, int32_t num_docs, int32_t num_topics,
                       const double *gamma);

/* ------------------------------------------------------------------------ */
/* Post-LDA event scoring (flow_post_lda.scala:227-245,                      */
/* dns_post_lda.scala:312-326).                                               */
/* theta [D][K] and pwz [V][K] (word-major) are the tables lda_post.py writes */
/* (doc_results.csv / word_results.csv, lda_post.py:35-59,70-122).  Events    */
/* carry integer indices; index < 0 = lookup miss -> the constant vector      */
/* (0.05 flow, 0.1 DNS).  Score = sum_{i<K} theta[i]*p[i] with one rounded    */
/* multiply and one rounded add per term (JVM semantics, no FMA).             */
/* Survivors (key < tol, strict) come back ascending by key, ties by event    */
/* index.  out_idx/out_key have room for out_cap entries; *out_n is the total */
/* number of survivors (may exceed out_cap; then only the first out_cap       */
/* ranked entries are written).  out_src_score / out_dst_score receive the    */
/* two per-side scores of the ranked survivors.  All pointers are HOST memory. */
int oni_score_flow(int32_t device, int32_t num_topics, int32_t num_docs, int32_t num_terms,
                   const double *theta, const double *pwz, int64_t num_events,
                   const int32_t *src_doc, const int32_t *dst_doc, const int32_t *src_word,
                   const int32_t *dst_word, double tol, int64_t out_cap, int64_t *out_idx,
                   double *out_key, double *out_src_score /* [out_cap], ranked; may be NULL */,
                   double *out_dst_score /* [out_cap], ranked; may be NULL */, int64_t *out_n);
int oni_score_dns(int32_t device, int32_t num_topics, int32_t num_docs, int32_t num_terms,
                  const double *theta, const double *pwz, int64_t num_events,
                  const int32_t *doc, const int32_t *word, double tol, int64_t out_cap,
                  int64_t *out_idx, double *out_key, int64_t *out_n);

/* The same two calls over several GPUs of this process (the Spark stage's events
 * are independent): the events are cut into `ngpus` contiguous shares, devices 0..ngpus-1 each
 * hold a copy of the two tables and filter and rank their share, and the host merges the ranked
 * shares on (key, event index) -- the result is the single-GPU result, bit for bit.  No
 * collective is involved. */
int oni_score_flow_multi(int32_t ngpus, int32_t num_topics, int32_t num_docs, int32_t num_terms,
                         const double *theta, const double *pwz, int64_t num_events,
                         const int32_t *src_doc, const int32_t *dst_doc, const int32_t *src_word,
                         const int32_t *dst_word, double tol, int64_t out_cap, int64_t *out_idx,
                         double *out_key, double *out_src, double *out_dst, int64_t *out_n);
int oni_score_dns_multi(int32_t ngpus, int32_t num_topics, int32_t num_docs, int32_t num_terms,
                        const double *theta, const double *pwz, int64_t num_events,
                        const int32_t *doc, const int32_t *word, double tol, int64_t out_cap,
                        int64_t *out_idx, double *out_key, int64_t *out_n);

/* The same path with the event indices already on the device (what the host  */
/* entry points above are built on; a day of events is pushed chunk by chunk). */
typedef struct oni_score_ctx oni_score_ctx;
/* Uploads theta [D][K] and pwz [V][K] (host). */
int oni_score_create(int32_t device, int32_t num_topics, int32_t num_docs, int32_t num_terms,
                     const double *theta, const double *pwz, oni_score_ctx **out);
void oni_score_destroy(oni_score_ctx *ctx);
/* Empty the survivor list. */
int oni_score_begin(oni_score_ctx *ctx);
/* Score `num_events` events (device index arrays; b_* ignored for DNS) and     */
/* append those with key < tol; idx_base = global index of the first event.   */
int oni_score_events_device(oni_score_ctx *ctx, int32_t is_flow, int64_t num_events,
                            int64_t idx_base, const int32_t *a_doc, const int32_t *b_doc,
                            const int32_t *a_word, const int32_t *b_word, double tol);
/* The same from HOST index arrays (what oni_score_flow / oni_score_dns run on a context of  */
/* their own): the events go up in chunks on a copy stream while the previous chunk is scored, */
/* so a caller that scores day after day against the same tables keeps ONE context and pays   */
/* the table upload once.  Pinned arrays travel at PCIe speed, pageable ones slower.  An index */
/* >= num_docs / num_terms fails with ONI_ERR_ARG (checked by the kernel; -1 = lookup miss ->  */
/* the scripts' constant vector, flow_post_lda.scala:228-231).                                 */
int oni_score_events_host(oni_score_ctx *ctx, int32_t is_flow, int64_t num_events, int64_t idx_base,
                          const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                          const int32_t *b_word, double tol);
/* Rank the survivors: ascending key, ties by event index. */
int oni_score_rank(oni_score_ctx *ctx, int64_t *n_survivors);
/* First `cap` ranked survivors to host arrays. */
int oni_score_fetch(oni_score_ctx *ctx, int64_t cap, int64_t *out_idx, double *out_key);
/* src / dst score of the first `cap` ranked survivors (the two columns the    */
/* Scala stage appends, flow_post_lda.scala:238); event arrays as above.       */
int oni_score_fetch_sides(oni_score_ctx *ctx, int32_t is_flow, int64_t cap, int64_t idx_base,
                          const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                          const int32_t *b_word, double *out_a, double *out_b);
/* The same when the events live in HOST arrays (any number of chunks): the index quadruples */
/* of the `cap` ranked survivors are gathered on the host and scored once more.              */
int oni_score_fetch_sides_host(oni_score_ctx *ctx, int32_t is_flow, int64_t cap, int64_t idx_base,
                               const int32_t *a_doc, const int32_t *b_doc, const int32_t *a_word,
                               const int32_t *b_word, double *out_a, double *out_b);
/* CUDA-event times of the last run and kernels launched so far. */
int oni_score_get_stats(oni_score_ctx *ctx, double *ms_score, double *ms_sort, int64_t *launches);

/* Flow scoring and ranking for callers that must not allocate.
 * (No counterpart in the reference: an addition of this library.)
 *
 * Meant for hosts whose device memory is spoken for before they call us: an
 * executor that was given a fixed slice of the GPU by a pool allocator, a
 * serving process that reserved everything at start-up, or one that runs
 * latency-critical streams beside this call.  For them cudaMalloc/cudaFree can
 * fail late (the pool holds the memory) and stall the whole device while they
 * run, so this entry point calls neither, nor any routine that allocates.
 *
 * Same result as oni_score_flow (survivors with key < tol, ascending by key,
 * ties by event index), but
 *   - every array is DEVICE memory the caller owns, on `device`: theta [D][K],
 *     pwz [V][K], the four event index arrays [num_events] (index < 0 = lookup
 *     miss; indices are NOT range-checked here), and the outputs d_out_idx /
 *     d_out_key [out_cap] (either may be NULL); *out_n (host) receives the
 *     number of survivors;
 *   - the working memory is a fixed array inside the library image (about
 *     8.2 MB, present on every GPU that loads the library until the process
 *     exits) with room for ONI_NOALLOC_MAX_SURVIVORS events below the
 *     threshold.  More survivors than that, or out_cap above it, is refused
 *     with ONI_ERR_ARG and a message; use oni_score_flow then;
 *   - the first call on a device looks up the array's device address
 *     (cudaGetSymbolAddress); make that first call before starting any stream
 *     capture on that device;
 *   - all calls on a device share the array, which keeps the previous call's
 *     contents (the call clears what has to start at zero).  Calls that overlap
 *     on one GPU race with each other: issue them one after the other.
 * `cuda_stream` is the cudaStream_t to run on (NULL = the default stream); the
 * call returns after the stream has finished its work. */
#define ONI_NOALLOC_MAX_SURVIVORS (1 << 18)
int oni_score_flow_noalloc(int32_t device, int32_t num_topics, const double *d_theta,
                           const double *d_pwz, int64_t num_events, const int32_t *d_src_doc,
                           const int32_t *d_dst_doc, const int32_t *d_src_word,
                           const int32_t *d_dst_word, double tol, int64_t out_cap,
                           int64_t *d_out_idx, double *d_out_key, int64_t *out_n,
                           void *cuda_stream);

/* Result shaping of lda_post.py on the device: theta = gamma/sum(gamma)
 * (lda_post.py:44-54) and pwz = exp(beta)/sum_w exp(beta) (lda_post.py:88-96),
 * both passed through 12-significant-digit decimal text when text_digits = 12
 * (0 = keep full doubles).  gamma [D][K], logbeta [K][V], host; outputs host. */
int oni_lda_post(int32_t device, int32_t num_docs, int32_t num_topics, int32_t num_terms,
                 const double *gamma, const double *logbeta, int32_t text_digits,
                 double *theta /* [D][K] */, double *pwz /* [V][K] */);

/* lda_post.py as a whole (ml_ops.sh:84): reads <dir>final.gamma, final.beta, doc.dat and
 * words.dat, writes <dir>doc_results.csv ("ip,t0 t1 ...", lda_post.py:35-59; an all-zero
 * gamma row gets the script's twenty "0.0") and <dir>word_results.csv ("word,p0 p1 ...",
 * word names joined with '_' and cut to 20 bytes like the script's S20 array,
 * lda_post.py:70-122).  Numbers are printed the way Python-2 str() prints them: 12
 * significant digits, ".0" appended to integral values.  `dir` is a prefix ("path/"). */
int oni_lda_post_files(const char *dir, int32_t device);

/* ---- raw events -> scored rows: the Spark scripts around the score loop ------------------
 *
 * Replaces flow_post_lda.scala:48-248 and dns_post_lda.scala:55-330 as a whole (SURVEY.md
 * 8(f) N3).  `open` does the host work on all cores: header removal and the field-count
 * filter (flow :128-135, dns :138), the empirical-CDF cut vectors (flow :67-103,148-158; dns
 * :214-266), binning (flow :161-180, dns :201-204), the port -> word rule with the script's
 * own column naming (flow :182-218), the sub-domain, entropy and top-domain features (dns
 * :154-190,206-212,272), and the lookups of IPs and words in doc_results.csv /
 * word_results.csv (a miss becomes index -1 = the script's constant vector).
 * `score_write` runs the score loop, threshold and ranking on the GPU (oni_score_flow /
 * oni_score_dns) and writes the surviving rows, original fields first, then the derived
 * columns, then the score(s), comma separated, one file instead of Spark's part-* directory.
 *
 * Inputs: flow = the nfdump CSV the script reads (27 columns, first line = header);
 * dns = one line per event holding the eight columns the script selects from its parquet
 * files, comma separated (frame_time, unix_tstamp, frame_len, ip_dst, dns_qry_name,
 * dns_qry_class, dns_qry_type, dns_qry_rcode), plus the Alexa-style "rank,domain" list the
 * script opens as top-1m.csv.
 *
 * Fixed here because the reference leaves them to the JVM: numbers are printed as the
 * shortest decimal that reads back to the same double, in Double.toString's layout; rows of
 * equal score keep their input order; the entropy sum runs over the distinct bytes in
 * ascending order.  A field that does not parse as a number is an error (the script dies
 * with a NumberFormatException). */
typedef struct oni_events oni_events;
int oni_flow_events_open(const char *flow_csv, const char *doc_results_csv,
                         const char *word_results_csv, oni_events **out);
int oni_dns_events_open(const char *dns_csv, const char *top_domains_csv,
                        const char *doc_results_csv, const char *word_results_csv,
                        oni_events **out);
/* The pre-LDA scripts also read the analysts' feedback (flow_pre_lda.scala:147-263 flow_scores.csv,
 * dns_pre_lda.scala:70-138 dns_scores.csv): rows marked severity 3 are turned back into event
 * rows (flow: every column the word does not use becomes "##") and appended dup_factor times
 * each, behind the events of the day.  First line of the feedback file = header. */
int oni_flow_events_open_feedback(const char *flow_csv, const char *feedback_csv, int32_t dup_factor,
                                  const char *doc_results_csv, const char *word_results_csv,
                                  oni_events **out);
int oni_dns_events_open_feedback(const char *dns_csv, const char *top_domains_csv,
                                 const char *feedback_csv, int32_t dup_factor,
                                 const char *doc_results_csv, const char *word_results_csv,
                                 oni_events **out);
void oni_events_close(oni_events *ev);
/* lines read, rows kept by the filters, distinct keys of the two result tables */
int oni_events_info(const oni_events *ev, int64_t *lines_in, int64_t *rows, int32_t *num_docs,
                    int32_t *num_words);
/* cut vector `which`: flow 0 time, 1 ibyt, 2 ipkt; dns 0 unix_tstamp, 1 frame_len,
 * 2 subdomain.length, 3 subdomain.entropy, 4 num.periods.  *n = 10 or 5. */
int oni_events_cuts(const oni_events *ev, int32_t which, double *out, int32_t cap, int32_t *n);
/* the indices handed to the scorer, [rows] each: flow (src_doc, dst_doc, src_word, dst_word);
 * dns (doc, -, word, -).  Any pointer may be NULL. */
int oni_events_indices(const oni_events *ev, int32_t *a_doc, int32_t *b_doc, int32_t *a_word,
                       int32_t *b_word);
/* text of kept row `row` as the script holds it just before scoring (no score columns).
 * Returns its length (the copy is cut to cap-1 bytes and NUL terminated), -1 on a bad row. */
int64_t oni_events_row_text(const oni_events *ev, int64_t row, char *buf, int64_t cap);
/* The pre-LDA stage (flow_pre_lda.scala:371-382, dns_pre_lda.scala:326-330): one line
 * "ip,word,count" per distinct (ip, word) -- the doc_wc.dat that lda_pre.py / oni_lda_pre read.
 * Flow writes the source-side list, then the destination-side list (a pair can occur in both,
 * as in the script).  Line order (Spark: hash partitioner) is first appearance here.  For this
 * stage the handle may be opened with doc_results_csv = word_results_csv = NULL.  The scripts'
 * analyst-feedback rows come in through the *_open_feedback calls above. */
int oni_events_word_counts(const oni_events *ev, const char *out_path, int64_t *num_lines);
/* GPU: score every kept row, keep score < tol, rank ascending, write the rows to out_path */
int oni_events_score_write(oni_events *ev, double tol, const char *out_path, int32_t device,
                           int64_t *num_scored);

#ifdef __cplusplus
}
#endif
#endif /* ONI_LDA_H */
