// estep.cuh -- per-document variational inference (the E-step) on sm_100a.
//
// Stands in for lda_inference + compute_likelihood (oni-lda-c/lda-inference.c:
// 27-127) and the per-document part of doc_e_step (lda-estimate.c:28-46).
//
// Mapping.  A document is owned by a GROUP of G adjacent lanes; each lane holds
// T topics (topic k = j + G*t for lane-in-group j), so a warp runs 32/G
// documents at once and all 32 lanes carry live topics when G*T == K (K=20:
// G=4, T=5 -> 8 documents per warp).  Groups pull documents from a global
// ticket counter in longest-first order; the result of a document does not
// depend on which group ran it, so the dynamic schedule is deterministic in its
// outputs.  The document's words are visited strictly in order and gamma /
// digamma(gamma) are refreshed after every word, exactly like the reference's
// Gauss-Seidel sweep (lda-inference.c:50-75); only the K-wide work inside one
// word visit is parallel.
//
// Arithmetic (fp64 throughout), restated rather than transliterated:
//   * phi_nk = exp(x_k) / sum_j exp(x_j), x_k = digamma(gamma_k) + logbeta_kw.
//     The reference forms log(sum) with a left-to-right log_sum chain
//     (lda-inference.c:53-64, utils.c:8-21) and then exp(x_k - logsum); the two
//     agree to a few ulp.  A max-shifted variant (SHIFT) is used when alpha is
//     so small that exp(x) could underflow for every topic.
//   * digamma (utils.c:55-64): same asymptotic series and literal coefficients;
//     the six recurrence terms 1/(x-1)..1/(x-6) = sum_{j<6} 1/(g+j) are folded
//     into one rational P(g)/Q(g) (one reciprocal instead of seven divisions).
//   * exp, log and the reciprocal are the branch-free table versions of
//     devmath.cuh (~1 ulp), so the T topic chains of a lane interleave.
//   * the bound (lda-inference.c:93-127).  With S_n = sum_k exp(x_k) at visit n
//     and psi_k^(n) the digamma values used at that visit,
//         log phi_nk = psi_k^(n) + logbeta_kw - log S_n ,
//     and after a full sweep gamma_k - alpha = sum_n c_n phi_nk, so the
//     (dig_k - digsum) terms of the reference's expression cancel and
//         L = lgamma(K alpha) - K lgamma(alpha) - lgamma(sum gamma)
//             + sum_k lgamma(gamma_k)
//             + sum_n c_n ( log S_n - sum_k phi_nk psi_k^(n) ).
//     This needs one log per word visit instead of K, and no second sweep over
//     phi.  Terms with phi == 0 vanish in both forms (the reference skips them,
//     lda-inference.c:119).
//   * convergence test and iteration cap exactly as lda-inference.c:46-47,79-80,
//     including (0 - L)/0 = +inf on the first pass.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "devmath.cuh"

namespace oni {

// exp/log tables (6 KB), copied to shared memory by every block of the E-step
static __device__ const MathTables g_math_tables = ONI_MATH_TABLES_INIT;

// ------------------------------------------------------------ special functions
// Sum_{j=0..5} 1/(g+j) as P(g)/Q(g); all coefficients positive, so Horner on
// g > 0 has no cancellation.
__device__ __forceinline__ double recurrence6(double g)
{
    double q = g + 15.0;
    q = fma(q, g, 85.0);
    q = fma(q, g, 225.0);
    q = fma(q, g, 274.0);
    q = fma(q, g, 120.0);
    q = q * g;
    double p = fma(6.0, g, 75.0);
    p = fma(p, g, 340.0);
    p = fma(p, g, 675.0);
    p = fma(p, g, 548.0);
    p = fma(p, g, 120.0);
    return p / q;
}

// utils.c:55-64
__device__ __forceinline__ double dev_digamma(double g)
{
    const double x = g + 6.0;
    const double r = 1.0 / x;
    const double p = r * r;
    double ser = fma(0.004166666666667, p, -0.003968253986254);
    ser = fma(ser, p, 0.008333333333333);
    ser = fma(ser, p, -0.083333333333333);
    ser = ser * p;
    return ((ser + log(x)) - 0.5 * r) - recurrence6(g);
}

// utils.c:32-47 (only the alpha Newton uses it; kept literal)
__device__ inline double dev_trigamma(double x)
{
    x = x + 6.0;
    double p = 1.0 / (x * x);
    p = (((((0.075757575757576 * p - 0.033333333333333) * p + 0.0238095238095238) * p -
           0.033333333333333) * p + 0.166666666666667) * p + 1.0) / x + 0.5 * p;
    for (int i = 0; i < 6; i++) {
        x = x - 1.0;
        p = 1.0 / (x * x) + p;
    }
    return p;
}

// ------------------------------------------------------------------ parameters
struct EstepParams {
    const int2 *tok;           // [NNZ] (word, count)
    const int64_t *doc_ptr;    // [D+1]
    const int32_t *order;      // [D] documents, longest first
    const double *doc_total;   // [D] sum of counts (lda-data.c:47)
    const double *logbeta_t;   // [V][KP] word-major log p(w|k)
    double *phi;               // [NNZ][KP]
    double *gamma;             // [D][K]
    double *doc_lik;           // [D]
    double *doc_ass;           // [D] sum_k digamma(gamma_k) - K digamma(sum gamma)
    int32_t *doc_iters;        // [D]
    const double *alpha;       // device scalar
    unsigned int *ticket;      // next position in `order`
    int32_t *nan_flag;
    int32_t D, K, KP;
    int32_t var_max_iter;      // -1 = no cap
    double var_converged;
    int32_t no_exact;          // 1: not the exact-fit build (this launch may hold one-pair documents)
    const double *psi_tab;     // long-document kernel only (psi_table.h): [PSI_N][6]
    const double *exk_tab;     // ... [EXK_N]
};

__device__ __forceinline__ void prefetch_l2(const void *p)
{
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

// How many visits ahead the phi row is prefetched into L2 (0 = off; 3 and 6 measured 3 % slower).
#ifndef ONI_ESTEP_PF_AHEAD
#define ONI_ESTEP_PF_AHEAD 0
#endif

template <int G>
__device__ __forceinline__ double group_sum(double v, unsigned gmask)
{
#pragma unroll
    for (int o = 1; o < G; o <<= 1) v += __shfl_xor_sync(gmask, v, o);
    return v;
}
template <int G>
__device__ __forceinline__ double group_max(double v, unsigned gmask)
{
#pragma unroll
    for (int o = 1; o < G; o <<= 1) v = fmax(v, __shfl_xor_sync(gmask, v, o));
    return v;
}

// Asynchronous 8-byte copy global -> shared (LDGSTS): nothing waits on a register, so a row can
// be requested many visits ahead without a scoreboard stall when it is next touched.
__device__ __forceinline__ void cp_async8(void *smem_dst, const void *gsrc)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---------------------------------------------------------------- the kernel
// A group that reaches the end of its sweep waits up to this many word visits for the
// other groups of its warp, so that their end-of-sweep sections (bound, convergence
// test, refill) run together instead of one group at a time.  0 = never wait.
#ifndef ONI_ESTEP_WAIT
#define ONI_ESTEP_WAIT 256
#endif

// Latency-tuned arithmetic in the one-document-per-warp instantiation (0 = off).
#ifndef ONI_ESTEP_LAT
#define ONI_ESTEP_LAT 1
#endif

// Long-document build: rows and pairs arrive through a shared-memory ring filled by asynchronous
// copies requested 8/T visits ahead (0 = the one-visit-ahead register loads of the main kernel).
// With the main kernel streaming phi through HBM a row takes ~1700 cycles to arrive, two visits
// of a lone document; register loads cannot run further ahead than the first instruction that
// moves them, the ring can.
#ifndef ONI_ESTEP_LAT_RING
#define ONI_ESTEP_LAT_RING 1
#endif

// Block size and the resident blocks per SM the register allocation is tuned for.
#ifndef ONI_ESTEP_THREADS
#define ONI_ESTEP_THREADS 128
#endif
#ifndef ONI_ESTEP_MINB
#define ONI_ESTEP_MINB 3
#endif

#if defined(ONI_ESTEP_MAXNREG)
#define ONI_ESTEP_BOUNDS __maxnreg__(ONI_ESTEP_MAXNREG)
#elif ONI_ESTEP_MINB > 0
#define ONI_ESTEP_BOUNDS __launch_bounds__(ONI_ESTEP_THREADS, ONI_ESTEP_MINB)
#else
#define ONI_ESTEP_BOUNDS __launch_bounds__(ONI_ESTEP_THREADS)
#endif
// EXACT: the host promises K == G*T, so every topic slot of every lane is live and the
// per-slot validity tests (2 % of the visit's instructions at K = 20) compile away.
template <int G, int T, bool SHIFT, bool LONGDOC = false, bool EXACT = false>
__global__ void ONI_ESTEP_BOUNDS estep_kernel(const EstepParams p)
{
    __shared__ MathTables tabs;
    // one document per warp (G = 32): a long document is a serial chain of visits, so this
    // instantiation trades a few instructions for a shorter dependency chain per visit
    // (LONGDOC launches hold only documents of >= 1024 pairs)
    constexpr bool LAT = LONGDOC && (G == 32) && (ONI_ESTEP_LAT != 0);
    __shared__ double red_buf[LAT ? ONI_ESTEP_THREADS : 1];  // per-warp exchange buffer of the sum below
    constexpr bool RING = LAT && (ONI_ESTEP_LAT_RING != 0) && T <= 4;
    constexpr int QD = (T == 1) ? 8 : (T == 2) ? 4 : 2;  // visits ahead for the rows
    constexpr int QT = 2 * QD;                           // ... and for the pairs (their addresses)
    constexpr int NW = ONI_ESTEP_THREADS / 32;
    __shared__ double ring_lb[RING ? NW * QD * T * 32 : 1];   // [warp][slot][t][lane]
    __shared__ double ring_old[RING ? NW * QD * T * 32 : 1];
    __shared__ int2 ring_tk[RING ? NW * QT * 32 : 1];         // [warp][slot][lane] (every lane its own copy)
    {
        const double *src = reinterpret_cast<const double *>(&g_math_tables);
        double *dst = reinterpret_cast<double *>(&tabs);
        for (int i = threadIdx.x; i < (int)(sizeof(MathTables) / sizeof(double)); i += blockDim.x)
            dst[i] = src[i];
    }
#ifdef ONI_ESTEP_REPL
    // Experiment (profiles/r2_estep_experiments.txt): bank-interleaved copies of the two tables in
    // dynamic shared memory.  A table read is one lane = one random entry; the 16 lanes of a
    // half-warp (LDS.64) or the 8 lanes of a quarter-warp (LDS.128) then collide in banks ~3-deep.
    // With copy r = lane mod 16 (mod 8) laid out entry-major, every lane of a phase owns a bank.
    extern __shared__ __align__(16) unsigned char repl_smem[];
    double *const ex_rep = reinterpret_cast<double *>(repl_smem);                                  // [256][16]
    LogEntry *const lg_rep = reinterpret_cast<LogEntry *>(repl_smem + sizeof(double) * ONI_MT_N * 16);  // [256][8]
    for (int i = threadIdx.x; i < ONI_MT_N * 16; i += blockDim.x) ex_rep[i] = g_math_tables.ex[i >> 4];
    for (int i = threadIdx.x; i < ONI_MT_N * 8; i += blockDim.x) lg_rep[i] = g_math_tables.lg[i >> 3];
#endif
    __syncthreads();

    // Topic of slot t of lane j.  Ordinary builds: k = j + G t (the G lanes of a group read G
    // consecutive doubles per slot).  The exact K = 20 build (4 lanes x 5 topics) pairs the slots so
    // that a lane's topics are (2j, 2j+1), (8+2j, 9+2j) and 16+j: two 16-byte accesses and one
    // 8-byte access per 160-byte row instead of five 8-byte ones (same sectors, 3 instructions for 5).
#ifdef ONI_ESTEP_NO_PAIRED_SLOTS
    constexpr bool V2 = false;
#else
    constexpr bool V2 = EXACT && G == 4 && T == 5 && !LONGDOC;
#endif
    const int lane = threadIdx.x & 31;
#ifdef ONI_ESTEP_REPL
    constexpr int XS = 16, LS = 8;
    const double *const ex_l = ex_rep + (lane & 15);
    const LogEntry *const lg_l = lg_rep + (lane & 7);
#else
    constexpr int XS = 1, LS = 1;
    const double *const ex_l = tabs.ex;
    const LogEntry *const lg_l = tabs.lg;
#endif
    const int j = lane & (G - 1);
    const int lead = lane & ~(G - 1);
    const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << lead);
    const int K = p.K, KP = p.KP;
    double *const my_lb = ring_lb + (RING ? (threadIdx.x >> 5) * QD * T * 32 + lane : 0);
    double *const my_old = ring_old + (RING ? (threadIdx.x >> 5) * QD * T * 32 + lane : 0);
    int2 *const my_tk = ring_tk + (RING ? (threadIdx.x >> 5) * QT * 32 + lane : 0);
    int vi = 0;  // RING: visits of the current document so far (ring slots are vi mod QD / QT)
    const double alpha = *p.alpha;
    const double inv_k = 1.0 / (double)K;
    // lda-inference.c:105-108
    const double lik_const = lgamma(alpha * K) - K * lgamma(alpha);
    // T = ceil(K/G): only the last topic slot of a lane can fall outside [0,K)
    const bool last_ok = EXACT || (j + G * (T - 1)) < K;

    double gam[T], dg[T];
    double lbc[T], oldc[T];           // rows of the visit at `pos`, loaded one visit ahead
    double acc_a = 0.0, acc_b = 0.0;  // sum c*phi*psi (this lane's topics), sum c*log S
    double acc_s = 0.0;               // SHIFT only: sum c*shift (same in every lane)
    double s_keep = 1.0, c_keep = 0.0;
    double lik_old = 0.0;
    // pair positions fit 32 bits (a shard holds fewer than 2^31 pairs, checked when it is indexed)
    uint32_t beg = 0, end = 0, pos = 0;
    int32_t doc = -1, pass = 0;
    int2 tk = make_int2(0, 0);   // (word, count) of the pair at `pos`
    int2 tk1 = make_int2(0, 0);  // ... and of the visit after it
    bool have = false, exhausted = false;
    int waited = 0;

    while (true) {
        if (!have && !exhausted) {
            unsigned i = 0;
            if (j == 0) i = atomicAdd(p.ticket, 1u);
            i = __shfl_sync(gmask, i, lead);
            if (i >= (unsigned)p.D) {
                exhausted = true;
            } else {
                doc = p.order[i];
                beg = (uint32_t)p.doc_ptr[doc];
                end = (uint32_t)p.doc_ptr[doc + 1];
                pos = beg;
                pass = 0;
                lik_old = 0.0;
                acc_a = acc_b = acc_s = 0.0;
                s_keep = 1.0; c_keep = 0.0;
                // lda-inference.c:37-43
                const double g0 = alpha + p.doc_total[doc] / ((double)K);
                const double d0 = digamma_ref<LS>(g0, lg_l);
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const bool ok = (t < T - 1) || last_ok;
                    gam[t] = ok ? g0 : 1.0;
                    dg[t] = ok ? d0 : 0.0;
                }
                if (RING) {
                    // (a long document has far more than QT pairs)
                    cp_async_wait<0>();  // nothing of the previous document may still land in the ring
                    vi = 0;
#pragma unroll
                    for (int i = 0; i < QT; i++) cp_async8(my_tk + i * 32, p.tok + beg + i);
                    cp_async_commit();
                    cp_async_wait<0>();
#pragma unroll
                    for (int i = 0; i < QD; i++) {
                        const int2 ti = my_tk[i * 32];
                        const double *r = p.logbeta_t + (size_t)ti.x * KP;
#pragma unroll
                        for (int t = 0; t < T; t++)
                            cp_async8(my_lb + (i * T + t) * 32, r + (((t < T - 1) || last_ok) ? j + G * t : 0));
                        cp_async_commit();  // one group per visit: visit 0 waits for the first only
                    }
                } else if (beg < end) {
                    // prime the one-visit-ahead pipeline (first sweep: old phi = 1/K)
                    tk = __ldg(p.tok + beg);
                    tk1 = __ldg(p.tok + ((beg + 1 < end) ? beg + 1 : beg));
                    const double *lbrow = p.logbeta_t + (size_t)tk.x * KP;
                    if constexpr (V2) {
                        const double2 a = __ldg(reinterpret_cast<const double2 *>(lbrow + 2 * j));
                        const double2 b = __ldg(reinterpret_cast<const double2 *>(lbrow + 8 + 2 * j));
                        lbc[0] = a.x; lbc[1] = a.y; lbc[2] = b.x; lbc[3] = b.y; lbc[4] = __ldg(lbrow + 16 + j);
#pragma unroll
                        for (int t = 0; t < T; t++) oldc[t] = inv_k;
                    } else {
#pragma unroll
                    for (int t = 0; t < T; t++) {
                        const bool ok = (t < T - 1) || last_ok;
                        lbc[t] = ok ? __ldg(lbrow + j + G * t) : 0.0;
                        oldc[t] = ok ? inv_k : 0.0;
                    }
                    }
                }
                have = true;
            }
        }
        if (__all_sync(0xffffffffu, !have)) break;

        if (have && pos < end) {
            // ---- one word visit (lda-inference.c:50-75)
            if (RING) {
                // this visit's pair and rows out of the ring (requested QD visits ago), then the
                // request for the visit QD ahead into the slots just read
                cp_async_wait<QD - 1>();
                const int st = vi & (QT - 1), sr = vi & (QD - 1);
                tk = my_tk[st * 32];
#pragma unroll
                for (int t = 0; t < T; t++) {
                    lbc[t] = my_lb[(sr * T + t) * 32];
                    oldc[t] = my_old[(sr * T + t) * 32];  // (not used during the first sweep)
                }
                const uint32_t len = end - beg;
                const uint32_t rpos = (pos + QD < end) ? pos + QD : pos + QD - len;
                const uint32_t tpos = (pos + QT < end) ? pos + QT : pos + QT - len;
                const int2 rtk = my_tk[((vi + QD) & (QT - 1)) * 32];
                const double *nlb = p.logbeta_t + (size_t)rtk.x * KP;
                const double *nph = p.phi + (size_t)rpos * KP;
                const bool row_first_sweep = (pass == 0) && (pos + QD < end);
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const int k = ((t < T - 1) || last_ok) ? j + G * t : 0;
                    cp_async8(my_lb + (sr * T + t) * 32, nlb + k);
                    if (!row_first_sweep) cp_async8(my_old + (sr * T + t) * 32, nph + k);
                }
                cp_async8(my_tk + st * 32, p.tok + tpos);
                cp_async_commit();
                vi++;
            }
            const double c = (double)tk.y;
            double *phrow = p.phi + (size_t)pos * KP;
            // Software pipeline: the rows of the NEXT visit (wrapping to the first word
            // for the next sweep) are requested now and consumed one visit later, so
            // their latency hides behind this visit's arithmetic.
            const uint32_t npos = (pos + 1 < end) ? pos + 1 : beg;
            const uint32_t n2pos = (npos + 1 < end) ? npos + 1 : beg;
            int2 tk2 = make_int2(0, 0);
            if (!RING) tk2 = __ldg(p.tok + n2pos);
            if (ONI_ESTEP_PF_AHEAD > 0 && pass > 0 && j * 128 < KP * 8 + 127) {
                // the phi rows stream from HBM (every sweep rereads what the last one wrote);
                // pull the row of a later visit into L2 so the register load above finds it there
                const uint32_t fpos = pos + ONI_ESTEP_PF_AHEAD;
                if (fpos < end) prefetch_l2(reinterpret_cast<const char *>(p.phi + (size_t)fpos * KP) + j * 128);
            }
            const bool next_first_sweep = (pass == 0) && (npos != beg);
            double lbn[T], oldn[T] = {};
            if (!RING && V2) {
                const double *nlb = p.logbeta_t + (size_t)tk1.x * KP;
                const double *nph = p.phi + (size_t)npos * KP;
                const double2 a = __ldg(reinterpret_cast<const double2 *>(nlb + 2 * j));
                const double2 b = __ldg(reinterpret_cast<const double2 *>(nlb + 8 + 2 * j));
                lbn[0] = a.x; lbn[1] = a.y; lbn[2] = b.x; lbn[3] = b.y; lbn[4] = __ldg(nlb + 16 + j);
                if (next_first_sweep) {
#pragma unroll
                    for (int t = 0; t < T; t++) oldn[t] = inv_k;
                } else {
                    const double2 c2 = *reinterpret_cast<const double2 *>(nph + 2 * j);
                    const double2 d2 = *reinterpret_cast<const double2 *>(nph + 8 + 2 * j);
                    oldn[0] = c2.x; oldn[1] = c2.y; oldn[2] = d2.x; oldn[3] = d2.y; oldn[4] = nph[16 + j];
                }
            } else if (!RING) {
                const double *nlb = p.logbeta_t + (size_t)tk1.x * KP;
                const double *nph = p.phi + (size_t)npos * KP;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const int k = j + G * t;
                    const bool ok = (t < T - 1) || last_ok;
                    if (LAT) {
                        // unconditional loads from a clamped address: a select on the loaded
                        // VALUE would wait for the load right here instead of one visit later
                        // (20 % of a lone long document's time); the selects move to the use
                        lbn[t] = __ldg(nlb + (ok ? k : 0));
                        if (!next_first_sweep) oldn[t] = nph[ok ? k : 0];
                    } else {
                        lbn[t] = ok ? __ldg(nlb + k) : 0.0;
                        oldn[t] = ok ? (next_first_sweep ? inv_k : nph[k]) : 0.0;
                    }
                    // (a one-word document rereads the row this visit is about to write;
                    //  that case is patched below)
                }
            }
            double x[T];
#pragma unroll
            for (int t = 0; t < T; t++) {
                x[t] = dg[t] + lbc[t];
            }
            double shift = 0.0;
            if (SHIFT) {
                double m = -1.0e300;
#pragma unroll
                for (int t = 0; t < T; t++) m = ((t < T - 1) || last_ok) ? fmax(m, x[t]) : m;
                shift = group_max<G>(m, gmask);
            }
            double s = 0.0;
#pragma unroll
            for (int t = 0; t < T; t++) {
                const double e = exp_tab<XS>(SHIFT ? (x[t] - shift) : x[t], ex_l);
                x[t] = ((t < T - 1) || last_ok) ? e : 0.0;
                s += x[t];
            }
            if (LAT) {
                // all 32 partial sums through shared memory and a register tree: one LDS
                // latency + 5 add levels instead of 5 shuffle+add levels
                double *rb = red_buf + (threadIdx.x >> 5) * 32;
                rb[lane] = s;
                __syncwarp();
                double v[32];
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const double2 two = reinterpret_cast<const double2 *>(rb)[i];
                    v[2 * i] = two.x; v[2 * i + 1] = two.y;
                }
#pragma unroll
                for (int w = 16; w > 0; w >>= 1)
#pragma unroll
                    for (int i = 0; i < w; i++) v[i] += v[i + w];
                s = v[0];
                __syncwarp();
            } else {
                s = group_sum<G>(s, gmask);
            }
            const double rs = rcp_pos(s);
            double a = 0.0;
#pragma unroll
            for (int t = 0; t < T; t++) {
                const double ph = x[t] * rs;
                a = fma(ph, dg[t], a);
                // old phi: 1/K during the first sweep (lda-inference.c:42), else the row loaded
                // one visit ahead; nothing for a topic slot beyond K
                const bool ok = (t < T - 1) || last_ok;
                const double old = LAT ? (ok ? ((pass == 0) ? inv_k : oldc[t]) : 0.0) : oldc[t];
                if (LAT) gam[t] = fma(c * x[t], rs, fma(-c, old, gam[t]));  // same value, one op after rs
                else gam[t] = fma(c, ph - old, gam[t]);                     // lda-inference.c:69-70
                x[t] = ph;
            }
#pragma unroll
            for (int t = 0; t < T; t++)
                dg[t] = LAT ? digamma_ref_lat(gam[t], tabs.lg) : digamma_ref<LS>(gam[t], lg_l);
            if constexpr (V2) {
                *reinterpret_cast<double2 *>(phrow + 2 * j) = make_double2(x[0], x[1]);
                *reinterpret_cast<double2 *>(phrow + 8 + 2 * j) = make_double2(x[2], x[3]);
                phrow[16 + j] = x[4];
            } else {
#pragma unroll
            for (int t = 0; t < T; t++)
                if ((t < T - 1) || last_ok) phrow[j + G * t] = x[t];
            }
            // rotate the pipeline
            if (!RING) {
#pragma unroll
                for (int t = 0; t < T; t++) {
                    lbc[t] = lbn[t];
                    oldc[t] = oldn[t];
                }
            }
            // one-word document: the row just written is the next "old".  (Compiled out of the
            // long-document kernel: it becomes a select on the in-flight load and cost a lone
            // long document a quarter of its time.)
            // (the exact-fit build is never given one-pair documents: the engine launches those
            //  separately, which takes ten selects out of every visit of everything else)
            if (!LONGDOC && !EXACT && npos == pos) {
#pragma unroll
                for (int t = 0; t < T; t++) oldc[t] = x[t];
            }
            if (!RING) {
                tk = tk1;
                tk1 = tk2;
            }
            acc_a = fma(c, a, acc_a);
            if (SHIFT) acc_s = fma(c, shift, acc_s);  // log S_n = shift + log(sum)
            // log S_n is needed once per visit: lane (n mod G) keeps S_n and the
            // group takes the logs of G visits in one go.
            const int slot = (int)(pos - beg) & (G - 1);
            if (slot == j) { s_keep = s; c_keep = c; }
            pos++;
            if (slot == G - 1 || pos == end) {
                acc_b = fma(c_keep, log_tab<LS>(s_keep, lg_l), acc_b);
                s_keep = 1.0; c_keep = 0.0;
            }
        }
        // groups still inside their sweep (every lane of the warp votes here)
        const bool others_stepping =
            (ONI_ESTEP_WAIT > 0) && __any_sync(0xffffffffu, have && pos < end);
        if (have && pos == end && others_stepping && waited < ONI_ESTEP_WAIT) waited++;
        else if (have && pos == end) {
            waited = 0;
            // ---- end of a sweep: the bound (lda-inference.c:93-127, folded)
            double gs = 0.0, part = 0.0;
#pragma unroll
            for (int t = 0; t < T; t++) {
                const bool ok = (t < T - 1) || last_ok;
                gs += ok ? gam[t] : 0.0;
                part += ok ? lgamma_pos<LS>(gam[t], lg_l) : 0.0;
            }
            part += acc_b - acc_a;
#pragma unroll
            for (int o = 1; o < G; o <<= 1) {
                gs += __shfl_xor_sync(gmask, gs, o);
                part += __shfl_xor_sync(gmask, part, o);
            }
            const double lik = (lik_const - lgamma_pos<LS>(gs, lg_l)) + part + (SHIFT ? acc_s : 0.0);
            pass++;
            // lda-inference.c:79-80 and the loop condition :46-47
            const double conv = (lik_old - lik) / lik_old;
            lik_old = lik;
            bool more = (conv > p.var_converged) &&
                        ((pass < p.var_max_iter) || (p.var_max_iter == -1));
            if (lik != lik) {  // assert(!isnan(likelihood)), lda-inference.c:78
                if (j == 0) atomicExch(p.nan_flag, 1);
                more = false;
            }
            if (more) {
                pos = beg;
                acc_a = acc_b = acc_s = 0.0;
            } else {
                // ---- document done: gamma row, bound, alpha statistic
                // (lda-estimate.c:40-46)
                double ds = 0.0;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const int k = V2 ? ((t < 4) ? (t >> 1) * 8 + 2 * j + (t & 1) : 16 + j) : j + G * t;
                    const bool ok = (t < T - 1) || last_ok;
                    ds += ok ? dg[t] : 0.0;
                    if (ok) p.gamma[(size_t)doc * K + k] = gam[t];
                }
                ds = group_sum<G>(ds, gmask);
                if (j == 0) {
                    p.doc_lik[doc] = lik;
                    p.doc_ass[doc] = ds - K * digamma_ref<LS>(gs, lg_l);
                    p.doc_iters[doc] = pass;
                }
                have = false;
            }
        }
    }
}

}  // namespace oni
