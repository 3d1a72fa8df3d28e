// estep_long.cuh -- the E-step of LONG documents: one document per warp, built for a short
// dependency chain per word visit.
//
// Same computation as estep_kernel (lda_inference + compute_likelihood, oni-lda-c/
// lda-inference.c:27-127): the words of a document are visited strictly in order and gamma /
// digamma(gamma) move after every word, so a document of N pairs is a serial chain of
// N x sweeps visits and a shard's longest document bounds its E-step from below -- it IS the
// E-step as soon as the shard is small (one flow day split over 8 GPUs: ~12 ms of bulk work
// against a 4000-pair document).  estep_kernel's one-document-per-warp build spends ~820 cycles
// per visit on that chain (profiles/r2_longdoc.txt); this kernel is organised around the chain:
//
//   gamma' = fma(c e r0, w, ...)  (the Newton correction of 1/S folded into the update)
//   psi(gamma') + log beta   interval polynomial (psi_table.h)  table read + 4 dependent fp64 operations
//   e = exp(.)               2048-entry table, cubic in 2 levels 6
//   S = sum over topics      every lane adds the K values out of shared memory (one store, K/2
//                            broadcast 16-byte loads, a 5-level tree): ~100 cycles, against ~160
//                            for five shuffle+add levels
//   1/S                      MUFU seed + 3
//
// ~15 dependent fp64 operations (8.4 cycles each on B200, profiles/micro/lat.cu), a table index
// five integer operations deep and two shared-memory round trips per visit: ~360-390 cycles measured
// (0.19-0.20 us; the one-warp build of estep_kernel: 820).  A lone warp issues in order and pays
// ~20 cycles for every branch, so the visit body is straight-line code and its SOURCE ORDER is the
// schedule: the new gamma's table reads first, then -- in their shadow -- the previous visit's phi
// store and bound sums, the requests for later rows and the next visit's ring reads; dead topic
// slots are handled by zero-fill copies and a zero count instead of tests.  Rows and pairs arrive
// through a shared-memory ring filled by cp.async QD visits ahead.  Topic k lives in lane k mod 32
// (T = ceil(K/32) per lane).  History, per-instruction stall samples and what was tried and
// dropped (twelve warps per SM, a helper warp per document, a tensor-core sum): profiles/r2_longdoc.txt.
//
// The psi table covers gamma in [2^-8, 2^23).  A document that leaves it (alpha below 2^-8, or
// more than 8M tokens on one topic) is run again from its start by the SLOW instantiation
// (digamma_ref in place of the table); the flag is looked at once per sweep.  The max-shifted
// (tiny alpha) variant stays with estep_kernel.
#pragma once
#include "estep.cuh"
#include "psi_table.h"

namespace oni {

// (double)n for 0 <= n < 2^32 as one fixed-latency addition: I2F.F64 goes through a scoreboard the
// compiler shares with the asynchronous copies, and a lone warp then waits for those
__device__ __forceinline__ double count_to_double(int n)
{
    return __hiloint2double(0x43300000, n) - 4503599627370496.0;  // 2^52 + n, minus 2^52
}

// Sum of one double per lane, delivered to every lane, by the fp64 tensor instruction: with B = 1 an
// m8n8k4 product adds the four values of each quad (row g of A = the lanes of quad g), one
// exchange brings quad sums k and k+4 to lane position k of every quad, and a second product adds
// those four.  2 DMMA (26 cycles each on B200) + one shuffle + one addition: ~85 cycles and 8
// instructions on paper, against ~105 cycles and 32 instructions for the exchange through shared
// memory with a 20-value tree.  MEASURED (build -DONI_LONG_SUM_DMMA): 0.206-0.208 us per visit
// against 0.204 for the shared-memory tree beside a 125k-document shard -- no gain, so the tree
// is the default and this stays as the record of the experiment.
__device__ __forceinline__ double warp_sum_dmma(double v, int lane)
{
    const double one = 1.0, zero = 0.0;
    double q0, q1;
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                 : "=d"(q0), "=d"(q1) : "d"(v), "d"(one), "d"(zero), "d"(zero));
    // quad sum g is now in every lane of quad g; fetch sums (lane & 3) and (lane & 3) + 4
    const int src = (lane & 3) * 4;
    int lo = __double2loint(q0), hi = __double2hiint(q0), alo, ahi, blo, bhi;
    asm volatile("shfl.sync.idx.b32 %0, %4, %6, 0x1f, 0xffffffff;\n\t"
                 "shfl.sync.idx.b32 %1, %5, %6, 0x1f, 0xffffffff;\n\t"
                 "shfl.sync.idx.b32 %2, %4, %7, 0x1f, 0xffffffff;\n\t"
                 "shfl.sync.idx.b32 %3, %5, %7, 0x1f, 0xffffffff;"
                 : "=r"(alo), "=r"(ahi), "=r"(blo), "=r"(bhi) : "r"(lo), "r"(hi), "r"(src), "r"(src + 16));
    const double u = __hiloint2double(ahi, alo) + __hiloint2double(bhi, blo);
    double s0, s1;
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%4,%5};"
                 : "=d"(s0), "=d"(s1) : "d"(u), "d"(one), "d"(zero), "d"(zero));
    return s0 + 0.0 * (q1 + s1);  // (the second output registers are part of the instruction)
}

template <int N>
__device__ __forceinline__ double tree_sum(double *v)
{
    if constexpr (N == 1) {
        return v[0];
    } else {
        constexpr int H = (N + 1) / 2;
#pragma unroll
        for (int i = 0; i < N / 2; i++) v[i] += v[i + H];
        return tree_sum<H>(v);
    }
}

// 8-byte asynchronous copy that writes zeros instead when `bytes` is 0
__device__ __forceinline__ void cp_async8_zfill(void *smem_dst, const void *gsrc, int bytes)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gsrc), "r"(bytes) : "memory");
}

// Warps (= documents) per block: one per scheduler.  More chains per scheduler were measured and
// dropped: twelve warps per SM run at ~1360 cycles per visit each (the table reads, the sum's
// broadcast loads and the ring take ~50 shared-memory wavefronts per visit, so twelve chains
// saturate the SM's load/store pipe), a worse throughput than four warps at 383.
constexpr int LONG_NW = 4;
template <int T, int NW>
struct LongLayout {
    static constexpr int QD = (T == 1) ? 8 : (T == 2) ? 4 : 2;  // visits ahead for the rows
    static constexpr int QT = 2 * QD;                           // ... and for the pairs
    static constexpr size_t psi = 0;
    static constexpr size_t exk = psi + sizeof(double) * PSI_N * PSI_COEF;
    static constexpr size_t lg = exk + sizeof(double) * EXK_N;
    static constexpr size_t red = lg + sizeof(LogEntry) * ONI_MT_N;
    static constexpr size_t keep = red + sizeof(double) * NW * 2 * 32;   // (S, c) of 32 visits
    static constexpr size_t lb = keep + sizeof(double) * NW * 2 * 32;
    static constexpr size_t old = lb + sizeof(double) * NW * QD * T * 32;
    static constexpr size_t tk = old + sizeof(double) * NW * QD * T * 32;
    static constexpr size_t bytes = tk + sizeof(int2) * NW * QT * 32;
};

struct LongShared {
    const double *psi, *exk;
    const LogEntry *lg;
    double *red;      // this warp's [2][32]
    double *keep;     // this warp's [32][2]
    double *lb, *old; // this warp's rings, already offset by the lane
    int2 *tk;
};

// One document.  Returns true when gamma left the psi table (SLOW == false only): the caller
// then runs the document again with SLOW == true.
template <int T, int NSUM, bool SLOW>
__device__ __forceinline__ bool long_document(const EstepParams &p, const LongShared &sh, int doc, int lane,
                                              const bool (&ok)[T], const int (&kk)[T], double alpha,
                                              double inv_k, double lik_const)
{
    constexpr int QD = LongLayout<T, 4>::QD, QT = LongLayout<T, 4>::QT;
#ifndef ONI_LONG_SUM_DMMA
    constexpr bool MASK = !(T == 1 && NSUM < 32);  // dead slots could reach the sum
#else
    constexpr bool MASK = true;                    // every lane's value enters the tensor-core sum
#endif
    const unsigned full = 0xffffffffu;
    const int K = p.K, KP = p.KP;
    const int64_t beg = p.doc_ptr[doc];
    const int len = (int)(p.doc_ptr[doc + 1] - beg);
    const int2 *const tok = p.tok + beg;
    double *const phi = p.phi + (size_t)beg * KP;
    const double g0 = alpha + p.doc_total[doc] / ((double)K);  // lda-inference.c:37-43

    double gam[T];
    int okb[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        gam[t] = ok[t] ? g0 : 1.0;
        okb[t] = ok[t] ? 8 : 0;
    }
    double acc_a = 0.0, acc_b = 0.0;   // sum c phi psi (this lane's topics), sum c log S
    double lik_old = 0.0, lik = 0.0;
    int pass = 0;
    unsigned bad = 0;                  // largest table index seen (as unsigned: below the table = huge)
    unsigned v = 0;                    // visits so far

    // ---- prime the ring.  old-phi slots start at 1/K (lda-inference.c:42) and are only
    // overwritten by rows of later sweeps; dead topic slots read as zero throughout.
    cp_async_wait<0>();  // nothing of the previous document may still land in it
#pragma unroll
    for (int i = 0; i < QD * T; i++) sh.old[i * 32] = ok[i % T] ? inv_k : 0.0;
#pragma unroll
    for (int i = 0; i < QT; i++) cp_async8(sh.tk + i * 32, tok + (i % len));
    cp_async_commit();
    cp_async_wait<0>();
#pragma unroll
    for (int i = 0; i < QD; i++) {
        const int2 ti = sh.tk[i * 32];
        const double *r = p.logbeta_t + (size_t)ti.x * KP;
#pragma unroll
        for (int t = 0; t < T; t++) cp_async8_zfill(sh.lb + (i * T + t) * 32, r + kk[t], okb[t]);
        cp_async_commit();  // one group per visit
    }
    // positions of the requests: rows go out for the visit QD ahead, pairs for the visit QT ahead.
    // Offsets are 32-bit element counts (a shard holds < 2^31 pairs and the launcher checks
    // pairs * KP < 2^32 for this kernel's documents).
    const unsigned kp = (unsigned)KP, row_bytes = kp * 8u;
    int nd = QD % len, nt = QT % len;
    unsigned od = (unsigned)nd * kp;  // phi offset of the row QD ahead
    bool wrapped = (QD >= len);       // ... which belongs to a later sweep: its phi exists
    const char *const lb_base = reinterpret_cast<const char *>(p.logbeta_t);
    // registers of the visit about to run (read one visit ahead): its pair count as a double, its
    // two rows, and the word of the visit QD ahead (whose row this visit requests)
    cp_async_wait<QD - 1>();
    double c = count_to_double(sh.tk[0].y);
    int rword = sh.tk[(QD & (QT - 1)) * 32].x;
    double lb[T], old[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        lb[t] = sh.lb[t * 32];
        old[t] = sh.old[t * 32];
    }
    double2 *const keep = reinterpret_cast<double2 *>(sh.keep);
    // the psi-table coefficients of the interval each topic's gamma is in (none yet)
    double2 c01[T], c23[T], c45[T];
    unsigned held[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        held[t] = 0xffffffffu;
        c01[t] = c23[t] = c45[t] = make_double2(0.0, 0.0);
    }
    const unsigned psi_saddr = (unsigned)__cvta_generic_to_shared(sh.psi);

    for (;;) {  // sweeps
        unsigned pr = 0;  // phi offset of the row of this visit
        for (int n = 0; n < len;) {
            const int m = min(32, len - n);
            // what the previous visit leaves for the shadow of this visit's table read
            double e_p[T], psi_p[T], cl_p[T], rs_p = 0.0, S_p = 1.0, c_p = 0.0;
            unsigned pr_p = 0;
#pragma unroll
            for (int t = 0; t < T; t++) e_p[t] = psi_p[t] = cl_p[t] = 0.0;
#pragma unroll 4
            for (int i = 0; i < m; i++) {
                // ---- one word visit (lda-inference.c:50-75); straight-line.  A lone warp issues
                // in order, so the order below is the schedule: first what the chain needs next
                // (the table rows of the new gamma), then, while those are on their way, the
                // leftovers of the previous visit and the requests for later ones.
                double r[T];
#pragma unroll
                for (int t = 0; t < T; t++) {
                    if constexpr (!SLOW) {
                        const unsigned idx_raw = (unsigned)psi_index(gam[t]);
                        bad = max(bad, idx_raw);  // (negative indices are huge as unsigned)
                        const unsigned idx = min(idx_raw, (unsigned)(PSI_N - 1));
                        // The six coefficients stay in registers from visit to visit and are read again
                        // only when gamma has moved to another interval (1.6 % wide: after the first
                        // sweep or two a rare event), by loads predicated per lane -- no branch, and the
                        // polynomial below does not wait for shared memory when nothing was read.
                        const unsigned addr = psi_saddr + idx * (unsigned)(PSI_COEF * sizeof(double));
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %6, %7;\n\t"
                                     "@p ld.shared.v2.f64 {%0, %1}, [%8];\n\t"
                                     "@p ld.shared.v2.f64 {%2, %3}, [%8+16];\n\t"
                                     "@p ld.shared.v2.f64 {%4, %5}, [%8+32];\n\t}"
                                     : "+d"(c01[t].x), "+d"(c01[t].y), "+d"(c23[t].x), "+d"(c23[t].y), "+d"(c45[t].x),
                                       "+d"(c45[t].y)
                                     : "r"(idx), "r"(held[t]), "r"(addr));
                        held[t] = idx;
                        r[t] = gam[t] - psi_centre(gam[t]);
                    }
                }
                // leftovers of the previous visit: phi row, the bound's sums, (S, c) for the logs
                {
                    double *const prow = phi + (pr_p + (unsigned)lane);
#pragma unroll
                    for (int t = 0; t < T; t++) {
                        const double ph = e_p[t] * rs_p;
                        acc_a = fma(cl_p[t] * ph, psi_p[t], acc_a);
                        if (ok[t] && i > 0) prow[32 * t] = ph;
                    }
                    if (lane == 0 && i > 0) keep[i - 1] = make_double2(S_p, c_p);
                }
                // requests into the slots of this visit (its values are in registers already)
                const int sr = v & (QD - 1), st = v & (QT - 1);
                {
                    const double *nlb = reinterpret_cast<const double *>(lb_base + (size_t)((unsigned)rword * row_bytes));
                    const double *nph = phi + od;
#pragma unroll
                    for (int t = 0; t < T; t++) {
                        cp_async8_zfill(sh.lb + (sr * T + t) * 32, nlb + kk[t], okb[t]);
                        if (wrapped) cp_async8_zfill(sh.old + (sr * T + t) * 32, nph + kk[t], okb[t]);
                    }
                    cp_async8(sh.tk + st * 32, tok + nt);
                    cp_async_commit();
                }
                // the registers of the next visit
                cp_async_wait<QD - 1>();
                const double c_n = count_to_double(sh.tk[((v + 1) & (QT - 1)) * 32].y);
                const int rword_n = sh.tk[((v + 1 + QD) & (QT - 1)) * 32].x;
                double lb_n[T], old_n[T];
#pragma unroll
                for (int t = 0; t < T; t++) {
#ifdef ONI_RING_CHECK_BREAK
                    lb_n[t] = sh.lb[(((v + 2) & (QD - 1)) * T + t) * 32];  // negative control of the self-check: the wrong slot
#else
                    lb_n[t] = sh.lb[(((v + 1) & (QD - 1)) * T + t) * 32];
#endif
                    old_n[t] = sh.old[(((v + 1) & (QD - 1)) * T + t) * 32];
                }

                // positions of the next visit's requests (integer work, out of the chain's way)
                const unsigned pr_now = pr;
                pr += kp;
                nd++; nt++; od += kp;
                if (nd == len) { nd = 0; od = 0; wrapped = true; }
                if (nt == len) nt = 0;

#ifdef ONI_RING_CHECK
                // Self-check build (profiles/ring_check.sh; compute-sanitizer is closed on this pool):
                // what the ring delivered for THIS visit against plain loads of the same words --
                // a copy that lands late, lands in the wrong slot or is overwritten early shows here.
                {
                    const int2 tk_direct = tok[pr_now / kp];
                    bool okv = (c == (double)tk_direct.y);
#pragma unroll
                    for (int t = 0; t < T; t++) {
                        if (!ok[t]) continue;
                        okv = okv && (lb[t] == p.logbeta_t[(size_t)tk_direct.x * KP + lane + 32 * t]);
                        const double want_old = (pass == 0) ? inv_k : phi[pr_now + lane + 32 * t];
                        okv = okv && (old[t] == want_old);
                    }
                    if (!okv) atomicExch(p.nan_flag, 2);
                }
#endif
                double cl[T], psi[T], e[T];
                double sl = 0.0;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    cl[t] = ok[t] ? c : 0.0;  // dead slots never move
                    double x;
                    if constexpr (!SLOW) {
                        const double r2 = r[t] * r[t];
                        const double q = fma(fma(c45[t].y, r[t], c45[t].x), r2, fma(c23[t].y, r[t], c23[t].x));
                        x = fma(q, r2, fma(c01[t].y, r[t], c01[t].x + lb[t]));   // psi + log beta in one chain
                        psi[t] = x - lb[t];
                    } else {
                        psi[t] = digamma_ref_lat(gam[t], sh.lg);
                        x = fmax(psi[t] + lb[t], -5.0e5);
                    }
                    e[t] = exp_k(x, sh.exk);
                    if (MASK) e[t] = ok[t] ? e[t] : 0.0;
                    sl += e[t];
                }
                // S in every lane (lanes without a live topic contribute zero)
#ifndef ONI_LONG_SUM_DMMA
                double *rb = sh.red + (v & 1) * 32;  // two buffers: one __syncwarp per visit
                rb[lane] = sl;
                __syncwarp();
                double vs[NSUM];
#pragma unroll
                for (int j = 0; j < NSUM / 2; j++) {
                    const double2 two = reinterpret_cast<const double2 *>(rb)[j];
                    vs[2 * j] = two.x; vs[2 * j + 1] = two.y;
                }
                const double S = tree_sum<NSUM>(vs);
#else
                const double S = warp_sum_dmma(sl, lane);
#endif
                // 1/S = r0 (1 + w), w = err + err^2; gamma + c (phi - old) (lda-inference.c:69-70)
                // takes r0 and w directly, one operation sooner than through 1/S
                double r0;
                asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r0) : "d"(S));
                const double err = fma(-S, r0, 1.0);
                const double wv = fma(err, err, err);
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const double pe = (cl[t] * e[t]) * r0;
                    gam[t] = fma(pe, wv, pe + fma(-cl[t], old[t], gam[t]));
                }
                // ---- hand over to the next visit
                rs_p = fma(r0, wv, r0); S_p = S; c_p = c; pr_p = pr_now;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    e_p[t] = e[t]; psi_p[t] = psi[t]; cl_p[t] = cl[t];
                    lb[t] = lb_n[t]; old[t] = old_n[t];
                }
                c = c_n; rword = rword_n;
                v++;
            }
            {
                // leftovers of the chunk's last visit
                double *const prow = phi + (pr_p + (unsigned)lane);
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const double ph = e_p[t] * rs_p;
                    acc_a = fma(cl_p[t] * ph, psi_p[t], acc_a);
                    if (ok[t]) prow[32 * t] = ph;
                }
                if (lane == 0) keep[m - 1] = make_double2(S_p, c_p);
            }
            __syncwarp();
            const double2 kp2 = (lane < m) ? keep[lane] : make_double2(1.0, 0.0);
            __syncwarp();
            const double s_keep = kp2.x, c_keep = kp2.y;
            acc_b = fma(c_keep, log_tab(s_keep, sh.lg), acc_b);
            n += m;
        }

        // ---- end of a sweep: the bound (lda-inference.c:93-127, folded as in estep.cuh)
        double gs = 0.0, part = 0.0;
#pragma unroll
        for (int t = 0; t < T; t++) {
            gs += ok[t] ? gam[t] : 0.0;
            part += ok[t] ? lgamma_pos(gam[t], sh.lg) : 0.0;
        }
        part += acc_b - acc_a;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            gs += __shfl_xor_sync(full, gs, o);
            part += __shfl_xor_sync(full, part, o);
        }
        lik = (lik_const - lgamma_pos(gs, sh.lg)) + part;
        pass++;
        if (!SLOW && __any_sync(full, bad >= (unsigned)PSI_N)) return true;  // left the table: start over, slowly
        const double conv = (lik_old - lik) / lik_old;         // lda-inference.c:79-80, :46-47
        lik_old = lik;
        bool more = (conv > p.var_converged) && ((pass < p.var_max_iter) || (p.var_max_iter == -1));
        if (lik != lik) {  // assert(!isnan(likelihood)), lda-inference.c:78
            if (lane == 0) atomicExch(p.nan_flag, 1);
            more = false;
        }
        acc_a = acc_b = 0.0;
        if (more) continue;

        // ---- document done (lda-estimate.c:40-46)
        double ds = 0.0;
#pragma unroll
        for (int t = 0; t < T; t++) {
            if (ok[t]) {
                p.gamma[(size_t)doc * K + lane + 32 * t] = gam[t];
                ds += digamma_ref(gam[t], sh.lg);
            }
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) ds += __shfl_xor_sync(full, ds, o);
        if (lane == 0) {
            p.doc_lik[doc] = lik;
            p.doc_ass[doc] = ds - K * digamma_ref(gs, sh.lg);
            p.doc_iters[doc] = pass;
        }
        return false;
    }
}

// NSUM: how many per-lane partial sums the visit adds (32, or K when T == 1 and the host
// promises K == NSUM: lanes >= K hold nothing).
template <int T, int NSUM, int NW>
__global__ void __launch_bounds__(NW * 32, 1) estep_long_kernel(const EstepParams p)
{
    using L = LongLayout<T, NW>;
    constexpr int QD = L::QD, QT = L::QT;
    extern __shared__ __align__(16) unsigned char smem[];
    {
        // tables: psi | exk are contiguous in shared memory and come from two global arrays
        const double2 *a = reinterpret_cast<const double2 *>(p.psi_tab);
        double2 *d = reinterpret_cast<double2 *>(smem + L::psi);
#pragma unroll 4
        for (int i = threadIdx.x; i < PSI_N * PSI_COEF / 2; i += blockDim.x) d[i] = __ldg(a + i);
        const double2 *b = reinterpret_cast<const double2 *>(p.exk_tab);
        d = reinterpret_cast<double2 *>(smem + L::exk);
#pragma unroll 4
        for (int i = threadIdx.x; i < EXK_N / 2; i += blockDim.x) d[i] = __ldg(b + i);
        const double *src = reinterpret_cast<const double *>(&g_math_tables.lg[0]);
        double *dst = reinterpret_cast<double *>(smem + L::lg);
        for (int i = threadIdx.x; i < 2 * ONI_MT_N; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    LongShared sh;
    sh.psi = reinterpret_cast<const double *>(smem + L::psi);
    sh.exk = reinterpret_cast<const double *>(smem + L::exk);
    sh.lg = reinterpret_cast<const LogEntry *>(smem + L::lg);
    sh.red = reinterpret_cast<double *>(smem + L::red) + w * 64;
    sh.keep = reinterpret_cast<double *>(smem + L::keep) + w * 64;
    sh.lb = reinterpret_cast<double *>(smem + L::lb) + w * QD * T * 32 + lane;
    sh.old = reinterpret_cast<double *>(smem + L::old) + w * QD * T * 32 + lane;
    sh.tk = reinterpret_cast<int2 *>(smem + L::tk) + w * QT * 32 + lane;

    bool ok[T];
    int kk[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        ok[t] = (lane + 32 * t) < p.K;
        kk[t] = ok[t] ? lane + 32 * t : 0;  // copies of dead slots read nothing (zero fill)
    }
    const double alpha = *p.alpha;
    const double inv_k = 1.0 / (double)p.K;
    const double lik_const = lgamma(alpha * p.K) - p.K * lgamma(alpha);  // lda-inference.c:105-108

    for (;;) {
        unsigned ti = 0;
        if (lane == 0) ti = atomicAdd(p.ticket, 1u);
        ti = __shfl_sync(0xffffffffu, ti, 0);
        if (ti >= (unsigned)p.D) break;
        const int doc = p.order[ti];
        if (p.doc_ptr[doc + 1] <= p.doc_ptr[doc]) continue;  // (empty documents never come here)
        if (long_document<T, NSUM, false>(p, sh, doc, lane, ok, kk, alpha, inv_k, lik_const))
            long_document<T, NSUM, true>(p, sh, doc, lane, ok, kk, alpha, inv_k, lik_const);
    }
    cp_async_wait<0>();
}

}  // namespace oni
