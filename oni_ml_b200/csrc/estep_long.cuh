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
//   gamma' = fma(c e, 1/S, gamma - c old)                       1 dependent fp64 operation
//   psi(gamma') + log beta   interval polynomial (psi_table.h)  table read + 4
//   e = exp(.)               2048-entry table, cubic in 2 levels 6
//   S = sum over topics      every lane adds the K values out of shared memory (one store, K/2
//                            broadcast 16-byte loads, a 5-level tree): ~100 cycles, against ~160
//                            for five shuffle+add levels
//   1/S                      MUFU seed + 3
//
// ~15 dependent fp64 operations (8.4 cycles each on B200) + two shared-memory round trips per
// visit.  Everything that is not on the chain (the bound's sums, phi stores, the requests for
// later rows) fills the issue slots the chain leaves empty; rows and pairs arrive through a
// shared-memory ring filled by cp.async QD visits ahead, as in estep_kernel's ring build.
// Topic k lives in lane k mod 32 (T = ceil(K/32) topics per lane).
//
// The psi table covers gamma in [2^-8, 2^23).  A document that leaves it (alpha below 2^-8, or
// more than 8M tokens on one topic) is run again from its start with digamma_ref in place of the
// table; the flag is looked at once per sweep, so the common path pays two integer
// instructions for it.  The max-shifted (tiny alpha) variant stays with estep_kernel.
#pragma once
#include "estep.cuh"
#include "psi_table.h"

namespace oni {

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

constexpr int LONG_NW = 4;  // warps (documents) per block

template <int T>
struct LongLayout {
    static constexpr int QD = (T == 1) ? 8 : (T == 2) ? 4 : 2;  // visits ahead for the rows
    static constexpr int QT = 2 * QD;                           // ... and for the pairs
    static constexpr size_t psi = 0;
    static constexpr size_t exk = psi + sizeof(double) * PSI_N * PSI_COEF;
    static constexpr size_t lg = exk + sizeof(double) * EXK_N;
    static constexpr size_t red = lg + sizeof(LogEntry) * ONI_MT_N;
    static constexpr size_t lb = red + sizeof(double) * LONG_NW * 2 * 32;
    static constexpr size_t old = lb + sizeof(double) * LONG_NW * QD * T * 32;
    static constexpr size_t tk = old + sizeof(double) * LONG_NW * QD * T * 32;
    static constexpr size_t bytes = tk + sizeof(int2) * LONG_NW * QT * 32;
};

// NSUM: how many per-lane partial sums the visit adds (32, or K when T == 1 and the host
// promises K == NSUM: lanes >= K hold nothing).
template <int T, int NSUM>
__global__ void __launch_bounds__(LONG_NW * 32, 1) estep_long_kernel(const EstepParams p)
{
    using L = LongLayout<T>;
    constexpr int QD = L::QD, QT = L::QT;
    extern __shared__ __align__(16) unsigned char smem[];
    double *const s_psi = reinterpret_cast<double *>(smem + L::psi);
    double *const s_exk = reinterpret_cast<double *>(smem + L::exk);
    LogEntry *const s_lg = reinterpret_cast<LogEntry *>(smem + L::lg);
    {
        // tables: psi | exk are contiguous in shared memory and come from two global arrays
        const double2 *a = reinterpret_cast<const double2 *>(p.psi_tab);
        double2 *d = reinterpret_cast<double2 *>(s_psi);
#pragma unroll 8
        for (int i = threadIdx.x; i < PSI_N * PSI_COEF / 2; i += blockDim.x) d[i] = __ldg(a + i);
        const double2 *b = reinterpret_cast<const double2 *>(p.exk_tab);
        d = reinterpret_cast<double2 *>(s_exk);
#pragma unroll 8
        for (int i = threadIdx.x; i < EXK_N / 2; i += blockDim.x) d[i] = __ldg(b + i);
        const double *src = reinterpret_cast<const double *>(&g_math_tables.lg[0]);
        double *dst = reinterpret_cast<double *>(s_lg);
        for (int i = threadIdx.x; i < 2 * ONI_MT_N; i += blockDim.x) dst[i] = src[i];
    }
    __syncthreads();

    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned full = 0xffffffffu;
    const int K = p.K, KP = p.KP;
    double *const my_red = reinterpret_cast<double *>(smem + L::red) + w * 64;
    double *const my_lb = reinterpret_cast<double *>(smem + L::lb) + w * QD * T * 32 + lane;
    double *const my_old = reinterpret_cast<double *>(smem + L::old) + w * QD * T * 32 + lane;
    int2 *const my_tk = reinterpret_cast<int2 *>(smem + L::tk) + w * QT * 32 + lane;

    bool ok[T];
    int kk[T];
#pragma unroll
    for (int t = 0; t < T; t++) {
        ok[t] = (lane + 32 * t) < K;
        kk[t] = ok[t] ? lane + 32 * t : 0;  // loads of dead slots go to a clamped address
    }
    const double alpha = *p.alpha;
    const double inv_k = 1.0 / (double)K;
    const double lik_const = lgamma(alpha * K) - K * lgamma(alpha);  // lda-inference.c:105-108

    for (;;) {
        unsigned ti = 0;
        if (lane == 0) ti = atomicAdd(p.ticket, 1u);
        ti = __shfl_sync(full, ti, 0);
        if (ti >= (unsigned)p.D) break;
        const int doc = p.order[ti];
        const int64_t beg = p.doc_ptr[doc];
        const int len = (int)(p.doc_ptr[doc + 1] - beg);
        const int2 *const tok = p.tok + beg;
        double *const phi = p.phi + (size_t)beg * KP;
        const double g0 = alpha + p.doc_total[doc] / ((double)K);  // lda-inference.c:37-43
        if (len <= 0) {
            // (an empty document never reaches this kernel; kept so that nothing hangs if one does)
            continue;
        }

        bool slow = false;  // second attempt: digamma_ref instead of the table
        for (;;) {
            double gam[T];
#pragma unroll
            for (int t = 0; t < T; t++) gam[t] = ok[t] ? g0 : 1.0;
            double acc_a = 0.0, acc_b = 0.0;   // sum c phi psi (this lane's topics), sum c log S
            double s_keep = 1.0, c_keep = 0.0;
            double lik_old = 0.0, lik = 0.0;
            int pass = 0;
            unsigned bad = 0;
            unsigned v = 0;                    // visits so far (ring slots are v mod QD / QT)
            int n = 0;                         // pair of this visit
            int nd = QD % len, nt = QT % len;  // pairs of the visits QD and QT ahead
            bool more = true;

            // ---- prime the ring
            cp_async_wait<0>();  // nothing of the previous document may still land in it
#pragma unroll
            for (int i = 0; i < QT; i++) cp_async8(my_tk + i * 32, tok + (i % len));
            cp_async_commit();
            cp_async_wait<0>();
#pragma unroll
            for (int i = 0; i < QD; i++) {
                const int2 ti2 = my_tk[i * 32];
                const double *r = p.logbeta_t + (size_t)ti2.x * KP;
#pragma unroll
                for (int t = 0; t < T; t++) cp_async8(my_lb + (i * T + t) * 32, r + kk[t]);
                cp_async_commit();  // one group per visit
            }

            while (more) {
                // ---- one word visit (lda-inference.c:50-75)
                cp_async_wait<QD - 1>();
                const int sr = v & (QD - 1), st = v & (QT - 1);
                const int2 tk = my_tk[st * 32];
                double lb[T], old[T];
#pragma unroll
                for (int t = 0; t < T; t++) {
                    lb[t] = fmax(my_lb[(sr * T + t) * 32], -2.0e5);  // (exp_k's argument range)
                    // old phi: 1/K during the first sweep (lda-inference.c:42)
                    old[t] = ok[t] ? ((pass == 0) ? inv_k : my_old[(sr * T + t) * 32]) : 0.0;
                }
                {
                    // requests for the visit QD ahead (rows) and QT ahead (pair), into the slots just read
                    const int2 rtk = my_tk[((v + QD) & (QT - 1)) * 32];
                    const double *nlb = p.logbeta_t + (size_t)rtk.x * KP;
                    const double *nph = phi + (size_t)nd * KP;
                    const bool want_old = (pass > 0) || (n + QD >= len);
#pragma unroll
                    for (int t = 0; t < T; t++) {
                        cp_async8(my_lb + (sr * T + t) * 32, nlb + kk[t]);
                        if (want_old) cp_async8(my_old + (sr * T + t) * 32, nph + kk[t]);
                    }
                    cp_async8(my_tk + st * 32, tok + nt);
                    cp_async_commit();
                }
                const double c = (double)tk.y;

                double psi[T], e[T];
                double sl = 0.0;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    double x;
                    if (!slow) {
                        int idx = psi_index(gam[t]);
                        bad |= ((unsigned)idx >= (unsigned)PSI_N) ? 1u : 0u;
                        idx = min(max(idx, 0), PSI_N - 1);
                        const double2 *cf = reinterpret_cast<const double2 *>(s_psi + idx * PSI_COEF);
                        const double2 c01 = cf[0], c23 = cf[1], c45 = cf[2];
                        const double r = gam[t] - psi_centre(gam[t]);
                        const double r2 = r * r;
                        const double q = fma(fma(c45.y, r, c45.x), r2, fma(c23.y, r, c23.x));
                        psi[t] = fma(q, r2, fma(c01.y, r, c01.x));
                        x = fma(q, r2, fma(c01.y, r, c01.x + lb[t]));   // psi + log beta, same depth
                    } else {
                        psi[t] = digamma_ref_lat(gam[t], s_lg);
                        x = fmax(psi[t] + lb[t], -5.0e5);
                    }
                    e[t] = ok[t] ? exp_k(x, s_exk) : 0.0;
                    sl += e[t];
                }
                // S: every lane adds all the partial sums
                double *rb = my_red + (v & 1) * 32;  // two buffers: one __syncwarp per visit
                rb[lane] = sl;
                __syncwarp();
                double vs[NSUM];
#pragma unroll
                for (int i = 0; i < NSUM / 2; i++) {
                    const double2 two = reinterpret_cast<const double2 *>(rb)[i];
                    vs[2 * i] = two.x; vs[2 * i + 1] = two.y;
                }
                const double S = tree_sum<NSUM>(vs);
                const double rs = rcp_pos(S);
                double a = 0.0;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    const double ph = e[t] * rs;
                    // gamma + c (phi - old), lda-inference.c:69-70, one operation after 1/S
                    gam[t] = fma(c * e[t], rs, fma(-c, old[t], gam[t]));
                    a = fma(ph, psi[t], a);
                    if (ok[t]) phi[(size_t)n * KP + lane + 32 * t] = ph;
                }
                acc_a = fma(c, a, acc_a);
                // log S once per visit: lane (v mod 32) keeps S, the warp takes 32 logs at once
                if ((int)(v & 31) == lane) { s_keep = S; c_keep = c; }
                v++; n++; nd++; nt++;
                if (nd == len) nd = 0;
                if (nt == len) nt = 0;
                if ((v & 31) == 0 || n == len) {
                    acc_b = fma(c_keep, log_tab(s_keep, s_lg), acc_b);
                    s_keep = 1.0; c_keep = 0.0;
                }
                if (n < len) continue;

                // ---- end of a sweep: the bound (lda-inference.c:93-127, folded as in estep.cuh)
                n = 0;
                double gs = 0.0, part = 0.0;
#pragma unroll
                for (int t = 0; t < T; t++) {
                    gs += ok[t] ? gam[t] : 0.0;
                    part += ok[t] ? lgamma_pos(gam[t], s_lg) : 0.0;
                }
                part += acc_b - acc_a;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    gs += __shfl_xor_sync(full, gs, o);
                    part += __shfl_xor_sync(full, part, o);
                }
                lik = (lik_const - lgamma_pos(gs, s_lg)) + part;
                pass++;
                if (!slow && __any_sync(full, bad != 0)) break;  // left the table: start over, slowly
                const double conv = (lik_old - lik) / lik_old;   // lda-inference.c:79-80, :46-47
                lik_old = lik;
                more = (conv > p.var_converged) && ((pass < p.var_max_iter) || (p.var_max_iter == -1));
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
                        ds += digamma_ref(gam[t], s_lg);
                    }
                }
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) ds += __shfl_xor_sync(full, ds, o);
                if (lane == 0) {
                    p.doc_lik[doc] = lik;
                    p.doc_ass[doc] = ds - K * digamma_ref(gs, s_lg);
                    p.doc_iters[doc] = pass;
                }
            }
            if (!more) break;  // finished
            slow = true;       // (left through the table check)
        }
    }
    cp_async_wait<0>();
}

}  // namespace oni
