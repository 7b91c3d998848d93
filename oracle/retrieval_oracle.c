/* Retrieval oracle, C twin of oracle/retrieval_oracle.py -- TEST INFRASTRUCTURE ONLY.
 *
 * Same stated rule (see the Python file's header; parity unpinned by the
 * reference, whose search is a substring test at
 * pilott/memory/enhanced_memory.py:110):
 *   bf16 inputs, fp64 accumulation in eight interleaved partial sums (partial j takes
 *   d = j, j+8, ... in ascending order) combined ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)), cosine with
 *   den = sqrt(nq)*sqrt(nm), score = den > 0 ? dot/den : 0, rank by
 *   (score desc, insertion sequence asc), filter as enhanced_memory.py:118-131
 *   and expiry as :20-21.
 * Built with -ffp-contract=off; the bf16*bf16 products are exact in fp64 so
 * contraction could not change a bit anyway.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline double bf2d(uint16_t b) {
    uint32_t u = ((uint32_t)b) << 16;
    float f;
    memcpy(&f, &u, 4);
    return (double)f;
}

static inline double tree8(const double* p) {
    return ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
}

/* scores[q*N + n] for all rows; ld = row stride of m in elements. */
void oracle_exact_scores(const uint16_t* q, int64_t Q, const uint16_t* m, int64_t N,
                         int D, int64_t ld_q, int64_t ld_m, double* out) {
    double* qd = (double*)malloc(sizeof(double) * (size_t)D);
    for (int64_t qi = 0; qi < Q; ++qi) {
        double pq[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int d = 0; d < D; ++d) {
            qd[d] = bf2d(q[qi * ld_q + d]);
            pq[d & 7] += qd[d] * qd[d];
        }
        const double sq = sqrt(tree8(pq));
        for (int64_t n = 0; n < N; ++n) {
            const uint16_t* row = m + n * ld_m;
            double pd[8] = {0, 0, 0, 0, 0, 0, 0, 0}, pm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int d = 0; d < D; ++d) {
                const double v = bf2d(row[d]);
                pd[d & 7] += qd[d] * v;
                pm[d & 7] += v * v;
            }
            const double den = sq * sqrt(tree8(pm));
            out[qi * N + n] = den > 0.0 ? tree8(pd) / den : 0.0;
        }
    }
    free(qd);
}

/* (score desc, seq asc): returns 1 when a ranks before b. */
static inline int before(double sa, int64_t qa, double sb, int64_t qb) {
    return sa > sb || (sa == sb && qa < qb);
}

/* Full search over a slot-ordered store.  seq[n] < 0 marks a never-written slot. */
void oracle_search(const uint16_t* q, int64_t Q, const uint16_t* m, int64_t N, int D,
                   int64_t ld_q, int64_t ld_m, const int64_t* seq, const int32_t* prio,
                   const uint64_t* tagmask, const int64_t* expires_us, int32_t min_prio,
                   uint64_t req_tags, int64_t now_us, int k, int64_t* out_idx,
                   double* out_score) {
    double* sc = (double*)malloc(sizeof(double) * (size_t)N);
    for (int64_t qi = 0; qi < Q; ++qi) {
        oracle_exact_scores(q + qi * ld_q, 1, m, N, D, ld_q, ld_m, sc);
        int64_t* bi = out_idx + qi * k;
        double* bs = out_score + qi * k;
        int have = 0;
        for (int j = 0; j < k; ++j) { bi[j] = -1; bs[j] = -INFINITY; }
        for (int64_t n = 0; n < N; ++n) {
            if (seq[n] < 0) continue;
            if (prio[n] < min_prio) continue;
            if ((tagmask[n] & req_tags) != req_tags) continue;
            if (expires_us[n] != 0 && now_us > expires_us[n]) continue;
            if (have == k && !before(sc[n], seq[n], bs[k - 1], bi[k - 1])) continue;
            int j = have < k ? have : k - 1;
            while (j > 0 && before(sc[n], seq[n], bs[j - 1], bi[j - 1])) {
                bs[j] = bs[j - 1]; bi[j] = bi[j - 1]; --j;
            }
            bs[j] = sc[n]; bi[j] = seq[n];
            if (have < k) ++have;
        }
    }
    free(sc);
}

/* ---- the same search, organised for long corpora (bench.py's parity sample over 1 M - 10 M rows):
 * one pass over a RANGE of rows scores all Q queries against each row, so a row is converted to
 * fp64 and its norm summed once instead of once per query, and callers split the corpus into row
 * ranges across host threads and merge the per-range lists (oracle/retrieval_c.py).  Arithmetic and
 * order are exactly those of oracle_exact_scores / oracle_search above -- eight interleaved partial
 * sums per quantity, the fixed combining tree, den = sqrt(nq)*sqrt(nm), (score desc, seq asc) --
 * and tests/test_oracle_twin.py holds this function to them and to the numpy statement bit for bit.
 * The inner loops are written eight lanes wide so that the compiler can keep one partial sum per
 * vector lane; target_clones picks the widest vector unit of the host at load time (the result
 * does not depend on it: every lane performs the same IEEE operations in the same order). */
#if defined(__x86_64__) && defined(__GNUC__)
#define ORACLE_CLONES __attribute__((target_clones("avx512f", "avx2", "default")))
#else
#define ORACLE_CLONES
#endif

ORACLE_CLONES
void oracle_search_rows(const uint16_t* q, int64_t Q, const uint16_t* m, int64_t n_lo, int64_t n_hi,
                        int D, int64_t ld_q, int64_t ld_m, const int64_t* seq, const int32_t* prio,
                        const uint64_t* tagmask, const int64_t* expires_us, int32_t min_prio,
                        uint64_t req_tags, int64_t now_us, int k, int64_t* out_idx, double* out_score) {
    const int D8 = (D + 7) & ~7;                       /* padded with zeros: they add exactly nothing */
    double* qd = (double*)calloc((size_t)Q * D8, sizeof(double));
    double* sq = (double*)malloc(sizeof(double) * (size_t)Q);
    double* vd = (double*)calloc((size_t)D8, sizeof(double));
    int* have = (int*)calloc((size_t)Q, sizeof(int));
    for (int64_t qi = 0; qi < Q; ++qi) {
        double pq[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int d = 0; d < D; ++d) qd[qi * D8 + d] = bf2d(q[qi * ld_q + d]);
        for (int d = 0; d < D8; d += 8)
            for (int j = 0; j < 8; ++j) pq[j] += qd[qi * D8 + d + j] * qd[qi * D8 + d + j];
        sq[qi] = sqrt(tree8(pq));
        for (int j = 0; j < k; ++j) { out_idx[qi * k + j] = -1; out_score[qi * k + j] = -INFINITY; }
    }
    for (int64_t n = n_lo; n < n_hi; ++n) {
        if (seq[n] < 0) continue;
        if (prio[n] < min_prio) continue;
        if ((tagmask[n] & req_tags) != req_tags) continue;
        if (expires_us[n] != 0 && now_us > expires_us[n]) continue;
        const uint16_t* row = m + n * ld_m;
        double pm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int d = 0; d < D; ++d) vd[d] = bf2d(row[d]);
        for (int d = 0; d < D8; d += 8)
            for (int j = 0; j < 8; ++j) pm[j] += vd[d + j] * vd[d + j];
        const double sm = sqrt(tree8(pm));
        for (int64_t qi = 0; qi < Q; ++qi) {
            const double* qq = qd + qi * D8;
            double pd[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int d = 0; d < D8; d += 8)
                for (int j = 0; j < 8; ++j) pd[j] += qq[d + j] * vd[d + j];
            const double den = sq[qi] * sm;
            const double s = den > 0.0 ? tree8(pd) / den : 0.0;
            int64_t* bi = out_idx + qi * k;
            double* bs = out_score + qi * k;
            if (have[qi] == k && !before(s, seq[n], bs[k - 1], bi[k - 1])) continue;
            int j = have[qi] < k ? have[qi] : k - 1;
            while (j > 0 && before(s, seq[n], bs[j - 1], bi[j - 1])) {
                bs[j] = bs[j - 1]; bi[j] = bi[j - 1]; --j;
            }
            bs[j] = s; bi[j] = seq[n];
            if (have[qi] < k) ++have[qi];
        }
    }
    free(qd); free(sq); free(vd); free(have);
}
