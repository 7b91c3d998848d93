// finish.cu -- K4: merge the per-slot candidate lists, re-score the survivors exactly in
// fp64 in the oracle's order, sort by (score desc, insertion sequence asc), gap-check.
// Also: the brute-force exact search (fallback / large k) and the multi-GPU top-k merge.
#include "common.cuh"
#include "internal.h"

namespace pilott {

// (score desc, seq asc): true when a ranks strictly before b.
__device__ __forceinline__ bool ranks_before(double sa, int64_t qa, double sb, int64_t qb) {
    return sa > sb || (sa == sb && qa < qb);
}

__device__ __forceinline__ double shfl_xor_d(double v, int m) {
    return __longlong_as_double((long long)shfl_xor64((uint64_t)__double_as_longlong(v), m));
}

// Bitonic sort of (score, seq) pairs, one per lane; lane 0 ends with the best.
__device__ __forceinline__ void warp_sort_pairs(double& s, int64_t& q, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const double os = shfl_xor_d(s, j);
            const int64_t oq = (int64_t)shfl_xor64((uint64_t)q, j);
            const bool desc_block = ((lane & k) == 0);        // best-first in this block
            const bool lower = (lane & j) == 0;
            const bool want_best = (lower == desc_block);
            const bool other_better = ranks_before(os, oq, s, q);
            if (want_best == other_better) { s = os; q = oq; }
        }
    }
}

// THE STATED SCORING RULE (oracle/retrieval_oracle.py): fp64, eight interleaved partial sums --
// partial j accumulates the terms d = j, j+8, j+16, ... in ascending order -- combined as
// ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)).  bf16*bf16 products are exact in fp64, so FMA == mul+add.
// Eight independent chains of length D/8 instead of one of length D: the same determinism with
// an eighth of the dependent latency.  Padding columns are zeros and add exactly nothing.
struct Acc8 {
    double p[8];
    __device__ __forceinline__ Acc8() {
#pragma unroll
        for (int j = 0; j < 8; ++j) p[j] = 0.0;
    }
    __device__ __forceinline__ double total() const {
        return __dadd_rn(__dadd_rn(__dadd_rn(p[0], p[1]), __dadd_rn(p[2], p[3])),
                         __dadd_rn(__dadd_rn(p[4], p[5]), __dadd_rn(p[6], p[7])));
    }
};

// One 16-byte chunk = elements d .. d+7 (d a multiple of 8): element e feeds partial e.
__device__ __forceinline__ void acc_chunk(const uint4& qv, const uint4& mv, Acc8& dot, Acc8& nq, Acc8& nm) {
    const uint32_t qw[4] = {qv.x, qv.y, qv.z, qv.w}, mw[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const double q0 = (double)__uint_as_float(qw[j] << 16);
        const double q1 = (double)__uint_as_float(qw[j] & 0xffff0000u);
        const double m0 = (double)__uint_as_float(mw[j] << 16);
        const double m1 = (double)__uint_as_float(mw[j] & 0xffff0000u);
        dot.p[2 * j] = __fma_rn(q0, m0, dot.p[2 * j]);
        nq.p[2 * j] = __fma_rn(q0, q0, nq.p[2 * j]);
        nm.p[2 * j] = __fma_rn(m0, m0, nm.p[2 * j]);
        dot.p[2 * j + 1] = __fma_rn(q1, m1, dot.p[2 * j + 1]);
        nq.p[2 * j + 1] = __fma_rn(q1, q1, nq.p[2 * j + 1]);
        nm.p[2 * j + 1] = __fma_rn(m1, m1, nm.p[2 * j + 1]);
    }
}

__device__ __forceinline__ double cosine_from(const Acc8& dot, const Acc8& nq, const Acc8& nm, double* nq_out) {
    const double nqt = nq.total();
    if (nq_out) *nq_out = nqt;
    const double den = __dmul_rn(__dsqrt_rn(nqt), __dsqrt_rn(nm.total()));
    return den > 0.0 ? __ddiv_rn(dot.total(), den) : 0.0;
}

// Exact cosine of one query / one row.  Every lane computes its own row; the query is read
// through the read-only path (same address in all lanes -> broadcast).
__device__ __forceinline__ double exact_cosine(const __nv_bfloat16* __restrict__ qrow,
                                               const __nv_bfloat16* __restrict__ mrow, int ld,
                                               double* nq_out = nullptr) {
    Acc8 dot, nq, nm;
    for (int c = 0; c < ld; c += 32) {          // ld is a multiple of 64; 8 loads in flight per step
        uint4 qv[4], mv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            qv[u] = __ldg(reinterpret_cast<const uint4*>(qrow + c + 8 * u));
            mv[u] = __ldg(reinterpret_cast<const uint4*>(mrow + c + 8 * u));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) acc_chunk(qv[u], mv[u], dot, nq, nm);
    }
    return cosine_from(dot, nq, nm, nq_out);
}

// Same arithmetic, operands staged in shared memory (generic pointers).
__device__ __forceinline__ double exact_cosine_smem(const unsigned char* qrow, const unsigned char* mrow,
                                                    int ld, double* nq_out) {
    Acc8 dot, nq, nm;
    for (int c = 0; c < ld; c += 8)
        acc_chunk(*reinterpret_cast<const uint4*>(qrow + 2 * c), *reinterpret_cast<const uint4*>(mrow + 2 * c), dot, nq, nm);
    return cosine_from(dot, nq, nm, nq_out);
}

// Sorts a bitonic sequence (one key per lane) into descending order: 5 compare-exchange steps.
__device__ __forceinline__ uint64_t warp_bitonic_merge_desc(uint64_t v, int lane) {
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor64(v, j);
        const bool lower = (lane & j) == 0;
        const uint64_t mn = v < o ? v : o, mx = v < o ? o : v;
        v = lower ? mx : mn;
    }
    return v;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}

constexpr int FIN_WARPS = 4;      // warps per query in finish_kernel
constexpr int FIN_MAXLOADS = 8;   // 32-entry loads each warp prefetches at once

// One CTA (4 warps) per query.  Warps merge disjoint subsets of the candidate lists in
// parallel (all their loads issued up front), warp 0 merges the four pools, re-scores the
// survivors exactly and writes the result.
__global__ void __launch_bounds__(FIN_WARPS * 32)
finish_kernel(const uint64_t* __restrict__ partials, int slots_alloc, int list_len, int pool_n, int rows_per_group,
              int lists_base, int groups_rem, int lists_extra,
              const __nv_bfloat16* __restrict__ qb,
              const __nv_bfloat16* __restrict__ vec, int ld, int64_t nq, int64_t capacity,
              int64_t head, int64_t base_seq, int64_t seq_mul, int64_t seq_add, int k, double err,
              int64_t* __restrict__ out_idx,
              float* __restrict__ out_score, double* __restrict__ out_score64,
              uint8_t* __restrict__ flags, int32_t* __restrict__ inexact, int stage_rows,
              const uint32_t* __restrict__ gthr, unsigned char* const* __restrict__ peers, int world, int rank,
              int64_t block_bytes, int64_t flags_off, int64_t step, uint32_t* __restrict__ done_counter,
              unsigned long long* __restrict__ flagged_total) {
    extern __shared__ __align__(16) unsigned char s_rows[];     // [pool_n + 1][ld*2 + 16] when stage_rows
    __shared__ uint64_t s_pool[FIN_WARPS][32];
    __shared__ double s_part[3][FIN_WARPS][32];              // per-warp pair sums of (dot, nq, nm) per candidate
    __shared__ uint64_t s_bound[FIN_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t q = blockIdx.x;
    // 1. top-32 of this warp's lists by approximate key.  `bound` = best key that is NOT in the
    //    pool: a FULL list (all list_len entries used) may have dropped rows no better than its
    //    own worst entry; every merge drops the elementwise minima of its two inputs, and the
    //    largest of those is tracked in `drop` (tighter than "the pool's worst entry", and what
    //    lets k = 32 = pool size be proven).
    const int per_load = 32 / list_len;                    // lists fetched per 32-lane load (1 or 2)
    const int slots = lists_base + ((q / rows_per_group) < groups_rem ? lists_extra : 0);   // lists this query has
    const uint64_t* p = partials + q * (int64_t)slots_alloc * list_len;
    const int nloads = (slots + per_load - 1) / per_load;
    uint64_t best = 0ull, bound = 0ull, drop = 0ull;
    for (int i0 = warp; i0 < nloads; i0 += FIN_WARPS * FIN_MAXLOADS) {
        uint64_t x[FIN_MAXLOADS];
#pragma unroll
        for (int u = 0; u < FIN_MAXLOADS; ++u) {           // all loads in flight before any sort
            const int i = i0 + u * FIN_WARPS;
            const int li = i * per_load + lane / list_len; // list this lane reads from
            x[u] = (i < nloads && li < slots) ? __ldcg(p + (int64_t)li * list_len + (lane % list_len)) : 0ull;
        }
#pragma unroll
        for (int u = 0; u < FIN_MAXLOADS; ++u) {
            if (i0 + u * FIN_WARPS >= nloads) break;
            uint64_t mn = x[u];                            // per-list minimum (segmented butterfly)
            for (int j = list_len / 2; j > 0; j >>= 1) {
                const uint64_t o = shfl_xor64(mn, j);
                mn = mn < o ? mn : o;
            }
            if (mn != 0ull && mn > bound) bound = mn;      // list is full: it may have dropped rows <= mn
            const uint64_t xs = warp_sort64<false>(x[u], lane);
            const uint64_t lo = best < xs ? best : xs;     // the 32 keys this merge drops
            drop = drop > lo ? drop : lo;
            best = warp_bitonic_merge_desc(best > xs ? best : xs, lane);
        }
    }
    bound = bound > drop ? bound : drop;
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor64(bound, j);
        bound = bound > o ? bound : o;
    }
    s_pool[warp][lane] = best;
    if (lane == 0) s_bound[warp] = bound;
    __syncthreads();
    // 2. merge the four pools (each sorted descending; read reversed = ascending)
    if (warp == 0) {
        drop = 0ull;
#pragma unroll
        for (int w = 1; w < FIN_WARPS; ++w) {
            const uint64_t xs = s_pool[w][31 - lane];
            const uint64_t lo = best < xs ? best : xs;
            drop = drop > lo ? drop : lo;
            best = warp_bitonic_merge_desc(best > xs ? best : xs, lane);
            if (s_bound[w] > bound) bound = s_bound[w];
        }
        if (pool_n < 32) {                                 // entries beyond pool_n are dropped too
            const uint64_t cut = shfl64(best, pool_n);
            drop = drop > cut ? drop : cut;
        }
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) {
            const uint64_t o = shfl_xor64(drop, j);
            drop = drop > o ? drop : o;
        }
        bound = bound > drop ? bound : drop;
    }
    // pool_n = how many survivors get re-scored.
    if (lane >= pool_n) best = 0ull;
    const size_t row_bytes = (size_t)ld * 2 + 16;           // +16: rows start on different bank groups
    if (stage_rows) {
        // 2b. all four warps copy the 32 surviving rows (and the query) into shared memory with
        //     16-byte async copies, so the sequential fp64 chains below never wait on HBM.
        __syncthreads();
        if (warp == 0) s_pool[0][lane] = best;
        __syncthreads();
        const int chunks = ld / 8;                          // 16-byte pieces per row
        for (int r = warp; r <= pool_n; r += FIN_WARPS) {
            const __nv_bfloat16* src;
            if (r == pool_n) src = qb + q * ld;
            else {
                const uint64_t key = s_pool[0][r];
                if (key == 0ull) continue;
                int64_t slot = head + key_age(key);
                if (slot >= capacity) slot -= capacity;
                src = vec + slot * ld;
            }
            for (int c = lane; c < chunks; c += 32) cp_async16(s_rows + r * row_bytes + 16 * c, src + 8 * c);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        // 3a. exact re-score, split over the four warps: lane = candidate, warp w owns partial sums
        //     2w and 2w+1 of the stated rule (elements d = 2w, 2w+1 mod 8, i.e. word w of every
        //     16-byte chunk), so each warp walks a quarter of the elements.  The pair sum
        //     p[2w] + p[2w+1] is formed here; warp 0 finishes the fixed tree (s0+s1)+(s2+s3).
        {
            const unsigned char* qrow_s = s_rows + pool_n * row_bytes;
            const unsigned char* mrow_s = s_rows + lane * row_bytes;
            double d0 = 0.0, d1 = 0.0, a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
            if (s_pool[0][lane] != 0ull) {
                for (int c = 0; c < ld; c += 8) {
                    const uint32_t qw = *reinterpret_cast<const uint32_t*>(qrow_s + 2 * c + 4 * warp);
                    const uint32_t mw = *reinterpret_cast<const uint32_t*>(mrow_s + 2 * c + 4 * warp);
                    const double q0 = (double)__uint_as_float(qw << 16), q1 = (double)__uint_as_float(qw & 0xffff0000u);
                    const double m0 = (double)__uint_as_float(mw << 16), m1 = (double)__uint_as_float(mw & 0xffff0000u);
                    d0 = __fma_rn(q0, m0, d0); a0 = __fma_rn(q0, q0, a0); b0 = __fma_rn(m0, m0, b0);
                    d1 = __fma_rn(q1, m1, d1); a1 = __fma_rn(q1, q1, a1); b1 = __fma_rn(m1, m1, b1);
                }
            }
            s_part[0][warp][lane] = __dadd_rn(d0, d1);
            s_part[1][warp][lane] = __dadd_rn(a0, a1);
            s_part[2][warp][lane] = __dadd_rn(b0, b1);
        }
        __syncthreads();
    }
    if (warp != 0) return;
    if (gthr) {                                            // rows dropped by a threshold scored below it
        const uint32_t g = __ldcg(gthr + q);
        const uint64_t gk = ((uint64_t)g << 32) | 0xFFFFFFFFull;
        if (g != 0u && gk > bound) bound = gk;
    }
    const bool pool_full = bound != 0ull;                  // something was (or may have been) dropped
    // 3. exact re-score
    const bool valid = best != 0ull;
    double sc = -INFINITY;
    int64_t seq = INT64_MAX;
    const __nv_bfloat16* qrow = qb + q * ld;
    double nqn = 0.0;                                      // lane 0 is valid whenever pool_full
    if (valid) {
        const uint32_t age = key_age(best);
        int64_t slot = head + age;
        if (slot >= capacity) slot -= capacity;
        if (stage_rows) {
            const double dt = __dadd_rn(__dadd_rn(s_part[0][0][lane], s_part[0][1][lane]), __dadd_rn(s_part[0][2][lane], s_part[0][3][lane]));
            nqn = __dadd_rn(__dadd_rn(s_part[1][0][lane], s_part[1][1][lane]), __dadd_rn(s_part[1][2][lane], s_part[1][3][lane]));
            const double nmt = __dadd_rn(__dadd_rn(s_part[2][0][lane], s_part[2][1][lane]), __dadd_rn(s_part[2][2][lane], s_part[2][3][lane]));
            const double den = __dmul_rn(__dsqrt_rn(nqn), __dsqrt_rn(nmt));
            sc = den > 0.0 ? __ddiv_rn(dt, den) : 0.0;
        } else {
            sc = exact_cosine(qrow, vec + slot * ld, ld, &nqn);
        }
        seq = base_seq + age;
    }
    // 4. final order
    warp_sort_pairs(sc, seq, lane);
    if (lane < k) {
        const bool ok = seq != INT64_MAX;
        out_idx[q * k + lane] = ok ? seq * seq_mul + seq_add : -1;
        out_score[q * k + lane] = ok ? (float)sc : -INFINITY;
        if (out_score64) out_score64[q * k + lane] = ok ? sc : -INFINITY;
    }
    // 4b. row-sharded store: the result row goes straight into every rank's exchange buffer over
    // NVLink ([slot][rank] block = [2][nq][k] words: indices, fp64 score bits); the CTA that
    // finishes last releases the step's flag on every rank (fence / counter / fence: the pattern
    // of a last-block reduction, at system scope).
    if (peers) {
        const int slot = (int)(step & (XCHG_SLOTS - 1));
        if (lane < k) {
            const bool ok = seq != INT64_MAX;
            const int64_t iw = ok ? seq * seq_mul + seq_add : -1;
            const int64_t sw = __double_as_longlong(ok ? sc : -INFINITY);
            for (int p = 0; p < world; ++p) {
                int64_t* blk = reinterpret_cast<int64_t*>(peers[p] + ((int64_t)slot * world + rank) * block_bytes);
                blk[q * k + lane] = iw;
                blk[nq * k + q * k + lane] = sw;
            }
        }
        __syncwarp();                                              // orders the lanes' stores before lane 0's fence
        if (lane == 0) {
            __threadfence_system();                                // ONE fence per CTA (cumulative), not one per lane
            const uint32_t old = atomicAdd(done_counter, 1u);
            if (old == (uint32_t)(nq - 1)) {
                *done_counter = 0u;                                // ready for the next launch
                __threadfence_system();
                for (int p = 0; p < world; ++p) {
                    uint64_t* flag = reinterpret_cast<uint64_t*>(peers[p] + flags_off) + (int64_t)slot * world + rank;
                    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(PILOTT_XCHG_FLAG(step)) : "memory");
                }
            }
        }
    }
    // 5. gap check: every row dropped by the approximate pass has approximate score <=
    // the worst kept one; it cannot reach the exact k-th score if the gap exceeds err.
    const double e_k = __longlong_as_double((long long)shfl64((uint64_t)__double_as_longlong(sc), k - 1));
    if (lane == 0) {
        bool bad = false;
        if (pool_full) {
            const double a_min = nqn > 0.0 ? (double)key_score(bound) / sqrt(nqn) : 0.0;
            bad = !(a_min + err < e_k);
        }
        if (flags) flags[q] = bad ? 1 : 0;
        if (bad && inexact) atomicAdd(inexact, 1);
        if (bad && flagged_total) atomicAdd(flagged_total, 1ull);
    }
}

// ---------------------------------------------------------------------------------------
// finish_sel_kernel -- the default finish step.  Same inputs, same outputs and the same proof as
// finish_kernel above, but it re-scores only the survivors that CAN be in the top-k.
//
// finish_kernel re-scores all 32 survivors of every query with lane = survivor, so its cost is set
// by the 768-element fp64 chain of a full warp (and by the bf16 -> fp64 conversions: F2F.F64.F32
// issues at a quarter of the DFMA rate and every lane converted the same query element).  Here:
//   * the 32 survivors come out of the merge sorted by approximate score a_0 >= a_1 >= ...; with
//     |exact - approximate| <= err (the bound the gap check already relies on) the exact k-th best
//     is >= a_{k-1} - err, so a survivor with a_i < a_{k-1} - 2 err cannot be in the top-k.  Only
//     the R survivors above that line are fetched and re-scored (k = 10 on random data: R = 11..13
//     instead of 32: a third of the gather traffic and of the fp64 work);
//   * nothing is taken on trust: the best survivor that was NOT re-scored joins `bound`, and the
//     gap check (bound + err < exact k-th) proves the cut like it proves every other drop -- a
//     query it cannot prove is flagged and goes to the second-chance pass as before;
//   * work is spread as thread = (survivor, partial sum): 14 survivors x the 8 interleaved partial
//     sums of the stated rule per round of 128 threads, so a round costs ld/8 iterations of
//     {LDS.64 query (pre-converted to fp64 once per CTA), LDS.U16 row, one F2F, two DFMA}; the
//     query norm is accumulated once (survivor 0's threads), not once per survivor;
//   * the list length is a template parameter (no integer divisions, a fraction of the code).
// Scores and indices are bit-identical to finish_kernel's whenever the query is not flagged.
// ---------------------------------------------------------------------------------------
constexpr int SEL_ROWS = 14;              // survivors staged and re-scored per round: 14 rows + the query keep a CTA under 32 KB, so 7 CTAs share an SM and 1024 queries are one wave

template <int LL>
__global__ void __launch_bounds__(FIN_WARPS * 32, 7)
finish_sel_kernel(const __grid_constant__ FinishArgs a) {
    extern __shared__ __align__(16) unsigned char s_dyn[];      // [ld] fp64 query | [ld] bf16 query | [SEL_ROWS][ld*2+16] rows
    __shared__ uint64_t s_pool[FIN_WARPS][32];
    __shared__ uint64_t s_bound[FIN_WARPS];
    __shared__ double s_dot[32], s_nm[32];
    __shared__ double s_nq;
    __shared__ int s_R;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x;
    const int64_t q = blockIdx.x;
    const int ld = a.ld, k = a.k;
    double* q64 = reinterpret_cast<double*>(s_dyn);
    unsigned char* q16 = s_dyn + (size_t)ld * 8;
    unsigned char* rows = q16 + (size_t)ld * 2;
    const size_t row_bytes = (size_t)ld * 2 + 16;               // +16: consecutive rows start 4 banks apart
    const __nv_bfloat16* qb = reinterpret_cast<const __nv_bfloat16*>(a.qb);
    const __nv_bfloat16* vec = reinterpret_cast<const __nv_bfloat16*>(a.vec);
    const int chunks = ld / 8;                                  // 16-byte pieces per row
    pdl_trigger();                                              // next step's prepare kernel may start launching
    pdl_wait();                                                 // the scan's lists are complete and visible

    // 0. the query row is on its way while the lists are merged
    for (int c = tid; c < chunks; c += FIN_WARPS * 32) cp_async16(q16 + 16 * c, qb + q * ld + 8 * c);
    asm volatile("cp.async.commit_group;" ::: "memory");

    // 1. top-32 of this warp's lists by approximate key; `bound` = best key NOT kept (see finish_kernel)
    constexpr int PER = 32 / LL;                                // lists per 32-lane load
    const int slots = a.lists_base + ((q / a.rows_per_group) < a.groups_rem ? a.lists_extra : 0);
    const uint64_t* p = a.partials + q * (int64_t)a.slots * LL;
    const int nloads = (slots + PER - 1) / PER;
    auto load_keys = [&](int i) -> uint64_t {
        const int li = i * PER + lane / LL;
        return (i < nloads && li < slots) ? __ldcg(p + (int64_t)li * LL + (lane % LL)) : 0ull;
    };
    uint64_t best = 0ull, bound = 0ull, drop = 0ull;
    uint64_t cur = load_keys(warp), nxt = load_keys(warp + FIN_WARPS);
#pragma unroll 1
    for (int i = warp; i < nloads; i += FIN_WARPS) {
        const uint64_t nn = load_keys(i + 2 * FIN_WARPS);       // two loads ahead
        uint64_t mn = cur;                                      // per-list minimum (segmented butterfly)
#pragma unroll
        for (int j = LL / 2; j > 0; j >>= 1) {
            const uint64_t o = shfl_xor64(mn, j);
            mn = mn < o ? mn : o;
        }
        if (mn != 0ull && mn > bound) bound = mn;               // list is full: it may have dropped rows <= mn
        const uint64_t xs = warp_sort64<false>(cur, lane);
        const uint64_t lo = best < xs ? best : xs;              // the 32 keys this merge drops
        drop = drop > lo ? drop : lo;
        best = warp_bitonic_merge_desc(best > xs ? best : xs, lane);
        cur = nxt; nxt = nn;
    }
    bound = bound > drop ? bound : drop;
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor64(bound, j);
        bound = bound > o ? bound : o;
    }
    s_pool[warp][lane] = best;
    if (lane == 0) s_bound[warp] = bound;
    asm volatile("cp.async.wait_group 0;" ::: "memory");        // own pieces of the query row have landed
    __syncthreads();

    // 2. warp 0: merge the four pools and draw the line; warps 1-3: query -> fp64, once per CTA
    if (warp == 0) {
        drop = 0ull;
#pragma unroll
        for (int w = 1; w < FIN_WARPS; ++w) {
            const uint64_t xs = s_pool[w][31 - lane];           // sorted descending; reversed = ascending
            const uint64_t lo = best < xs ? best : xs;
            drop = drop > lo ? drop : lo;
            best = warp_bitonic_merge_desc(best > xs ? best : xs, lane);
            if (s_bound[w] > bound) bound = s_bound[w];
        }
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) {
            const uint64_t o = shfl_xor64(drop, j);
            drop = drop > o ? drop : o;
        }
        bound = bound > drop ? bound : drop;
        if (a.gthr) {                                           // rows dropped by a threshold scored below it
            const uint32_t g = __ldcg(a.gthr + q);
            const uint64_t gk = ((uint64_t)g << 32) | 0xFFFFFFFFull;
            if (g != 0u && gk > bound) bound = gk;
        }
        // |q| to fp32 accuracy: it only places the line, the gap check below is what proves the cut
        float n2 = 0.f;
        for (int c = lane; c < chunks; c += 32) {
            const uint4 v = *reinterpret_cast<const uint4*>(q16 + 16 * c);
            const uint32_t w4[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float x0 = __uint_as_float(w4[j] << 16), x1 = __uint_as_float(w4[j] & 0xffff0000u);
                n2 = fmaf(x0, x0, n2);
                n2 = fmaf(x1, x1, n2);
            }
        }
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) n2 += __shfl_xor_sync(FULL, n2, j);
        const int nvalid = __popc(__ballot_sync(FULL, best != 0ull));
        int R = 0;
        if (nvalid > 0) {
            const int kk = (k < nvalid ? k : nvalid) - 1;
            const float a_k = key_score(shfl64(best, kk));
            const float line = a_k - 2.0f * (float)a.err * sqrtf(n2) * 1.001f;
            // keys are sorted by score, so the survivors above the line are lanes [0, R); never fewer than k
            R = __popc(__ballot_sync(FULL, best != 0ull && !(key_score(best) < line)));
            if (R <= kk) R = kk + 1;
        }
        if (R < nvalid) {                                       // best survivor left out: it bounds the others
            const uint64_t nb = shfl64(best, R);
            if (nb > bound) bound = nb;
        }
        s_pool[0][lane] = lane < R ? best : 0ull;
        if (lane == 0) { s_R = R; s_bound[0] = bound; s_nq = 0.0; }
    } else {
        for (int d = tid - 32; d < ld; d += (FIN_WARPS - 1) * 32)
            q64[d] = (double)bf16_bits_to_float(*reinterpret_cast<const uint16_t*>(q16 + 2 * d));
    }
    __syncthreads();
    const int R = s_R;

    // 3. rounds of SEL_ROWS survivors: gather the rows, then thread (c, j) walks partial sum j of survivor c
    const int c = tid >> 3, j = tid & 7;
    for (int r0 = 0; r0 < R; r0 += SEL_ROWS) {
        const int nr = R - r0 < SEL_ROWS ? R - r0 : SEL_ROWS;
        for (int r = warp; r < nr; r += FIN_WARPS) {
            int64_t slot = a.head + key_age(s_pool[0][r0 + r]);
            if (slot >= a.capacity) slot -= a.capacity;
            const __nv_bfloat16* src = vec + slot * ld;
            for (int x = lane; x < chunks; x += 32) cp_async16(rows + r * row_bytes + 16 * x, src + 8 * x);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        __syncthreads();
        double pd = 0.0, pm = 0.0, pq = 0.0;
        if (c < nr) {
            const unsigned char* mrow = rows + c * row_bytes;
            if (r0 == 0 && warp == 0) {                         // survivor 0's threads also own |q|^2 (warp-uniform branch)
#pragma unroll 4
                for (int d = j; d < ld; d += 8) {
                    const double qv = q64[d];
                    const double mv = (double)bf16_bits_to_float(*reinterpret_cast<const uint16_t*>(mrow + 2 * d));
                    pd = __fma_rn(qv, mv, pd); pm = __fma_rn(mv, mv, pm); pq = __fma_rn(qv, qv, pq);
                }
            } else {
#pragma unroll 4
                for (int d = j; d < ld; d += 8) {
                    const double qv = q64[d];
                    const double mv = (double)bf16_bits_to_float(*reinterpret_cast<const uint16_t*>(mrow + 2 * d));
                    pd = __fma_rn(qv, mv, pd); pm = __fma_rn(mv, mv, pm);
                }
            }
        }
        // ((p0+p1)+(p2+p3))+((p4+p5)+(p6+p7)): fp64 addition is commutative, so three butterfly
        // steps leave exactly that sum in all eight lanes of the survivor
#pragma unroll
        for (int s = 1; s < 8; s <<= 1) {
            pd = __dadd_rn(pd, shfl_xor_d(pd, s));
            pm = __dadd_rn(pm, shfl_xor_d(pm, s));
            pq = __dadd_rn(pq, shfl_xor_d(pq, s));
        }
        if (c < nr && j == 0) {
            s_dot[r0 + c] = pd; s_nm[r0 + c] = pm;
            if (r0 == 0 && c == 0) s_nq = pq;
        }
        __syncthreads();                                        // results visible; rows free for the next round
    }
    if (warp != 0) return;

    // 4. scores, final order, outputs, gap check (lane = survivor)
    const uint64_t key = s_pool[0][lane];
    bound = s_bound[0];
    const bool pool_full = bound != 0ull;
    const double nqn = s_nq;
    double sc = -INFINITY;
    int64_t seq = INT64_MAX;
    if (lane < R) {
        const double den = __dmul_rn(__dsqrt_rn(nqn), __dsqrt_rn(s_nm[lane]));
        sc = den > 0.0 ? __ddiv_rn(s_dot[lane], den) : 0.0;
        seq = a.base_seq + key_age(key);
    }
    warp_sort_pairs(sc, seq, lane);
    if (lane < k) {
        const bool ok = seq != INT64_MAX;
        a.out_idx[q * k + lane] = ok ? seq * a.seq_mul + a.seq_add : -1;
        a.out_score[q * k + lane] = ok ? (float)sc : -INFINITY;
        if (a.out_score64) a.out_score64[q * k + lane] = ok ? sc : -INFINITY;
    }
    if (a.peers) {                                              // row-sharded store: see finish_kernel step 4b
        const int slot = (int)(a.step & (XCHG_SLOTS - 1));
        if (lane < k) {
            const bool ok = seq != INT64_MAX;
            const int64_t iw = ok ? seq * a.seq_mul + a.seq_add : -1;
            const int64_t sw = __double_as_longlong(ok ? sc : -INFINITY);
            for (int pr = 0; pr < a.world; ++pr) {
                int64_t* blk = reinterpret_cast<int64_t*>(a.peers[pr] + ((int64_t)slot * a.world + a.rank) * a.block_bytes);
                blk[q * k + lane] = iw;
                blk[a.nq * k + q * k + lane] = sw;
            }
        }
        __syncwarp();
        if (lane == 0) {
            __threadfence_system();
            const uint32_t old = atomicAdd(a.done_counter, 1u);
            if (old == (uint32_t)(a.nq - 1)) {
                *a.done_counter = 0u;
                __threadfence_system();
                for (int pr = 0; pr < a.world; ++pr) {
                    uint64_t* flag = reinterpret_cast<uint64_t*>(a.peers[pr] + a.flags_off) + (int64_t)slot * a.world + a.rank;
                    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(PILOTT_XCHG_FLAG(a.step)) : "memory");
                }
            }
        }
    }
    const double e_k = __longlong_as_double((long long)shfl64((uint64_t)__double_as_longlong(sc), k - 1));
    if (lane == 0) {
        bool bad = false;
        if (pool_full) {
            const double a_min = nqn > 0.0 ? (double)key_score(bound) / sqrt(nqn) : 0.0;
            bad = !(a_min + a.err < e_k);
        }
        if (a.flags) a.flags[q] = bad ? 1 : 0;
        if (bad && a.inexact) atomicAdd(a.inexact, 1);
        if (bad && a.flagged_total) atomicAdd(a.flagged_total, 1ull);
    }
}

static size_t finish_sel_smem(int ld) { return (size_t)ld * 10 + SEL_ROWS * ((size_t)ld * 2 + 16); }

template <int LL>
static cudaError_t launch_finish_sel(const FinishArgs& a, cudaStream_t st) {
    const size_t smem = finish_sel_smem(a.ld);
    if (smem > 48 * 1024) {                                     // dims above 1024 only (the attribute is per device)
        cudaError_t e = cudaFuncSetAttribute(finish_sel_kernel<LL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    return launch_pdl(finish_sel_kernel<LL>, dim3((unsigned)a.nq), dim3(FIN_WARPS * 32), smem, st, a);
}

cudaError_t launch_finish(const FinishArgs& a, cudaStream_t st) {
    if (a.nq <= 0) return cudaSuccess;
    // default: selective re-score; rows too long for its shared-memory staging (dim > 4096), or an
    // explicit request (option 9), take the original kernel
    if (!a.legacy && finish_sel_smem(a.ld) <= 200 * 1024) {
        if (a.list_len == 8) return launch_finish_sel<8>(a, st);
        if (a.list_len == 16) return launch_finish_sel<16>(a, st);
        return launch_finish_sel<32>(a, st);
    }
    // stage the surviving rows in shared memory when they fit (pool_n + 1 rows of ld*2+16 bytes)
    const size_t smem = (size_t)(a.pool_n + 1) * ((size_t)a.ld * 2 + 16);
    const int stage_rows = smem <= 100 * 1024 ? 1 : 0;
    if (stage_rows) {
        cudaError_t e = cudaFuncSetAttribute(finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    finish_kernel<<<(unsigned)a.nq, FIN_WARPS * 32, stage_rows ? smem : 0, st>>>(
        a.partials, a.slots, a.list_len, a.pool_n, a.rows_per_group, a.lists_base, a.groups_rem, a.lists_extra,
        (const __nv_bfloat16*)a.qb, (const __nv_bfloat16*)a.vec, a.ld, a.nq,
        a.capacity, a.head, a.base_seq, a.seq_mul, a.seq_add, a.k, a.err, a.out_idx, a.out_score, a.out_score64,
        a.flags, a.inexact, stage_rows, a.gthr, a.peers, a.world, a.rank, a.block_bytes, a.flags_off, a.step,
        a.done_counter, a.flagged_total);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Brute-force exact search.  Pass 1: fp64 score of every candidate row (NaN = filtered).
// Pass 2: k rounds of "best key strictly after the previous one", one CTA per query.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
exact_scores_kernel(const __nv_bfloat16* __restrict__ qb, const __nv_bfloat16* __restrict__ vec, int ld,
                    const int32_t* __restrict__ prio, const uint64_t* __restrict__ tagmask,
                    const int64_t* __restrict__ expires, int64_t nslots, const int32_t* __restrict__ qids,
                    int32_t min_prio, uint64_t req, int64_t now_us, double* __restrict__ scratch) {
    const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= nslots) return;
    const int sel = blockIdx.y;
    const int64_t q = qids ? qids[sel] : sel;
    const int64_t e = expires[n];
    const bool ok = prio[n] >= min_prio && (tagmask[n] & req) == req && !(e != 0 && now_us > e);
    double s = __longlong_as_double(0x7ff8000000000000ll);
    if (ok) s = exact_cosine(qb + q * ld, vec + n * ld, ld);
    scratch[(int64_t)sel * nslots + n] = s;
}

__global__ void __launch_bounds__(1024)
exact_select_kernel(const double* __restrict__ scratch, int64_t nslots, int64_t capacity, int64_t head,
                    int64_t base_seq, int64_t seq_mul, int64_t seq_add, const int32_t* __restrict__ qids, int k,
                    int64_t* __restrict__ out_idx, float* __restrict__ out_score,
                    double* __restrict__ out_score64) {
    __shared__ double s_sc[32];
    __shared__ int64_t s_age[32];
    __shared__ double prev_sc;
    __shared__ int64_t prev_age;
    const int sel = blockIdx.x;
    const int64_t q = qids ? qids[sel] : sel;
    const double* sc = scratch + (int64_t)sel * nslots;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { prev_sc = INFINITY; prev_age = -1; }
    __syncthreads();
    for (int r = 0; r < k; ++r) {
        const double ps = prev_sc;
        const int64_t pa = prev_age;
        double bs = -INFINITY;
        int64_t ba = INT64_MAX;               // INT64_MAX == nothing found
        for (int64_t n = threadIdx.x; n < nslots; n += blockDim.x) {
            const double s = sc[n];
            if (s != s) continue;             // filtered
            int64_t age = n - head;
            if (age < 0) age += capacity;
            const bool after_prev = ranks_before(ps, pa, s, age);
            if (after_prev && (ba == INT64_MAX || ranks_before(s, age, bs, ba))) { bs = s; ba = age; }
        }
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) {
            const double os = shfl_xor_d(bs, j);
            const int64_t oa = (int64_t)shfl_xor64((uint64_t)ba, j);
            if (oa != INT64_MAX && (ba == INT64_MAX || ranks_before(os, oa, bs, ba))) { bs = os; ba = oa; }
        }
        if (lane == 0) { s_sc[warp] = bs; s_age[warp] = ba; }
        __syncthreads();
        if (warp == 0) {
            bs = s_sc[lane]; ba = s_age[lane];
#pragma unroll
            for (int j = 16; j > 0; j >>= 1) {
                const double os = shfl_xor_d(bs, j);
                const int64_t oa = (int64_t)shfl_xor64((uint64_t)ba, j);
                if (oa != INT64_MAX && (ba == INT64_MAX || ranks_before(os, oa, bs, ba))) { bs = os; ba = oa; }
            }
            if (lane == 0) {
                const bool ok = ba != INT64_MAX;
                out_idx[q * k + r] = ok ? (base_seq + ba) * seq_mul + seq_add : -1;
                out_score[q * k + r] = ok ? (float)bs : -INFINITY;
                if (out_score64) out_score64[q * k + r] = ok ? bs : -INFINITY;
                // once nothing is found, nothing will be found again: park the cursor at the end
                prev_sc = ok ? bs : -INFINITY;
                prev_age = ok ? ba : INT64_MAX;
            }
        }
        __syncthreads();
    }
}

cudaError_t launch_exact(const void* qb, const void* vec, int ld, const int32_t* prio,
                         const uint64_t* tagmask, const int64_t* expires, int64_t nslots,
                         int64_t capacity, int64_t head, int64_t base_seq, int64_t seq_mul, int64_t seq_add,
                         const int32_t* qids, int64_t nsel, int k, int32_t min_prio, uint64_t req, int64_t now_us,
                         double* scratch, int64_t* out_idx, float* out_score, double* out_score64,
                         cudaStream_t st) {
    if (nsel <= 0) return cudaSuccess;
    if (nslots > 0) {
        dim3 grid((unsigned)((nslots + 127) / 128), (unsigned)nsel);
        exact_scores_kernel<<<grid, 128, 0, st>>>((const __nv_bfloat16*)qb, (const __nv_bfloat16*)vec,
                                                  ld, prio, tagmask, expires, nslots, qids,
                                                  min_prio, req, now_us, scratch);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
    }
    exact_select_kernel<<<(unsigned)nsel, 1024, 0, st>>>(scratch, nslots, capacity, head, base_seq,
                                                         seq_mul, seq_add, qids, k, out_idx, out_score, out_score64);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Second chance for queries the gap check flagged (near-ties wider than a candidate list, bursts
// of near-duplicate rows).  The fused search already holds k rows with EXACT scores for such a
// query; the k-th of them, LB, is a lower bound of the true k-th best.  A row can only belong to
// the true top-k if its exact score is >= LB, hence (same error bound as the gap check) if its
// approximate score s~ satisfies s~/|q| + err >= LB.  refine_prepare turns that into one fp32
// threshold per query and packs the flagged queries into a dense query block; the scan kernel in
// COLLECT mode appends every row reaching the threshold; refine_finish re-scores those rows by
// the stated rule and ranks them.  A query with more than `cap` such rows is left to the
// brute-force path (overflow flag).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
refine_prepare_kernel(const __nv_bfloat16* __restrict__ qb, const int32_t* __restrict__ qids, int64_t nsel,
                      int64_t nsel_pad, int ld, const int64_t* __restrict__ prior_idx,
                      const double* __restrict__ prior_s64, int k, double err,
                      __nv_bfloat16* __restrict__ qb2, uint32_t* __restrict__ thr_bits,
                      uint32_t* __restrict__ cnt, int32_t* __restrict__ n_overflow) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i == 0 && lane == 0) *n_overflow = 0;
    if (i >= nsel_pad) return;
    uint4* dst = reinterpret_cast<uint4*>(qb2 + i * ld);
    if (i >= nsel) {                                   // padding row of the dense block
        for (int c = lane; c < ld / 8; c += 32) dst[c] = make_uint4(0, 0, 0, 0);
        if (lane == 0) { thr_bits[i] = __float_as_uint(INFINITY); cnt[i] = 0u; }
        return;
    }
    const int64_t q = qids[i];
    const uint4* src = reinterpret_cast<const uint4*>(qb + q * ld);
    double n2 = 0.0;                                   // |q|^2 (threshold only: any accurate sum will do)
    for (int c = lane; c < ld / 8; c += 32) {
        const uint4 v = src[c];
        dst[c] = v;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double a = (double)__uint_as_float(w[j] << 16), b = (double)__uint_as_float(w[j] & 0xffff0000u);
            n2 += a * a + b * b;
        }
    }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) n2 += shfl_xor_d(n2, j);
    if (lane == 0) {
        const bool have = prior_idx[q * k + (k - 1)] >= 0;          // fused search found k rows
        float t = -INFINITY;
        if (have) {
            // (LB - err) * |q|, pushed down past every rounding on the way (1e-6 relative + one ulp)
            double td = (prior_s64[q * k + (k - 1)] - err) * sqrt(n2);
            td -= fabs(td) * 1e-6;
            t = __double2float_rd(td);
            if (t == t && fabsf(t) < INFINITY)         // one fp32 step further down
                t = __uint_as_float(t > 0.f ? __float_as_uint(t) - 1u : (t < 0.f ? __float_as_uint(t) + 1u : 0x80000001u));
            else if (t != t) t = -INFINITY;
        }
        thr_bits[i] = __float_as_uint(t);
        cnt[i] = 0u;
    }
}

constexpr int REFINE_THREADS = 256;

__global__ void __launch_bounds__(REFINE_THREADS)
refine_finish_kernel(const uint32_t* __restrict__ cand, const uint32_t* __restrict__ cnt, int cap,
                     const __nv_bfloat16* __restrict__ qb2, const __nv_bfloat16* __restrict__ vec, int ld,
                     const int32_t* __restrict__ qids, int64_t capacity, int64_t head, int64_t base_seq,
                     int64_t seq_mul, int64_t seq_add, int k, int64_t* __restrict__ out_idx,
                     float* __restrict__ out_score, double* __restrict__ out_score64,
                     uint8_t* __restrict__ overflow, int32_t* __restrict__ n_overflow,
                     const double* __restrict__ tcos, double err, double* __restrict__ hist_t,
                     uint32_t* __restrict__ hist_c) {
    // status written to overflow[sel]: 0 settled, 1 more candidates than `cap`, 2 (only with tcos: the
    // large-k pass, whose thresholds are guesses) the k-th exact score is not provably above
    // every uncollected row -- the threshold was too high, try a lower one.
    extern __shared__ __align__(16) unsigned char sm_raw[];
    __shared__ double s_ek;
    double* s_sc = reinterpret_cast<double*>(sm_raw);                  // [cap]
    int64_t* s_age = reinterpret_cast<int64_t*>(s_sc + cap);           // [cap]
    const int64_t sel = blockIdx.x;
    const int64_t q = qids[sel];
    const uint32_t c = cnt[sel];
    if (hist_t && threadIdx.x == 0) { hist_t[q] = tcos[sel]; hist_c[q] = c; }      // (threshold, rows above it): one measured point of the tail
    if (c > (uint32_t)cap) {                           // too many near-ties: brute force decides
        if (threadIdx.x == 0) { overflow[sel] = 1; atomicAdd(n_overflow, 1); }
        return;
    }
    if (threadIdx.x == 0) { overflow[sel] = 0; s_ek = -INFINITY; }
    const __nv_bfloat16* qrow = qb2 + sel * ld;
    for (uint32_t i = threadIdx.x; i < c; i += REFINE_THREADS) {
        const int64_t slot = cand[sel * cap + i];
        s_sc[i] = exact_cosine(qrow, vec + slot * ld, ld);
        int64_t age = slot - head;
        if (age < 0) age += capacity;
        s_age[i] = age;
    }
    for (int j = threadIdx.x; j < k; j += REFINE_THREADS) {
        out_idx[q * k + j] = -1;
        out_score[q * k + j] = -INFINITY;
        if (out_score64) out_score64[q * k + j] = -INFINITY;
    }
    __syncthreads();
    // rank counting: ages are unique, so (score, age) pairs are and every rank is taken once
    for (uint32_t i = threadIdx.x; i < c; i += REFINE_THREADS) {
        const double ms = s_sc[i];
        const int64_t ma = s_age[i];
        int rank = 0;
        for (uint32_t j = 0; j < c; ++j) rank += ranks_before(s_sc[j], s_age[j], ms, ma) ? 1 : 0;
        if (rank < k) {
            out_idx[q * k + rank] = (base_seq + ma) * seq_mul + seq_add;
            out_score[q * k + rank] = (float)ms;
            if (out_score64) out_score64[q * k + rank] = ms;
        }
        if (rank == k - 1) s_ek = ms;
    }
    if (tcos) {
        // rows the scan did not collect have approximate cosine < tcos[sel], hence exact < tcos + err
        __syncthreads();
        if (threadIdx.x == 0) {
            const double t = tcos[sel];
            const bool all_rows = t == -INFINITY;      // open threshold: every candidate row was collected
            if (!all_rows && !(c >= (uint32_t)k && s_ek >= t + err)) { overflow[sel] = 2; atomicAdd(n_overflow, 1); }
        }
    }
}

// Large k (32 < k <= 1024): no candidate lists that long exist, so the threshold of the COLLECT
// pass is a GUESS -- the fused top-32 of the query gives s_4 and s_32; with the scores of the
// n_rows held rows taken as normal, those two order statistics fix mean and spread, and the guess
// is the score expected at rank factor * k -- and refine_finish_kernel CHECKS it (status 2 = too
// high).  factor <= 0, a target rank beyond the store, or a tiny store open the threshold.
__global__ void __launch_bounds__(128)
largek_prepare_kernel(const __nv_bfloat16* __restrict__ qb, const int32_t* __restrict__ qids, int64_t nsel,
                      int64_t nsel_pad, int ld, const int64_t* __restrict__ top_idx,
                      const double* __restrict__ top_s64, int k, double factor, double min_gap, double n_rows,
                      const double* __restrict__ hist_t, const uint32_t* __restrict__ hist_c, double cap_rows,
                      __nv_bfloat16* __restrict__ qb2, uint32_t* __restrict__ thr_bits, double* __restrict__ tcos,
                      uint32_t* __restrict__ cnt, int32_t* __restrict__ n_overflow) {
    const int lane = threadIdx.x & 31;
    const int64_t i = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (i == 0 && lane == 0) *n_overflow = 0;
    if (i >= nsel_pad) return;
    uint4* dst = reinterpret_cast<uint4*>(qb2 + i * ld);
    if (i >= nsel) {
        for (int c = lane; c < ld / 8; c += 32) dst[c] = make_uint4(0, 0, 0, 0);
        if (lane == 0) { thr_bits[i] = __float_as_uint(INFINITY); cnt[i] = 0u; }
        return;
    }
    const int64_t q = qids[i];
    const uint4* src = reinterpret_cast<const uint4*>(qb + q * ld);
    double n2 = 0.0;
    for (int c = lane; c < ld / 8; c += 32) {
        const uint4 v = src[c];
        dst[c] = v;
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const double a = (double)__uint_as_float(w[j] << 16), b = (double)__uint_as_float(w[j] & 0xffff0000u);
            n2 += a * a + b * b;
        }
    }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) n2 += shfl_xor_d(n2, j);
    if (lane == 0) {
        float t = -INFINITY;
        double tc = -INFINITY;
        const double nq = sqrt(n2);
        const double target = factor * (double)k;
        if (factor > 0.0 && nq > 0.0 && n_rows >= 256.0 && target < 0.5 * n_rows && top_idx[q * 32 + 31] >= 0) {
            const double s4 = top_s64[q * 32 + 3], s32 = top_s64[q * 32 + 31];
            const double z4 = -normcdfinv(4.0 / n_rows), z32 = -normcdfinv(32.0 / n_rows), zt = -normcdfinv(target / n_rows);
            const double sigma = fmax((s4 - s32) / (z4 - z32), 0.0);
            double gap = sigma * (z32 - zt);
            // A previous pass measured how many rows (hc) lie above ITS threshold (ht): a second
            // point further down the tail, much better than s_4 for the spread (secant step in
            // z-space).  Whatever the fit says, a threshold that proved too high is followed by a
            // lower one and a threshold that collected more than the pass can hold by a higher one.
            const double hc = (double)hist_c[q], ht = hist_t[q];
            if (hc > 0.0 && ht > -INFINITY) {
                const double gap_prev = s32 - ht;
                if (hc > 36.0 && hc < 0.5 * n_rows && gap_prev > 0.0) {
                    const double zh = -normcdfinv(hc / n_rows);
                    if (z32 - zh > 1e-6) gap = gap_prev / (z32 - zh) * (z32 - zt);
                } else if (gap_prev > 0.0) {
                    gap = 2.0 * gap_prev;
                }
                if (hc <= cap_rows) gap = fmax(gap, gap_prev + min_gap);
                else if (gap_prev - min_gap > 0.0) gap = fmin(gap, gap_prev - min_gap);
            }
            gap = fmax(gap, min_gap);
            t = __double2float_rd((s32 - gap) * nq);
            if (t == t && fabsf(t) < INFINITY) tc = (double)t / nq;               // what the scan really applies
            else t = -INFINITY;
        }
        thr_bits[i] = __float_as_uint(t);
        tcos[i] = tc;
        cnt[i] = 0u;
    }
}

cudaError_t launch_refine_prepare(const void* qb, const int32_t* qids, int64_t nsel, int64_t nsel_pad, int ld,
                                  const int64_t* prior_idx, const double* prior_s64, int k, double err, void* qb2,
                                  uint32_t* thr_bits, uint32_t* cnt, int32_t* n_overflow, cudaStream_t st) {
    const unsigned grid = (unsigned)((nsel_pad + 3) / 4);
    refine_prepare_kernel<<<grid, 128, 0, st>>>((const __nv_bfloat16*)qb, qids, nsel, nsel_pad, ld, prior_idx, prior_s64,
                                                k, err, (__nv_bfloat16*)qb2, thr_bits, cnt, n_overflow);
    return cudaGetLastError();
}

cudaError_t launch_refine_finish(const uint32_t* cand, const uint32_t* cnt, int cap, const void* qb2, const void* vec,
                                 int ld, const int32_t* qids, int64_t nsel, int64_t capacity, int64_t head,
                                 int64_t base_seq, int64_t seq_mul, int64_t seq_add, int k, int64_t* out_idx,
                                 float* out_score, double* out_score64, uint8_t* overflow, int32_t* n_overflow,
                                 const double* tcos, double err, double* hist_t, uint32_t* hist_c, cudaStream_t st) {
    if (nsel <= 0) return cudaSuccess;
    const size_t smem = (size_t)cap * 16;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(refine_finish_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    refine_finish_kernel<<<(unsigned)nsel, REFINE_THREADS, smem, st>>>(
        cand, cnt, cap, (const __nv_bfloat16*)qb2, (const __nv_bfloat16*)vec, ld, qids, capacity, head, base_seq,
        seq_mul, seq_add, k, out_idx, out_score, out_score64, overflow, n_overflow, tcos, err, hist_t, hist_c);
    return cudaGetLastError();
}

cudaError_t launch_largek_prepare(const void* qb, const int32_t* qids, int64_t nsel, int64_t nsel_pad, int ld,
                                  const int64_t* top_idx, const double* top_s64, int k, double factor, double min_gap,
                                  double n_rows, const double* hist_t, const uint32_t* hist_c, double cap_rows, void* qb2,
                                  uint32_t* thr_bits, double* tcos, uint32_t* cnt, int32_t* n_overflow, cudaStream_t st) {
    const unsigned grid = (unsigned)((nsel_pad + 3) / 4);
    largek_prepare_kernel<<<grid, 128, 0, st>>>((const __nv_bfloat16*)qb, qids, nsel, nsel_pad, ld, top_idx, top_s64, k,
                                                factor, min_gap, n_rows, hist_t, hist_c, cap_rows, (__nv_bfloat16*)qb2, thr_bits, tcos, cnt, n_overflow);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Multi-GPU merge: [shards, nq, k] exact (score, idx) lists -> global top-k per query by
// rank counting (shards*k <= 256 candidates; pairs are unique because idx is).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
topk_merge_kernel(const int64_t* __restrict__ idx, const double* __restrict__ sc, int shards, int64_t nq,
                  int k, int64_t shard_stride, int64_t* __restrict__ out_idx, float* __restrict__ out_score,
                  double* __restrict__ out_score64) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int total = shards * k;
    double* s_sc = reinterpret_cast<double*>(sm_raw);
    int64_t* s_idx = reinterpret_cast<int64_t*>(s_sc + total);
    const int64_t q = blockIdx.x;
    pdl_trigger();
    pdl_wait();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int sh = i / k, j = i % k;
        s_sc[i] = sc[(int64_t)sh * shard_stride + q * k + j];
        s_idx[i] = idx[(int64_t)sh * shard_stride + q * k + j];
    }
    for (int j = threadIdx.x; j < k; j += blockDim.x) {
        out_idx[q * k + j] = -1;
        out_score[q * k + j] = -INFINITY;
        if (out_score64) out_score64[q * k + j] = -INFINITY;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int64_t me = s_idx[i];
        if (me < 0) continue;
        const double ms = s_sc[i];
        int rank = 0;
        for (int j = 0; j < total; ++j)
            if (s_idx[j] >= 0 && ranks_before(s_sc[j], s_idx[j], ms, me)) ++rank;
        if (rank < k) {
            out_idx[q * k + rank] = me;
            out_score[q * k + rank] = (float)ms;
            if (out_score64) out_score64[q * k + rank] = ms;
        }
    }
}

cudaError_t launch_topk_merge(const int64_t* idx, const double* sc, int shards, int64_t nq, int k,
                              int64_t shard_stride, int64_t* out_idx, float* out_score,
                              double* out_score64, cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    const size_t smem = (size_t)shards * k * 16;
    return launch_pdl(topk_merge_kernel, dim3((unsigned)nq), dim3(256), smem, st, idx, sc, shards, nq, k, shard_stride,
                      out_idx, out_score, out_score64);
}

}  // namespace pilott
