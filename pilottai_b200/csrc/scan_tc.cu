// scan_tc.cu -- K2+K3: query x memory similarity on tcgen05 tensor cores with the
// norm/mask/score epilogue and the streaming top-KP select fused in.  sm_100a only.
//
// GEMM view: D[query, row] = sum_d Q[query, d] * M[row, d]   (both operands K-major bf16,
// fp32 accumulation in TMEM).  The Q x N score matrix never reaches HBM: each epilogue
// thread owns one query row of the accumulator tile (TMEM lane == query), multiplies by
// the row's scale (inverse norm, NaN when filtered) and compares against its running
// threshold; the rare survivors are appended to a shared-memory candidate buffer that a
// whole warp prunes back to KP entries when it fills (common.cuh).
//
// Work split: a CTA (CG=1) or CTA pair (CG=2) is bound to ONE query tile for its whole
// life and sweeps a contiguous range of 256-row corpus tiles ("slot"), so the selection
// state stays on chip; the per-slot candidate lists are merged by finish.cu.
//
// Warp roles (384 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator,
// 3 = idle, 4..11 = epilogue in two sets of four warps (TMEM lane quarter = warp % 4).
// Set s drains columns [128*s, 128*s+128) of EVERY accumulator tile with its own candidate
// buffers: the tile's epilogue time halves (an accumulator stage is busy for MMA time +
// epilogue time, so with two stages the epilogue of one tile must fit in one MMA time), and
// every SM sub-partition holds two epilogue warps that cover each other's latencies.
// Pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty x2 (MMA <-> epilogue).
#include <cstdio>

#include "common.cuh"
#include "internal.h"
#include "tc_ptx.cuh"

namespace pilott {

constexpr int BM = 128;        // query rows per CTA (TMEM lanes)
constexpr int BN = 256;        // corpus rows per tile (UMMA N, TMEM columns per stage)
constexpr int BK = 64;         // bf16 elements per k-block = one 128-byte swizzle atom
constexpr int UK = 16;         // UMMA K for 16-bit inputs
constexpr int ACC_STAGES = 2;  // 2 x 256 TMEM columns

constexpr int EPI_SETS = 2;    // one per accumulator stage
constexpr int TC_THREADS = 128 + EPI_SETS * 128;
constexpr int PF_TILES = 0;     // corpus tiles prefetched into L2 ahead of the TMA loads (0: off -- measured no gain)
constexpr int GMAX_LISTS = 32;  // lists per query that publish their first-tile maximum (cold-start threshold)
constexpr int SC_RING = 4;     // staged scale tiles in flight (full/empty mbarrier pair per slot)

// Shared-memory budget per (CTA group, list length).  8-entry lists need a 30-entry row buffer
// (8 survivors + 22 pending) instead of 46, which buys the pair kernel a fifth smem stage: the
// mainloop is bound by TMA supply, and 4 -> 5 stages takes it from 1.12 to 1.04 ms.
template <int CG, int KPT> struct TcCfg {
    static constexpr int STAGES = CG == 1 ? 2 : (KPT == 8 ? 5 : 4);
    static constexpr int CAPT = KPT == 8 ? 30 : 46;                // per-row candidate buffer: survivors + pending appends
    static constexpr int A_BYTES = BM * BK * 2;                    // 16 KB
    static constexpr int B_ROWS = BN / CG;                         // rows of B this CTA loads
    static constexpr int B_BYTES = B_ROWS * BK * 2;                // 32 KB / 16 KB
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BUF_BYTES = EPI_SETS * BM * CAPT * 8;     // 96 KB candidate buffers
    static constexpr int SC_BYTES = SC_RING * BN * 4;              // ring of staged scale tiles (TMA bulk copies)
    static constexpr int BAR_BYTES = 256;
    static constexpr int SMEM = 1024 /*align slack*/ + STAGES * STAGE_BYTES + BUF_BYTES + SC_BYTES + BAR_BYTES;
};

// Sorts 32 values held in registers into descending order (bitonic network, every index a
// compile-time constant) and returns the one of rank `kth` (0 = largest).
__device__ __forceinline__ uint32_t kth_largest_of_32(uint32_t (&v)[32], int kth) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int l = i ^ j;
                if (l > i) {
                    const bool desc = (i & k) == 0;
                    const uint32_t lo = min(v[i], v[l]), hi = max(v[i], v[l]);
                    v[i] = desc ? hi : lo;
                    v[l] = desc ? lo : hi;
                }
            }
        }
    }
    uint32_t r = v[0];
#pragma unroll
    for (int i = 1; i < 32; ++i) r = (i == kth) ? v[i] : r;
    return r;
}

// ---- the kernel ------------------------------------------------------------------------
// Grid: `units` work units (CTAs for CG=1, CTA pairs for CG=2).  Unit u serves query group
// u % groups (BM*CG query rows) and slot u / groups of that group.
// COLLECT = true is the second-chance pass for queries the gap check flagged: no lists, no
// pruning -- every row whose approximate score reaches the query's FIXED threshold
// (thresholds in `gthr` as float bits, one per query row) has its ring slot appended to the
// query's candidate array (`partials` viewed as uint32 [nq][krank], counters in `gmax`).
template <int CG, int KPT, bool COLLECT>
__global__ void __launch_bounds__(TC_THREADS, 1)
scan_tc_kernel(const __grid_constant__ CUtensorMap map_q, const __grid_constant__ CUtensorMap map_c,
               const float* __restrict__ scale, const float* __restrict__ tile_max,
               uint64_t* __restrict__ partials, uint32_t* __restrict__ gthr, uint32_t* __restrict__ gmax,
               uint32_t* __restrict__ progress, int window, int64_t nq, int64_t nslots, int64_t capacity, int64_t head, int kblocks, int slots, int slots_rem, int groups, int krank, int dbg_arg) {
    // Developer switches (profiling experiments, some of which give WRONG results on purpose) exist
    // only in -DPILOTT_DEBUG builds; in the production build `dbg` is the constant 0 and every
    // `dbg &` branch, counter and clock read below is compiled out.
#ifdef PILOTT_DEBUG
    const int dbg = dbg_arg;
#else
    constexpr int dbg = 0;
    (void)dbg_arg;
#endif
    using Cfg = TcCfg<CG, KPT>;
    constexpr int CAPT = Cfg::CAPT;
    extern __shared__ unsigned char smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;        // 1024-B aligned (swizzle atom)
    unsigned char* base_ptr = smem_dyn + (base - smem_u32(smem_dyn));
    const uint32_t stage_base = base;
    uint64_t* bufs = reinterpret_cast<uint64_t*>(base_ptr + Cfg::STAGES * Cfg::STAGE_BYTES);
    const float* ssc = reinterpret_cast<const float*>(base_ptr + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::BUF_BYTES);
    const uint32_t bar_base = base + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::BUF_BYTES + Cfg::SC_BYTES;
    // barrier layout (8 B each): full[STAGES], empty[STAGES], tfull[2], tempty[2], then tmem ptr
    const uint32_t bar_full = bar_base;
    const uint32_t bar_empty = bar_base + 8 * Cfg::STAGES;
    const uint32_t bar_tfull = bar_base + 16 * Cfg::STAGES;
    const uint32_t bar_tempty = bar_tfull + 8 * ACC_STAGES;
    const uint32_t bar_scfull = bar_tempty + 8 * ACC_STAGES;
    const uint32_t bar_scempty = bar_scfull + 8 * SC_RING;
    const uint32_t tmem_slot = bar_scempty + 8 * SC_RING;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(
        base_ptr + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::BUF_BYTES + Cfg::SC_BYTES + 16 * Cfg::STAGES + 16 * ACC_STAGES + 16 * SC_RING);
    const uint32_t ssc_u32 = base + Cfg::STAGES * Cfg::STAGE_BYTES + Cfg::BUF_BYTES;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
    const bool leader = rank == 0;
    const int unit = CG == 2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
    const int group = unit % groups, slot = unit / groups;
    // The first `slots_rem` query groups get one more work unit than the others, so that every
    // SM has a CTA even when the SM count is not a multiple of the group count.
    const int slots_g = slots + (group < slots_rem ? 1 : 0);
    const int max_lists = (slots + (slots_rem > 0 ? 1 : 0)) * EPI_SETS;  // candidate lists per query row
    const int64_t q0 = ((int64_t)group * CG + rank) * BM;               // this CTA's first query row
    const int64_t ntiles = (nslots + BN - 1) / BN;
    const int64_t t_lo = ntiles * slot / slots_g, t_hi = ntiles * (slot + 1) / slots_g;
    // Pair mode with at most 128 real queries in the group (single queries, small or ragged
    // batches): the second CTA's whole query tile is padding.  It zeroes its A stages once and never
    // loads them again -- a quarter less L2 -> SM traffic in the regime where that traffic is the
    // bound -- and its epilogue warps only hand the accumulator back.
    const bool a_pad = CG == 2 && q0 >= nq;                              // this CTA's tile has no query
    const bool pair_pad = CG == 2 && ((int64_t)group * CG + 1) * BM >= nq;   // ... the pair's second tile
    if (a_pad) {
        for (int st = 0; st < Cfg::STAGES; ++st) {
            uint4* a = reinterpret_cast<uint4*>(base_ptr + st * Cfg::STAGE_BYTES);
            for (int i = threadIdx.x; i < Cfg::A_BYTES / 16; i += TC_THREADS) a[i] = make_uint4(0, 0, 0, 0);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");    // generic stores -> visible to UMMA
    }

    pdl_trigger();                                                      // the finish step may start launching
    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_q) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < Cfg::STAGES; ++i) {
            mbar_init(bar_full + 8 * i, 1);
            mbar_init(bar_empty + 8 * i, 1);
        }
        for (int i = 0; i < ACC_STAGES; ++i) {
            mbar_init(bar_tfull + 8 * i, 1);
            mbar_init(bar_tempty + 8 * i, 4 * EPI_SETS * CG);   // one arrive per epilogue warp per CTA
        }
        for (int i = 0; i < SC_RING; ++i) {
            mbar_init(bar_scfull + 8 * i, 1);
            mbar_init(bar_scempty + 8 * i, 4 * EPI_SETS);      // every epilogue warp reads part of the slot
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) tmem_alloc<CG>(tmem_slot, 512);
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;
    // Everything above touched only this CTA's shared memory and TMEM; from here on the roles read
    // what the prepare kernel wrote (queries, scales, zeroed words) and overwrite lists the previous
    // finish step read: wait for the predecessor grid (no-op without programmatic launch).
    pdl_wait();

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            const uint32_t full0 = CG == 2 ? mapa_u32(bar_full, 0) : bar_full;   // leader's barriers
            int stage = 0; uint32_t phase = 0;
            for (int64_t t = t_lo; t < t_hi; ++t) {
                const int n0 = (int)(t * BN) + (int)rank * Cfg::B_ROWS;
                // Keep step with the other query groups that sweep this slot: a corpus tile is read from
                // HBM once and served to the other groups from L2 only while they are within what L2
                // holds of each other; over thousands of tiles they drift (10 M x 1024: the corpus came
                // from DRAM twice per batch).  Before issuing tile i the producer waits until every
                // group of the slot has issued tile i - window.  Only launched this way when the whole
                // grid is resident (every unit it waits for is running); the wait is bounded anyway.
                if (progress != nullptr) {
                    const uint32_t ti = (uint32_t)(t - t_lo);
                    if (leader) *reinterpret_cast<volatile uint32_t*>(progress + unit) = ti;
                    if (ti > (uint32_t)window) {
                        const uint32_t need = ti - (uint32_t)window;
                        for (int spins = 0; spins < (1 << 16); ++spins) {
                            uint32_t lo = 0xffffffffu;
                            for (int g = 0; g < groups; ++g) {
                                const uint32_t p = *reinterpret_cast<volatile uint32_t*>(progress + slot * groups + g);
                                lo = p < lo ? p : lo;
                            }
                            if (lo >= need) break;
                            __nanosleep(200);
                        }
                    }
                }
                for (int kb = 0; kb < kblocks; ++kb) {
                    mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                    if (kb == 0) {
                        // this tile's scale row -> ring slot (t - t_lo) % SC_RING
                        const int64_t ti = t - t_lo;
                        const int ss = (int)(ti % SC_RING);
                        mbar_wait(bar_scempty + 8 * ss, (uint32_t)(((ti / SC_RING) & 1) ^ 1));
                        mbar_arrive_expect_tx(bar_scfull + 8 * ss, BN * 4);
                        bulk_load_1d(ssc_u32 + ss * BN * 4, scale + t * BN, BN * 4, bar_scfull + 8 * ss);
                    }
                    const uint32_t sa = stage_base + stage * Cfg::STAGE_BYTES;
                    const uint32_t sb = sa + Cfg::A_BYTES;
                    // corpus k-block of the tile PF_TILES ahead -> L2, so its TMA load later sees L2 latency
                    if (PF_TILES > 0 && !(dbg & 8) && t + PF_TILES < t_hi) tma_prefetch_l2_2d(&map_c, kb * BK, n0 + PF_TILES * BN);
                    if (CG == 1) {
                        mbar_arrive_expect_tx(bar_full + 8 * stage, Cfg::STAGE_BYTES);
                        tma_load_2d(sa, &map_q, kb * BK, (int)q0, bar_full + 8 * stage);
                        tma_load_2d(sb, &map_c, kb * BK, n0, bar_full + 8 * stage);
                    } else {
                        // dbg & 64: traffic experiment -- skip the corpus (B) load on odd k-blocks (stale
                        // data, wrong results) to see how much of the mainloop time is L2 -> SM supply
                        const bool skip_b = (dbg & 64) && (kb & 1);
                        if (leader) mbar_arrive_expect_tx(bar_full + 8 * stage, (skip_b ? 2 * Cfg::A_BYTES : 2 * Cfg::STAGE_BYTES) -
                                                                                      (pair_pad ? Cfg::A_BYTES : 0));
                        if (!a_pad) tma_load_2d_cg2(sa, &map_q, kb * BK, (int)q0, full0 + 8 * stage);
                        if (!skip_b) tma_load_2d_cg2(sb, &map_c, kb * BK, n0, full0 + 8 * stage);
                    }
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only in pair mode) =====
        if (lane == 0 && leader) {
            constexpr uint32_t idesc = make_idesc(BM * CG, BN);
            int stage = 0; uint32_t phase = 0;
            int as = 0; uint32_t aphase = 0;
            long long w_tempty = 0, w_full = 0, t_start = (dbg & 4) ? clock64() : 0;
            for (int64_t t = t_lo; t < t_hi; ++t) {
                long long c0 = (dbg & 4) ? clock64() : 0;
                mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                if (dbg & 4) {
                    const long long w = clock64() - c0;
                    w_tempty += w;
                    if (blockIdx.x == 0 && w > 1500) printf("[mma cta 0] tile %d waited %lld cyc for a free accumulator (at %lld)\n", (int)(t - t_lo), w, clock64() - t_start);
                }
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + as * BN;
                for (int kb = 0; kb < kblocks; ++kb) {
                    c0 = (dbg & 4) ? clock64() : 0;
                    mbar_wait(bar_full + 8 * stage, phase);
                    if (dbg & 4) w_full += clock64() - c0;
                    tc_fence_after();
                    const uint32_t sa = stage_base + stage * Cfg::STAGE_BYTES;
                    const uint64_t adesc = make_desc(sa), bdesc = make_desc(sa + Cfg::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UK; ++k) {
                        // +32 bytes per UMMA_K step inside the 128-byte swizzle atom
                        umma_bf16<CG>(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    }
                    umma_commit<CG>(bar_empty + 8 * stage);       // frees the smem stage when the MMAs retire
                    if (++stage == Cfg::STAGES) { stage = 0; phase ^= 1; }
                }
                umma_commit<CG>(bar_tfull + 8 * as);              // accumulator tile complete
                if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
            }
            if ((dbg & 4) && (blockIdx.x == 0 || blockIdx.x == 70))
                printf("[mma cta %d] total %lld cyc, wait accumulator-free %lld, wait smem-full %lld, tiles %d\n",
                       (int)blockIdx.x, clock64() - t_start, w_tempty, w_full, (int)(t_hi - t_lo));
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===== epilogue: scale, threshold, append, prune =====
        const int set = (warp - 4) >> 2;                           // column half of every tile this set drains
        const int quarter = warp & 3;                              // TMEM lanes [32*quarter, +32)
        const int row = quarter * 32 + lane;                       // query row within the CTA tile
        uint64_t* rows = bufs + ((size_t)set * BM + quarter * 32) * CAPT;
        uint64_t* mine = bufs + ((size_t)set * BM + row) * CAPT;
        const uint32_t head32 = (uint32_t)head, cap32 = (uint32_t)capacity;
        const int64_t qrow = q0 + row;
        const bool has_q = qrow < nq;
        // Padding rows (no query behind them: ragged batches, or the second CTA of a pair when fewer
        // than 129 queries come in) score 0 everywhere; with an open threshold each of them would
        // collect and prune candidates all sweep long.  +inf keeps them on the fast path.
        // dbg & 128: fast path only (no candidate ever passes).
        float thr = (!has_q || (dbg & 128)) ? INFINITY : -INFINITY;
        if (COLLECT && has_q) thr = __uint_as_float(__ldg(gthr + qrow));
        uint32_t* const cbuf = reinterpret_cast<uint32_t*>(partials) + (has_q ? qrow : 0) * (int64_t)krank;
        int cnt = 0, conv = 0;
        uint32_t aphase = 0;
        int as = 0;
        // cold-start protocol needs at least `krank` publishing lists for this query (krank = rank of
        // the list maxima used as threshold: comfortably below the k results the caller wants)
        const int nlists_q = min(slots_g * EPI_SETS, GMAX_LISTS);
        const bool cold = !COLLECT && nlists_q >= krank && !(dbg & 16);
        const int cold_need = max(krank, nlists_q - 4);
        const int my_list = slot * EPI_SETS + set;
        float rowmax = -INFINITY, pubmax = -INFINITY;             // this list's best score so far for this row / what it last published
        long long w_tfull = 0, t_start = (dbg & 4) ? clock64() : 0;
        int n_prune = 0, n_slow = 0;
        long long t_prune = 0, t_slow = 0, t_loop = 0, t_early = 0;
        int n_early = 0;
        constexpr int CPS = BN / EPI_SETS;                         // columns per set
        for (int64_t t = t_lo; t < t_hi; ++t) {
            const uint32_t n0 = (uint32_t)(t * BN);
            // Any list's KPT-th best score for this query bounds the global top-KPT from below, so
            // all CTAs sweeping the same query share one monotone threshold through gthr[query].
            const uint32_t g = (!COLLECT && has_q) ? __ldcg(gthr + qrow) : 0u;
            const float smax = __ldg(tile_max + t);                 // largest valid scale in this tile
            const int64_t ti = t - t_lo;
            const float* scb = ssc + (ti % SC_RING) * BN;
            const long long c0 = (dbg & 4) ? clock64() : 0;
            mbar_wait(bar_scfull + 8 * (int)(ti % SC_RING), (uint32_t)((ti / SC_RING) & 1));
            mbar_wait(bar_tfull + 8 * as, aphase);
            if (dbg & 4) w_tfull += clock64() - c0;
            tc_fence_after();
            if (!COLLECT && g > f2ord(thr)) thr = ord2f(g);
            const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN;
            // ---- cold start (first tile only) -------------------------------------------------
            // With thr = -inf every element of the first tiles is a candidate and every row prunes
            // again and again while the MMA pipe waits.  Instead: pass A only takes this list's best
            // score over the tile (no candidates kept) and publishes it; the krank-th largest of the
            // maxima published by the other CTAs sweeping the same query is a valid lower bound of
            // the query's global krank-th best (krank distinct rows reach it), so it becomes the initial
            // threshold; pass B is the normal scan of the same (still resident) accumulator tile.
            if (ti == 0 && cold && !(dbg & 1) && !a_pad) {
                // pass A: this list's best score over its columns of the first tile
#pragma unroll 1
                for (int c = set * CPS; c < (set + 1) * CPS; c += 16) {
                    uint32_t v[16];
                    tmem_ld16_issue(taddr0 + c, v);
                    float sc[16];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float4 f = *reinterpret_cast<const float4*>(scb + c + 4 * j);
                        sc[4 * j] = f.x; sc[4 * j + 1] = f.y; sc[4 * j + 2] = f.z; sc[4 * j + 3] = f.w;
                    }
                    tmem_ld_wait(v);
#pragma unroll
                    for (int j = 0; j < 16; ++j) rowmax = fmaxf(rowmax, __uint_as_float(v[j]) * sc[j]);
                }
                if (my_list < GMAX_LISTS && has_q && rowmax > -INFINITY) {
                    gmax[qrow * GMAX_LISTS + my_list] = f2ord(rowmax);
                    pubmax = rowmax;
                    __threadfence();
                }
                // krank-th largest of the maxima the lists sweeping this query have published: krank
                // distinct rows reach it, so it bounds the query's global krank-th best from below.
                // (A cheaper, weaker bound here was measured and lost: the first tiles are where a low
                // threshold costs most -- every row collects candidates and prunes.)
                uint32_t vals[32];
                for (int attempt = 0; attempt < 64; ++attempt) {
                    int have = 32;
                    if (has_q) {
                        have = 0;
                        const uint4* gp = reinterpret_cast<const uint4*>(gmax + qrow * GMAX_LISTS);
#pragma unroll
                        for (int u = 0; u < 8; ++u) {
                            const uint4 w4 = __ldcg(gp + u);
                            vals[4 * u] = w4.x; vals[4 * u + 1] = w4.y; vals[4 * u + 2] = w4.z; vals[4 * u + 3] = w4.w;
                        }
#pragma unroll
                        for (int i = 0; i < 32; ++i) have += vals[i] != 0u;
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) vals[i] = 0u;
                    }
                    if (__all_sync(FULL, have >= cold_need)) break;      // enough lists have published
                    __nanosleep(256);
                }
                const uint32_t tk = kth_largest_of_32(vals, krank - 1);
                if (tk != 0u && ord2f(tk) > thr) thr = ord2f(tk);
            }
            // One 16-column chunk: bound test on the raw accumulators, exact test and append on the
            // rare hit, prune full rows.
            auto process = [&](const uint32_t (&v)[16], const int c) {
                // Fast path: no per-row scale needed.  Every score of the chunk is at most
                // max(raw accumulator) * (largest scale of the tile), or 0 when that maximum is
                // not positive; fmaxf drops NaN.  One FMNMX tree, one multiply, one vote.
                float mx[4];
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    mx[g] = fmaxf(fmaxf(__uint_as_float(v[4 * g]), __uint_as_float(v[4 * g + 1])),
                                  fmaxf(__uint_as_float(v[4 * g + 2]), __uint_as_float(v[4 * g + 3])));
                const float mxc = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3]));
                const bool lane_hit = fmaxf(mxc * smax, 0.f) >= thr;
                if (!__any_sync(FULL, lane_hit)) return;
                const long long cs = (dbg & 4) ? clock64() : 0;
                // Slow path.  Only the lanes whose bound reached their threshold do anything, and
                // they only look at the groups of four whose own bound reached it (typically one
                // lane, one group): per-row scales for those four columns (LDS.128, same address in
                // every active lane), exact approximate scores, append.  No further votes: the rest
                // of the warp waits at the reconvergence point.
                auto scan_groups = [&](const int g0) {
                    if (lane_hit) {
#pragma unroll
                        for (int g = g0; g < g0 + 2; ++g) {
                            if (fmaxf(mx[g] * smax, 0.f) >= thr) {
                                if (dbg & 4) ++n_slow;
                                const float4 f = *reinterpret_cast<const float4*>(scb + c + 4 * g);
                                const float sc4[4] = {f.x, f.y, f.z, f.w};
#pragma unroll
                                for (int j = 0; j < 4; ++j) {
                                    const float sj = __uint_as_float(v[4 * g + j]) * sc4[j];
                                    if (sj >= thr) {                      // NaN scale (filtered row): false
                                        if (COLLECT) {
                                            const uint32_t pos = atomicAdd(gmax + qrow, 1u);
                                            if (pos < (uint32_t)krank) cbuf[pos] = n0 + c + 4 * g + j;
                                        } else {
                                            mine[cnt++] = ((uint64_t)f2ord(sj) << 32) | (n0 + c + 4 * g + j);
                                            rowmax = fmaxf(rowmax, sj);
                                        }
                                    }
                                }
                            }
                        }
                    }
                    __syncwarp();
                };
                scan_groups(0);
                if (KPT == 32) {                               // 32-entry lists: fullness check every 8 columns
                    const unsigned full = __ballot_sync(FULL, cnt > CAPT - 8);
                    if (full) {
                        warp_prune_rows<CAPT, KPT>(full, rows, lane, cnt, conv, thr, head32, cap32);
                        if (((full >> lane) & 1u) && has_q) atomicMax(gthr + qrow, f2ord(thr));
                    }
                }
                scan_groups(2);
                if (KPT == 32) {
                    const unsigned full = __ballot_sync(FULL, cnt > CAPT - 8);
                    if (full) {
                        warp_prune_rows<CAPT, KPT>(full, rows, lane, cnt, conv, thr, head32, cap32);
                        if (((full >> lane) & 1u) && has_q) atomicMax(gthr + qrow, f2ord(thr));
                    }
                }
                if (dbg & 4) t_slow += clock64() - cs;
                if (KPT <= 16) {                               // 8/16-entry lists: once per 16 columns
                    const unsigned full = __ballot_sync(FULL, cnt > CAPT - 16);
                    if (full) {
                        const long long cp = (dbg & 4) ? clock64() : 0;
                        warp_prune_rows<CAPT, KPT>(full, rows, lane, cnt, conv, thr, head32, cap32);
                        if (((full >> lane) & 1u) && has_q) atomicMax(gthr + qrow, f2ord(thr));
                        if (dbg & 4) { n_prune += __popc(full); t_prune += clock64() - cp; }
                    }
                }
            };
            // Software pipeline over the set's 128 columns: the TMEM load of the next chunk is in
            // flight while the current one is processed (two register buffers, ping-pong).
            const long long cl = (dbg & 4) ? clock64() : 0;
            if (!(dbg & 1) && !a_pad) {
                uint32_t va[16], vb[16];
                const int c_lo = set * CPS, c_hi = (set + 1) * CPS;
                tmem_ld16_issue(taddr0 + c_lo, va);
#pragma unroll 1
                for (int c = c_lo; c < c_hi; c += 32) {
                    tmem_ld_wait(va);
                    tmem_ld16_issue(taddr0 + c + 16, vb);
                    if (!(dbg & 2)) process(va, c);
                    tmem_ld_wait(vb);
                    if (c + 32 < c_hi) tmem_ld16_issue(taddr0 + c + 32, va);
                    if (!(dbg & 2)) process(vb, c + 16);
                }
            }
            if (dbg & 4) t_loop += clock64() - cl;
            if (!COLLECT && cold && rowmax > pubmax && my_list < GMAX_LISTS && has_q) {   // a new best for this list: publish it
                gmax[qrow * GMAX_LISTS + my_list] = f2ord(rowmax);
                pubmax = rowmax;
            }
            // release this accumulator stage to the MMA issuer (in the leader CTA)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (CG == 2) mbar_arrive_cluster(bar_tempty + 8 * as, 0);
                else mbar_arrive(bar_tempty + 8 * as);
                mbar_arrive(bar_scempty + 8 * (int)(ti % SC_RING));     // scale slot may be refilled
            }
            if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
            // ---- threshold refresh (after tiles 1, 3, 7, 15, ...) ------------------------------
            // The bound keeps improving as the sweep goes on: every list publishes its running
            // maximum whenever it grows (above), and the krank-th largest of the 32 published maxima
            // is a lower bound of the query's global krank-th best at ANY time.  It is far tighter
            // than what the lists can tell each other through gthr (the best KPT-th entry of a single
            // list: one list sees 1/36 of the rows), and the fraction of rows that pass the threshold
            // falls like 1/(rows seen), so re-reading it at every power of two keeps the pass rate
            // within 2x of what a per-tile refresh would give, for a handful of sorts per sweep --
            // done here, after the accumulator stage went back to the MMA issuer, where the epilogue
            // would otherwise wait for the next tile.
            if (cold && ti > 0 && ((ti + 1) & ti) == 0 && !(dbg & (1 | 256)) && !a_pad) {
                uint32_t vals[32];
                if (has_q) {
                    const uint4* gp = reinterpret_cast<const uint4*>(gmax + qrow * GMAX_LISTS);
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        const uint4 w4 = __ldcg(gp + u);
                        vals[4 * u] = w4.x; vals[4 * u + 1] = w4.y; vals[4 * u + 2] = w4.z; vals[4 * u + 3] = w4.w;
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < 32; ++i) vals[i] = 0u;
                }
                const uint32_t tk = kth_largest_of_32(vals, krank - 1);
                if (tk != 0u && ord2f(tk) > thr) thr = ord2f(tk);
            }
            // Idle-time pruning: if the next accumulator is not ready yet, shrink the rows that are
            // filling up now, off the critical path, so that prunes inside the column loop (which
            // hold the accumulator stage and stall the MMA issuer) become rare.
            if (!(dbg & 32) && t + 1 < t_hi && !mbar_test(bar_tfull + 8 * as, aphase)) {
                const unsigned early = __ballot_sync(FULL, cnt > (KPT == 8 ? 10 : KPT + 8));
                if (early) {
                    const long long cp = (dbg & 4) ? clock64() : 0;
                    warp_prune_rows<CAPT, KPT>(early, rows, lane, cnt, conv, thr, head32, cap32);
                    if (((early >> lane) & 1u) && has_q) atomicMax(gthr + qrow, f2ord(thr));
                    if (dbg & 4) { n_early += __popc(early); t_early += clock64() - cp; }
                }
            }
        }
        if ((dbg & 4) && lane == 0 && (blockIdx.x == 0 || blockIdx.x == 70) && (warp == 4 || warp == 9))
            printf("[epi cta %d warp %d] total %lld cyc, wait accumulator-full %lld, column loop %lld (slow path %lld in %d groups, prune %lld for %d rows), idle-time prune %lld for %d rows\n",
                   (int)blockIdx.x, warp, clock64() - t_start, w_tfull, t_loop, t_slow, n_slow, t_prune, n_prune, t_early, n_early);
        // flush: final prune of every row (four at a time), write this (slot, set) candidate list
        if (!COLLECT) {
            warp_prune_all_rows<CAPT, KPT>(rows, lane, cnt, conv, thr, head32, cap32);
            // every row this list ever dropped scored below its final threshold: finish_kernel's gap
            // check needs the largest of them per query
            if (has_q && thr > -INFINITY) atomicMax(gthr + qrow, f2ord(thr));
            for (int r = 0; r < 32; ++r) {
                const int c = __shfl_sync(FULL, cnt, r);
                const int64_t q = q0 + quarter * 32 + r;
                if (q < nq && lane < KPT)
                    partials[(q * max_lists + slot * EPI_SETS + set) * KPT + lane] = lane < c ? rows[r * CAPT + lane] : 0ull;
            }
        }
    }

    // ---- teardown ----
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<CG>(tmem_base, 512);
    }
}

int scan_tc_lists_per_slot() { return EPI_SETS; }

int scan_tc_rows_per_group(int cg) { return BM * cg; }

// Work units per query group: `*rem` groups get one more than the returned base count.
int scan_tc_slots(int64_t nq, int64_t nslots, int cg, int ctas, int* rem) {
    const int64_t rows_per_unit = (int64_t)BM * cg;
    const int64_t groups = (nq + rows_per_unit - 1) / rows_per_unit;
    const int64_t units = ctas / cg;
    const int64_t ntiles = (nslots + BN - 1) / BN;
    int64_t s = units / groups;
    // Leftover units (148 SMs = 74 pairs = 4 groups x 18 + 2) stay idle on purpose: giving two
    // groups a 19th unit makes the groups sweep the corpus out of step, the L2 sharing of corpus
    // tiles between groups breaks, and the whole batch gets SLOWER (measured: 1.33 -> 1.39 ms,
    // mainloop alone 1.12 -> 1.19 ms).  The kernel supports uneven counts; the policy does not use them.
    *rem = 0;
    if (s >= ntiles) { s = ntiles; *rem = 0; }
    if (s < 1) { s = 1; *rem = 0; }
    return (int)s;
}

template <int CG, int KPT, bool COLLECT = false>
static cudaError_t launch_one(const ScanArgs& a, const CUtensorMap* map_q, const CUtensorMap* map_c,
                              cudaStream_t st) {
    const int rows_per_unit = BM * CG;
    const int groups = (int)((a.nq + rows_per_unit - 1) / rows_per_unit);
    const int units = groups * a.slots + a.slots_rem;
    const int kblocks = a.ld / BK;
    cudaLaunchConfig_t cfg{};
    cfg.blockDim = dim3(TC_THREADS);
    cfg.gridDim = dim3(units * CG);
    cfg.dynamicSmemBytes = TcCfg<CG, KPT>::SMEM;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;       // see common.cuh (pdl_trigger / pdl_wait)
    attr[1].val.programmaticStreamSerializationAllowed = pdl_enabled();
    cfg.attrs = attr; cfg.numAttrs = 2;
    return cudaLaunchKernelEx(&cfg, scan_tc_kernel<CG, KPT, COLLECT>, *map_q, *map_c, a.scale, a.tile_max, a.partials, a.gthr, a.gmax,
                              a.progress, a.window, a.nq, a.nslots, a.capacity, a.head, kblocks, a.slots, a.slots_rem, groups,
                              COLLECT ? a.collect_cap : a.krank, a.dbg);
}

cudaError_t scan_tc_init() {
    cudaError_t e;
#define PILOTT_TC_ATTR(CG, KPT)                                                                     \
    e = cudaFuncSetAttribute(scan_tc_kernel<CG, KPT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,  \
                             TcCfg<CG, KPT>::SMEM);                                                      \
    if (e != cudaSuccess) return e;
    PILOTT_TC_ATTR(1, 8) PILOTT_TC_ATTR(1, 16) PILOTT_TC_ATTR(1, 32)
    PILOTT_TC_ATTR(2, 8) PILOTT_TC_ATTR(2, 16) PILOTT_TC_ATTR(2, 32)
#undef PILOTT_TC_ATTR
    e = cudaFuncSetAttribute(scan_tc_kernel<2, 8, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TcCfg<2, 8>::SMEM);
    if (e != cudaSuccess) return e;
    return cudaSuccess;
}

// a.list_len selects how many candidates each (query, slot, set) list keeps: 16 or 32.
cudaError_t launch_scan_tc(const ScanArgs& a, const CUtensorMap* map_q, const CUtensorMap* map_c,
                           int cg, int ctas, cudaStream_t st) {
    (void)ctas;
    if (a.collect_cap > 0) return launch_one<2, 8, true>(a, map_q, map_c, st);     // pair kernel only
    if (cg == 1) {
        if (a.list_len == 8) return launch_one<1, 8>(a, map_q, map_c, st);
        return a.list_len == 16 ? launch_one<1, 16>(a, map_q, map_c, st) : launch_one<1, 32>(a, map_q, map_c, st);
    }
    if (a.list_len == 8) return launch_one<2, 8>(a, map_q, map_c, st);
    return a.list_len == 16 ? launch_one<2, 16>(a, map_q, map_c, st) : launch_one<2, 32>(a, map_q, map_c, st);
}

}  // namespace pilott

static_assert(pilott::TcCfg<1, 8>::SMEM <= 232448 && pilott::TcCfg<1, 32>::SMEM <= 232448 &&
              pilott::TcCfg<2, 8>::SMEM <= 232448 && pilott::TcCfg<2, 32>::SMEM <= 232448,
              "shared memory budget (227 KB per CTA)");
