// internal.h -- host-side launch wrappers shared between the translation units.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

namespace pilott {

// PILOTT_PDL=0 in the environment turns programmatic dependent launch off (plain stream order).
inline int pdl_enabled() {
    static const int on = [] { const char* e = getenv("PILOTT_PDL"); return (e && e[0] == '0') ? 0 : 1; }();
    return on;
}

// Launch with programmatic stream serialization (see common.cuh: pdl_trigger / pdl_wait).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled();
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

struct ScanArgs {
    const void* qb;          // bf16 [nq_pad, ld] converted queries
    const void* vec;         // bf16 [capacity, ld]
    const float* scale;      // [cap_pad]: inverse norm, NaN for masked / unused slots
    const float* tile_max;   // [cap_pad/256]: largest valid scale of each 256-row tile
    uint64_t* partials;      // [nq][slots][KP] candidate keys
    uint32_t* gthr;          // [nq] shared per-query threshold (ordered-uint score), zeroed per search
    uint32_t* gmax;          // [nq][32] first-tile maximum of each candidate list (ordered-uint), zeroed per search
    int64_t nq;
    int64_t nslots;          // physical slots holding rows (= min(total, capacity))
    int64_t capacity;
    int64_t head;            // physical slot of the oldest row
    int ld;                  // row stride in elements (multiple of 64)
    int slots;               // work units per query group (base count)
    int slots_rem;           // this many groups (the first ones) get one more unit
    int list_len;            // candidates per list the scan writes (16 or 32)
    int krank;               // rank of the first-tile list maxima used as cold-start threshold (16 or 32)
    int dbg;                 // developer switches for the tcgen05 kernel (0 in production)
    uint32_t* progress;      // [units] tiles issued by each work unit (zeroed per search), or NULL: the groups of a
                             // slot keep within `window` tiles of each other (long sweeps only)
    int window;
    int collect_cap;         // > 0: second-chance pass (pair kernel): partials = uint32 [nq][collect_cap] ring slots,
                             // gthr = float-bit thresholds [nq], gmax = uint32 counters [nq]
};

// generic CUDA-core scan (any shape).  Returns slots used via args.slots (caller sets).
cudaError_t launch_scan_generic(const ScanArgs& a, int num_sms, cudaStream_t st);
int scan_generic_slots(int64_t nq, int64_t nslots, int num_sms);

// tcgen05 scan.  cg = 1 or 2.
cudaError_t launch_scan_tc(const ScanArgs& a, const CUtensorMap* map_q, const CUtensorMap* map_c,
                           int cg, int ctas, cudaStream_t st);
int scan_tc_slots(int64_t nq, int64_t nslots, int cg, int ctas, int* rem);
int scan_tc_rows_per_group(int cg);
int scan_tc_lists_per_slot();   // candidate lists each (query, slot) writes; partials hold slots*this lists
cudaError_t scan_tc_init();   // opt-in shared memory attributes

cudaError_t launch_append(const float* x, int64_t n, int dim, int ld, void* vec, float* inv_norm,
                          int32_t* prio, uint64_t* tagmask, int64_t* expires,
                          const int32_t* in_prio, const uint64_t* in_tag, const int64_t* in_exp,
                          int64_t first_slot, int64_t capacity, cudaStream_t st);
cudaError_t launch_convert_queries(const float* q, int64_t nq, int dim, int ld, void* qb,
                                   cudaStream_t st);
cudaError_t launch_scale(const float* inv_norm, const int32_t* prio, const uint64_t* tagmask,
                         const int64_t* expires, int64_t nslots, int64_t cap_pad, int32_t min_prio,
                         uint64_t req, int64_t now_us, float* scale, float* tile_max, cudaStream_t st);

cudaError_t launch_prepare(const float* q, int64_t nq, int dim, int ld, void* qb, const float* inv_norm,
                           const int32_t* prio, const uint64_t* tagmask, const int64_t* expires, int64_t nslots,
                           int64_t capacity, int64_t cap_pad, int32_t min_prio, uint64_t req, int64_t now_us,
                           float* scale, float* tile_max, uint32_t* zero_words, int64_t n_zero_words,
                           int32_t* inexact, bool do_scale, cudaStream_t st);

struct FinishArgs {
    const uint64_t* partials; int slots; int list_len;   // slots = allocated lists per query
    int pool_n;                 // survivors the merge keeps per query (32)
    const uint32_t* gthr;       // [nq] largest threshold any list of the query ended with (may be NULL)
    int rows_per_group, lists_base, groups_rem, lists_extra;  // lists of query q: lists_base + (q/rows_per_group < groups_rem ? lists_extra : 0)
    const void* qb; const void* vec; int ld;
    int64_t nq, capacity, head, base_seq;
    int k; double err;
    int64_t seq_mul, seq_add;   // reported index = seq * seq_mul + seq_add
    int64_t* out_idx; float* out_score; double* out_score64;
    uint8_t* flags; int32_t* inexact;
    // Row-sharded store: when `peers` is set the kernel also stores every result row into each
    // rank's exchange buffer over NVLink (xchg.cu layout) and the last CTA releases the flags.
    unsigned char* const* peers = nullptr;
    int world = 1, rank = 0;
    int64_t block_bytes = 0, flags_off = 0, step = 0;
    uint32_t* done_counter = nullptr;
    unsigned long long* flagged_total = nullptr;   // handle-lifetime count of flagged queries (never zeroed by a search)
    bool legacy = false;        // true: the original one-kernel form that re-scores all pool_n survivors
};
cudaError_t launch_finish(const FinishArgs& a, cudaStream_t st);


// exchange over NVLink peer memory (xchg.cu)
// value of a slot's flag for step s: never 0 (the cleared state), exact within any window of 65536 steps
#define PILOTT_XCHG_FLAG(step) ((uint64_t)(((step) & 0xffff) + 1))
constexpr int XCHG_FLAG_RING = 65536;
constexpr int XCHG_SLOTS = 8;   // steps of the exchange that may be in flight before a slot is reused (power of two)
cudaError_t launch_xchg_publish(unsigned char* const* peers_dev, int world, int rank, int64_t block_bytes,
                                int64_t flags_off, const void* src, int64_t bytes, int64_t step, cudaStream_t st);
cudaError_t launch_xchg_merge(const void* mine, int world, int64_t block_bytes, int64_t flags_off, int64_t step,
                              int64_t nq, int k, int64_t* out_idx, float* out_score, double* out_score64,
                              int32_t* err_word, cudaStream_t st);

// second-chance pass for flagged queries (finish.cu); the COLLECT scan runs between the two
cudaError_t launch_refine_prepare(const void* qb, const int32_t* qids, int64_t nsel, int64_t nsel_pad, int ld,
                                  const int64_t* prior_idx, const double* prior_s64, int k, double err, void* qb2,
                                  uint32_t* thr_bits, uint32_t* cnt, int32_t* n_overflow, cudaStream_t st);
cudaError_t launch_refine_finish(const uint32_t* cand, const uint32_t* cnt, int cap, const void* qb2, const void* vec,
                                 int ld, const int32_t* qids, int64_t nsel, int64_t capacity, int64_t head,
                                 int64_t base_seq, int64_t seq_mul, int64_t seq_add, int k, int64_t* out_idx,
                                 float* out_score, double* out_score64, uint8_t* overflow, int32_t* n_overflow,
                                 const double* tcos, double err, double* hist_t, uint32_t* hist_c, cudaStream_t st);
cudaError_t launch_largek_prepare(const void* qb, const int32_t* qids, int64_t nsel, int64_t nsel_pad, int ld,
                                  const int64_t* top_idx, const double* top_s64, int k, double factor, double min_gap,
                                  double n_rows, const double* hist_t, const uint32_t* hist_c, double cap_rows, void* qb2,
                                  uint32_t* thr_bits, double* tcos, uint32_t* cnt, int32_t* n_overflow, cudaStream_t st);

cudaError_t launch_exact(const void* qb, const void* vec, int ld, const int32_t* prio,
                         const uint64_t* tagmask, const int64_t* expires, int64_t nslots,
                         int64_t capacity, int64_t head, int64_t base_seq, int64_t seq_mul, int64_t seq_add,
                         const int32_t* qids,
                         int64_t nsel, int k, int32_t min_prio, uint64_t req, int64_t now_us,
                         double* scratch, int64_t* out_idx, float* out_score, double* out_score64,
                         cudaStream_t st);

cudaError_t launch_topk_merge(const int64_t* idx, const double* sc, int shards, int64_t nq, int k,
                              int64_t shard_stride, int64_t* out_idx, float* out_score,
                              double* out_score64, cudaStream_t st);

cudaError_t launch_compact_flags(const int64_t* expires, int64_t count, int64_t head,
                                 int64_t capacity, int64_t now_us, uint8_t* keep, cudaStream_t st);
cudaError_t launch_gather_rows(const void* vec, const float* inv_norm, const int32_t* prio,
                               const uint64_t* tagmask, const int64_t* expires, int ld,
                               const int64_t* src_slot, int64_t n, void* dvec, float* dinv,
                               int32_t* dprio, uint64_t* dtag, int64_t* dexp, cudaStream_t st);

cudaError_t launch_route(const uint64_t* task_req, const uint64_t* task_tags,
                         const int32_t* task_type, const uint8_t* task_flags, int64_t nt,
                         const uint64_t* agent_caps,
                         const uint64_t* agent_spec, const uint8_t* agent_flags,
                         const double* agent_s12, const double* agent_t4, int na, int words,
                         const double* spec_u, int max_tags, int mode, int32_t* out_agent,
                         double* out_score, cudaStream_t st);
// tcgen05 form of the same scoring (route_tc.cu): bit sets -> bf16 feature matrices in `workspace` -> GEMM + epilogue
int64_t route_tc_workspace_bytes(int64_t nt, int na, int words);
bool route_tc_supported(int na, int words, int max_tags);
cudaError_t launch_route_tc(const uint64_t* task_req, const uint64_t* task_tags, const int32_t* task_type,
                            const uint8_t* task_flags, int64_t nt, const uint64_t* agent_caps, const uint64_t* agent_spec,
                            const uint8_t* agent_flags, const double* agent_s12, const double* agent_t4, int na, int words,
                            const double* spec_u, int max_tags, int mode, int32_t* out_agent, double* out_score,
                            void* workspace, int num_sms, int (*encode)(CUtensorMap*, const void*, int64_t, int, int),
                            cudaStream_t st);
cudaError_t launch_task_priority(const uint8_t* urgent, const int32_t* complexity,
                                 const int32_t* ndeps, int64_t nt, uint8_t* out, cudaStream_t st);

}  // namespace pilott
