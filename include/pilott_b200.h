/* pilott_b200.h -- C ABI of the B200-native PilottAI hot path.
 *
 * One shared library (libpilott_b200.so), plain pointers and sizes, no torch or
 * C++ types in any signature.  Every entry point returns 0 on success or a
 * negative PILOTT_E* code; pilott_last_error() then describes the failure
 * (thread-local string).  Pointers named *_dev are device pointers on the
 * store's GPU, *_host are ordinary host pointers.  `stream` is a cudaStream_t
 * passed as void* (NULL = the legacy default stream).  Calls on one handle must
 * be serialised by the caller (the reference serialises them with asyncio locks:
 * pilott/memory/enhanced_memory.py:85,103; pilott/core/router.py:37); different
 * handles are independent.
 *
 * Streams: entry points that take device pointers enqueue on the CALLER's `stream`; the
 * *_host entry points enqueue on a stream the handle owns.  The library orders the two
 * itself (an event behind every call, waited for by the next call of the other family), so a
 * device append followed by a host search sees the new rows, and a host search still in
 * flight (pilott_mem_search_host_begin) is not disturbed by a device search that follows.
 * Two device-pointer calls on DIFFERENT caller streams are the caller's to order: they
 * share the handle's scratch buffers.
 *
 * Citations are paths under the reference tree (appyjain21/pilottai).
 */
#ifndef PILOTT_B200_H
#define PILOTT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PILOTT_OK 0
#define PILOTT_EINVAL (-1)   /* bad argument */
#define PILOTT_ECUDA (-2)    /* CUDA runtime / driver error */
#define PILOTT_ENOMEM (-3)   /* device or host allocation failed */
#define PILOTT_EUNSUPPORTED (-4)

/* Largest k the fused similarity + top-k path answers; larger k (and queries the
 * fused path flags as not provably exact) go through pilott_mem_search_exact. */
#define PILOTT_MAX_FUSED_K 32
/* Largest k the threshold-collect path answers (pilott_mem_search_large); beyond it: brute force. */
#define PILOTT_MAX_LARGE_K 1024
/* Tags are stored as one 64-bit mask per row. */
#define PILOTT_MAX_TAGS 64

typedef struct pilott_mem pilott_mem_t;

const char* pilott_last_error(void);
int pilott_version(void);
/* Candidates the fused path keeps per query before the exact re-score (32). */
int pilott_kp(void);

/* ---- semantic store ----------------------------------------------------------------
 * Replaces EnhancedMemory.__init__'s bounded deque (pilott/memory/enhanced_memory.py:25-47):
 * a device-resident ring of `capacity` rows; the oldest row is overwritten when full
 * (deque(maxlen) semantics, :27,:86).  Per row: bf16[dim] vector (row stride = dim
 * rounded up to 64), fp32 inverse norm, int32 priority, uint64 tag mask, int64
 * expires_at in microseconds (0 = never), implicit int64 insertion sequence. */
int pilott_mem_create(int dim, int64_t capacity, int device, pilott_mem_t** out);
int pilott_mem_destroy(pilott_mem_t* h);
/* EnhancedMemory.clear, enhanced_memory.py:284-292 (semantic store part). */
int pilott_mem_clear(pilott_mem_t* h, void* stream);
/* total = rows ever appended, count = rows currently held. */
int pilott_mem_info(const pilott_mem_t* h, int64_t* total, int64_t* count, int64_t* capacity,
                    int* dim, int* row_stride);
/* which = 0: scan kernel (0 auto, 1 generic CUDA-core, 2 tcgen05 1-CTA, 3 tcgen05 CTA-pair).
 * which = 1: CTAs used by the scan (0 = one per SM).
 * which = 2: developer switches of the scan kernel; they exist only in a -DPILOTT_DEBUG build, the
 *            production library answers PILOTT_EUNSUPPORTED to any value but 0.
 * which = 3: candidates each partial list keeps: 0 auto (8 when k <= 10 and the pair scan produces >= 16 lists per query; 16 when k <= 12 or >= 8 lists; else 32), 8, 16, 32.
 * which = 4: 1 = record CUDA events around the scan kernel (read with pilott_mem_profile_read).
 * which = 5 / 6: reported index = sequence * stride (5) + offset (6); a row-sharded store that
 *            holds global rows i with i % world == rank uses stride = world, offset = rank.
 * which = 7: set `total` (rows ever appended, >= rows held): a restored snapshot continues the
 *            sequence numbering of the store it was taken from.
 * which = 9: form of the finish step: 0 selective (re-scores only the survivors that can reach the
 *            top-k and lets the gap check prove the cut; default), 1 the original kernel that
 *            re-scores all 32 survivors.  Same results. */
int pilott_mem_set_option(pilott_mem_t* h, int which, int64_t value);
/* Kernel launches enqueued by this handle since creation (bench bookkeeping). */
int64_t pilott_mem_launch_count(const pilott_mem_t* h);
/* How many queries the gap check of the fused path flagged ("not provably exact") in all searches on
 * this handle since creation or since the last call with reset != 0 -- unlike *out_inexact_dev,
 * which every search zeroes, nothing is lost between batches, so a pipeline that keeps several
 * searches in flight (and a row-sharded job, which sums it over ranks) can assert "no query went
 * unproven" once at the end.  Synchronises the device.  No counterpart in the reference. */
int pilott_mem_flagged_total(pilott_mem_t* h, int64_t* total, int reset);
/* Waits for the recorded scan kernels, returns their summed device time and count since the
 * last read (option 4 must be on), and resets the window. */
int pilott_mem_profile_read(pilott_mem_t* h, double* scan_ms_total, int64_t* n_scans);

/* Replaces EnhancedMemory.store_semantic + _update_indexes (enhanced_memory.py:60-91,
 * :133-144) for a batch of n rows: fp32 -> bf16 (RNE), fused sum of squares -> inverse
 * norm, metadata columns, ring head advance.  x is [n, dim] fp32 row-major.  prio /
 * tagmask / expires_us may be NULL (zeros).  *first_seq receives the insertion
 * sequence number of row 0 (host int64, may be NULL).  n must be <= capacity. */
int pilott_mem_append(pilott_mem_t* h, const float* x_dev, int64_t n, const int32_t* prio_dev,
                      const uint64_t* tagmask_dev, const int64_t* expires_us_dev,
                      int64_t* first_seq, void* stream);
/* Same with host buffers: stages through pinned memory on the handle's own stream and
 * returns after the rows are resident. */
int pilott_mem_append_host(pilott_mem_t* h, const float* x_host, int64_t n,
                           const int32_t* prio_host, const uint64_t* tagmask_host,
                           const int64_t* expires_us_host, int64_t* first_seq);

/* Replaces EnhancedMemory.semantic_search + _get_candidate_indexes + MemoryItem.is_expired
 * (enhanced_memory.py:93-131, :20-21) for a batch of nq queries, with cosine similarity in
 * place of the reference's substring test (:110):
 *   candidates: priority >= min_prio AND (tagmask & req_tagmask) == req_tagmask
 *               AND NOT (expires_us != 0 AND now_us > expires_us)
 *   score:      cosine over bf16-rounded inputs (exact fp64 re-score of the survivors of
 *               the tensor-core pass: eight interleaved fp64 partial sums, fixed combining tree,
 *               exactly as oracle/retrieval_oracle.py states)
 *   rank:       (score desc, insertion sequence asc); k results per query.
 * q is [nq, dim] fp32.  out_idx [nq,k] receives insertion sequence numbers (-1 = fewer than
 * k candidates), out_score [nq,k] fp32 (-inf padding), out_score64 [nq,k] fp64 may be NULL.
 * *out_inexact_dev (device int32, may be NULL) receives how many queries could NOT be
 * proven exact (their gap check failed) and out_flags_dev [nq] (uint8, may be NULL) marks
 * them; re-run those through pilott_mem_search_refine (or pilott_mem_search_exact).
 * 1 <= k <= PILOTT_MAX_FUSED_K. */
int pilott_mem_search(pilott_mem_t* h, const float* q_dev, int64_t nq, int k, int32_t min_prio,
                      uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev,
                      float* out_score_dev, double* out_score64_dev, int32_t* out_inexact_dev,
                      uint8_t* out_flags_dev, void* stream);
/* Host-buffer variant: H2D of the queries, the search, the second-chance pass for flagged
 * queries (brute force for k > PILOTT_MAX_FUSED_K), D2H of the results; synchronous.
 * *n_fallback (may be NULL) = queries that needed more than the fused pass. */
int pilott_mem_search_host(pilott_mem_t* h, const float* q_host, int64_t nq, int k,
                           int32_t min_prio, uint64_t req_tagmask, int64_t now_us,
                           int64_t* out_idx_host, float* out_score_host, double* out_score64_host,
                           int32_t* n_fallback);
/* The host-buffer search split in two, so that a caller can keep several batches in flight:
 * _begin copies the queries to the device on a copy stream, enqueues the search behind it on the
 * handle's compute stream and the copy of the results back on a second copy stream, and returns
 * without waiting; _end waits for that slot's results, runs the second-chance pass if the batch
 * flagged queries, and fills the outputs exactly like pilott_mem_search_host.  With two slots the
 * H2D copy of batch i+1 and the D2H copy of batch i-1 overlap the search of batch i.  4 slots per
 * handle; a slot is busy from _begin to _end; q_host must stay valid until _end when it is
 * page-locked memory (pageable memory is copied into a pinned staging buffer inside _begin). */
int pilott_mem_search_host_begin(pilott_mem_t* h, int slot, const float* q_host, int64_t nq, int k,
                                 int32_t min_prio, uint64_t req_tagmask, int64_t now_us);
int pilott_mem_search_host_end(pilott_mem_t* h, int slot, int64_t* out_idx_host, float* out_score_host,
                               double* out_score64_host, int32_t* n_fallback);
/* Top-k for PILOTT_MAX_FUSED_K < k <= PILOTT_MAX_LARGE_K without scoring every row in fp64 (any
 * other k is handed to the fused or the brute-force path).  The fused top-32 of each query gives
 * two points of its score tail; a per-query threshold extrapolated to rank ~1.5k is applied by the
 * tensor-core scan in COLLECT mode; the collected rows are re-scored by the stated rule, ranked,
 * and the guess is CHECKED (the k-th exact score must clear every row that was not collected);
 * queries whose guess was too high (or collected more than 4096 rows) repeat with a guess refitted
 * from the row count the pass measured, at most twice, then with an open threshold.  Same results as pilott_mem_search_exact, bit for bit.  *n_bruteforce (host, may
 * be NULL): queries that ended in the brute-force kernels (more than 4096 rows above their
 * threshold).  Synchronises `stream` between passes.  The host-buffer entry points use this path
 * for k > PILOTT_MAX_FUSED_K.  No counterpart in the reference. */
int pilott_mem_search_large(pilott_mem_t* h, const float* q_dev, int64_t nq, int k, int32_t min_prio,
                            uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev, float* out_score_dev,
                            double* out_score64_dev, int32_t* n_bruteforce, void* stream);
/* Brute-force fp64 search (any k >= 1): every candidate row scored in the oracle's order.
 * Slow path for flagged queries / large k.  query_ids_dev (int32[nsel], may be NULL = all
 * nq) selects which rows of q/out to process. */
int pilott_mem_search_exact(pilott_mem_t* h, const float* q_dev, int64_t nq, const int32_t* query_ids_dev,
                            int64_t nsel, int k, int32_t min_prio, uint64_t req_tagmask,
                            int64_t now_us, int64_t* out_idx_dev, float* out_score_dev,
                            double* out_score64_dev, void* stream);

/* Second chance for the queries pilott_mem_search flagged (no counterpart in the reference; it is
 * the fused path's own repair step).  Same selection arguments as pilott_mem_search_exact, same
 * bit-identical results, but instead of scoring every row it re-runs the tensor-core scan over the
 * flagged queries only, with ONE fixed threshold per query: out_idx_dev / out_score64_dev must still
 * hold what pilott_mem_search wrote for these queries (k rows with exact fp64 scores; the k-th is a
 * lower bound of the true k-th best), so every row that can belong to the true top-k has an
 * approximate score above (bound - error margin) and is collected, re-scored by the stated rule and
 * ranked.  Cost: about one scan of the corpus for up to 256 flagged queries, whatever made them
 * ambiguous (bursts of near-duplicate rows, exact ties wider than a candidate list).  Queries with
 * more than 1024 rows inside the margin, k > PILOTT_MAX_FUSED_K, out_score64_dev == NULL or
 * query_ids_dev == NULL fall through to the brute-force path; *n_bruteforce (host, may be NULL)
 * says how many did.  The filters must be the ones the flagged search used.  Synchronises `stream`
 * before returning (it reads an overflow counter). */
int pilott_mem_search_refine(pilott_mem_t* h, const float* q_dev, int64_t nq, const int32_t* query_ids_dev,
                             int64_t nsel, int k, int32_t min_prio, uint64_t req_tagmask,
                             int64_t now_us, int64_t* out_idx_dev, float* out_score_dev,
                             double* out_score64_dev, int32_t* n_bruteforce, void* stream);

/* Same search as pilott_mem_search_exact, same arguments and bit-identical results, but the
 * call allocates nothing.  It has no counterpart in the reference.
 *
 * Who needs it: code that must not (or cannot) reach the CUDA allocator while it runs --
 * PilottAI's single asyncio loop thread, where the device-wide synchronisation inside
 * cudaMalloc/cudaFree would stall every other stream and coroutine; and processes in which a
 * framework's caching allocator (PyTorch's, typically) already owns nearly all of HBM, so a
 * raw cudaMalloc from this library fails although the store itself is resident.
 * pilott_mem_search_exact grows per-handle scratch with cudaMalloc on demand; this variant
 * keeps its scratch (converted queries + one fp64 score per selected query per stored row) in
 * a fixed PILOTT_NOALLOC_SCRATCH_BYTES array inside the library's device image, which every
 * GPU that loads the library holds until the process exits.  It calls no cudaMalloc, cudaFree
 * or other allocating routine.
 *
 * Limit: round_up(nq * row_stride * 2, 256) + nsel * rows_held * 8 <= PILOTT_NOALLOC_SCRATCH_BYTES
 * (e.g. 20 selected queries against 100 000 rows), and nsel <= 65535; beyond that the call
 * fails with PILOTT_EINVAL and a message that states both numbers.
 * Set-up (the lookup of the array's device address with cudaGetSymbolAddress, which is also what
 * makes CUDA load the library's device image -- array included -- on that device) happens once
 * per device inside pilott_mem_create, never in this call; create your handles before starting
 * any stream capture.
 * The array is shared by every call on the device and keeps the previous call's contents.
 * Nothing in it needs to start at zero (each call rewrites what it reads), but two calls that
 * overlap on the GPU -- different streams or different handles on one device -- race on it:
 * serialise them. */
#define PILOTT_NOALLOC_SCRATCH_BYTES (16u << 20)
int pilott_mem_search_exact_noalloc(pilott_mem_t* h, const float* q_dev, int64_t nq,
                                    const int32_t* query_ids_dev, int64_t nsel, int k, int32_t min_prio,
                                    uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev,
                                    float* out_score_dev, double* out_score64_dev, void* stream);

/* Replaces EnhancedMemory.cleanup's semantic-store pass (enhanced_memory.py:258-272):
 * drops expired rows and rebuilds the ring densely in insertion order.  `total` is
 * unchanged; the i-th oldest survivor is renumbered total - kept + i (the reference
 * renumbers positions the same way when it rebuilds its indexes, :270-272).
 * *kept (host, may be NULL) = rows surviving.
 * keep_host (uint8[count], may be NULL) receives the keep flag of every held row in age
 * order so the host can drop the same MemoryItems.  Synchronous. */
int pilott_mem_compact(pilott_mem_t* h, int64_t now_us, int64_t* kept, uint8_t* keep_host);

/* Read back raw columns of the held rows in age order (tests, snapshots). Any pointer may
 * be NULL.  vec_bits_host is [count, dim] uint16 bf16 bit patterns. Synchronous. */
int pilott_mem_export(pilott_mem_t* h, uint16_t* vec_bits_host, float* inv_norm_host,
                      int32_t* prio_host, uint64_t* tagmask_host, int64_t* expires_us_host);

/* ---- multi-GPU exchange over NVLink peer memory ---------------------------------------
 * The exchange step of the row-sharded store without a collective library: every rank maps
 * every peer's receive buffer (CUDA IPC, one process per GPU on one node), a publish kernel
 * stores the rank's [2][nq][k] (index, fp64 score) block straight into each peer's buffer and
 * releases a per-(slot, rank) flag there, and the merge kernel acquires the flags of its slot
 * before it ranks the world*k candidates of each query.  No reference counterpart (the
 * reference is single-process).
 *   create : allocates [8 slots][world][block_bytes] + flags on `device`; handle_out receives
 *            PILOTT_IPC_HANDLE_BYTES bytes to hand to the other ranks (any transport).
 *   connect: handles = world * PILOTT_IPC_HANDLE_BYTES bytes in rank order (own entry ignored).
 *   publish: block_dev = this rank's block (bytes <= block_bytes, multiple of 16), step >= 0.
 *   publish_copy: the same block and flag, written by the copy engines (cudaMemcpyAsync into each
 *            peer's buffer on a stream the exchange owns, ordered behind `stream`): no kernel, so
 *            no SM is taken from the scan of the next batch (the pair scan fills 72 of a B200's 74
 *            SM pairs, one CTA per SM: a collective's kernel that lands on one of them delays that
 *            pair by the whole collective).  buffer_id (0..7) names the caller's block buffer: call
 *            pilott_xchg_wait_buffer(x, buffer_id, stream) before the next search overwrites it.
 *   merge  : waits for every rank's block of `step`, writes the global top-k like
 *            pilott_topk_merge_packed.  Per rank, up to three publishes may be enqueued ahead of
 *            the oldest outstanding merge (publish(s+1..s+3) before merge(s)), but publish(s+4)
 *            must follow merge(s): that rule is what makes the eight slots enough.  Every rank
 *            must take part in every step.
 *   check  : synchronises the device and reports whether any merge gave up waiting (2 s) for
 *            a peer -- such a merge wrote "no result" rows instead of hanging. */
#define PILOTT_IPC_HANDLE_BYTES 64
typedef struct pilott_xchg pilott_xchg_t;
int pilott_xchg_create(int device, int rank, int world, int64_t block_bytes, pilott_xchg_t** out,
                       void* handle_out);
int pilott_xchg_connect(pilott_xchg_t* x, const void* handles);
int pilott_xchg_publish(pilott_xchg_t* x, const void* block_dev, int64_t bytes, int64_t step, void* stream);
int pilott_xchg_publish_copy(pilott_xchg_t* x, const void* block_dev, int64_t bytes, int64_t step, int buffer_id,
                             void* stream);
int pilott_xchg_wait_buffer(pilott_xchg_t* x, int buffer_id, void* stream);
int pilott_xchg_merge(pilott_xchg_t* x, int64_t step, int64_t nq, int k, int64_t* out_idx_dev,
                      float* out_score_dev, double* out_score64_dev, void* stream);
/* pilott_mem_search + publish in one: finish_kernel stores each query's result row straight into
 * every rank's exchange buffer as it produces it, and the CTA that finishes last releases the
 * step's flags -- no separate publish launch, no staging copy.  Same arguments and outputs as
 * pilott_mem_search (1 <= k <= PILOTT_MAX_FUSED_K, nq >= 1); follow with pilott_xchg_merge(step).
 * Rows of flagged queries are published as they are: a caller that repairs them
 * (pilott_mem_search_refine) uses pilott_mem_search + pilott_xchg_publish instead. */
int pilott_mem_search_publish(pilott_mem_t* h, pilott_xchg_t* x, int64_t step, const float* q_dev, int64_t nq,
                              int k, int32_t min_prio, uint64_t req_tagmask, int64_t now_us,
                              int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev,
                              int32_t* out_inexact_dev, uint8_t* out_flags_dev, void* stream);
/* The whole sharded search in one call, for hosts that do not pipeline: local fused search of this
 * rank's rows (global indices: options 5 / 6 of the store) into block_dev ([2][nq][k] 64-bit words,
 * caller-owned, one per step in flight), publish by copy, wait for every rank's block of `step`,
 * merge into out_idx / out_score / out_score64 [nq, k] -- all enqueued on `stream`, nothing
 * synchronises.  Every rank calls it with the same step, queries and filters; the result is what
 * one store holding every row returns (1 <= k <= PILOTT_MAX_FUSED_K; *out_inexact_dev counts this
 * rank's flagged queries: sum pilott_mem_flagged_total over the ranks before trusting the rows, or
 * repair with pilott_mem_search_refine and the split calls).  This is the role of the
 * pilott_shard_init / pilott_mem_search_sharded pair SURVEY.md section 8(b) planned over
 * ncclComm_t; the exchange it rides on needs no collective library.  No reference counterpart. */
int pilott_mem_search_sharded(pilott_mem_t* h, pilott_xchg_t* x, int64_t step, const float* q_dev, int64_t nq,
                              int k, int32_t min_prio, uint64_t req_tagmask, int64_t now_us,
                              int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev,
                              int32_t* out_inexact_dev, uint8_t* out_flags_dev, void* block_dev, void* stream);
int pilott_xchg_check(pilott_xchg_t* x);
int pilott_xchg_destroy(pilott_xchg_t* x);

/* ---- multi-GPU merge ---------------------------------------------------------------
 * After an all-gather of per-shard results: idx/score64 are [n_shards, nq, k]; picks the
 * global top-k per query by (score desc, idx asc).  No reference counterpart: the reference
 * is single-process; this is the exchange step of the row-sharded
 * store. */
int pilott_topk_merge(const int64_t* idx_dev, const double* score64_dev, int n_shards, int64_t nq,
                      int k, int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev,
                      void* stream);

/* Same merge for ONE gathered buffer: packed is [n_shards][2][nq][k] 64-bit words, plane 0 the
 * int64 indices and plane 1 the fp64 scores of a shard -- what a single all-gather of each
 * rank's (index, score) block produces. */
int pilott_topk_merge_packed(const int64_t* packed_dev, int n_shards, int64_t nq, int k,
                             int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev, void* stream);

/* ---- routing -----------------------------------------------------------------------
 * One call scores nt tasks against na agents and returns, per task, the first agent with
 * the maximal score among the viable ones (-1 = none) -- the inner loops of
 *   TaskRouter._calculate_agent_scores + _attempt_routing   pilott/core/router.py:51-95
 *   LoadBalancer._find_best_agent + _calculate_agent_suitability
 *                                         pilott/orchestration/load_balancer.py:268-336
 *   TaskDelegator._find_best_agent + _calculate_total_score
 *                                         pilott/delegation/task_delegator.py:60-113
 * with BaseAgent.evaluate_task_suitability (pilott/core/agent.py:549-575) folded in.
 * The per-agent fp64 terms are precomputed by the host in the reference's own operation
 * order (see pilottai_b200/routing.py); the device does the task x agent matching and the
 * remaining fp64 additions with one rounding each, so scores are bit-identical to Python.
 *
 * words:            64-bit words per capability / specialization bit set (vocab <= 64*words)
 * task_req [nt,words]   required capabilities          task_tags [nt,words] task tags
 * task_type [nt]        specialization id of task.get('type','') (-1 = not in the vocabulary)
 * task_flags [nt]       bit0: 'required_capabilities' key present, bit1: 'type' key present,
 *                       bit2: some required capability is outside the vocabulary (no agent has it)
 * agent_caps/agent_spec [na,words]; agent_flags [na]: bit0 viable, bit1 has required-caps
 *                       attribute, bit2 has specializations
 * agent_s12 [na,3]      first two score terms for base in {missing-cap, no type match, type match}
 * agent_t4 [na]         last score term
 * spec_u [2][max_tags+1] router spec*0.2, indexed [type match][tag matches]; max_tags >= 64*words
 * mode: 0 router (s12 + u + t4), 1 load balancer / 2 delegator (s12 holds the full score;
 *       mode 2 additionally requires score > 0).
 * out_agent [nt] int32, out_score [nt] fp64 (NaN when no agent). */
int pilott_route_score(const uint64_t* task_req_dev, const uint64_t* task_tags_dev,
                       const int32_t* task_type_dev, const uint8_t* task_flags_dev, int64_t nt,
                       const uint64_t* agent_caps_dev,
                       const uint64_t* agent_spec_dev, const uint8_t* agent_flags_dev,
                       const double* agent_s12_dev, const double* agent_t4_dev, int na, int words,
                       const double* spec_u_dev, int max_tags, int mode, int32_t* out_agent_dev,
                       double* out_score_dev, void* stream);

/* The same scoring with the task x agent matching on the tcgen05 tensor cores (csrc/route_tc.cu):
 * the bit sets are expanded into two bf16 feature matrices inside `workspace_dev`, ONE GEMM yields,
 * per (task, agent), the packed counts {required capabilities held, attribute missing, type match,
 * tag matches} as an exact fp32 integer, and the epilogue that reads TMEM assembles the score and
 * keeps the first maximum.  Same arguments, same outputs, bit-identical results.  Takes
 * words <= 3 and agent tables of up to about 1500 agents (PILOTT_EUNSUPPORTED otherwise: use
 * pilott_route_score); workspace_bytes >= pilott_route_tc_workspace(nt, na, words). */
int64_t pilott_route_tc_workspace(int64_t nt, int na, int words);
int pilott_route_score_tc(const uint64_t* task_req_dev, const uint64_t* task_tags_dev,
                          const int32_t* task_type_dev, const uint8_t* task_flags_dev, int64_t nt,
                          const uint64_t* agent_caps_dev, const uint64_t* agent_spec_dev,
                          const uint8_t* agent_flags_dev, const double* agent_s12_dev,
                          const double* agent_t4_dev, int na, int words, const double* spec_u_dev,
                          int max_tags, int mode, int32_t* out_agent_dev, double* out_score_dev,
                          void* workspace_dev, int64_t workspace_bytes, void* stream);

/* TaskRouter.get_task_priority (pilott/core/router.py:135-145) for nt tasks:
 * 3 critical, 2 high, 1 medium, 0 low. */
int pilott_task_priority(const uint8_t* urgent_dev, const int32_t* complexity_dev,
                         const int32_t* n_deps_dev, int64_t nt, uint8_t* out_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PILOTT_B200_H */
