// abi.cu -- the extern "C" surface declared in include/pilott_b200.h: handle management,
// argument checking, scratch sizing, tensor-map encoding and kernel sequencing.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/pilott_b200.h"
#include "internal.h"

namespace {

constexpr int KP_HOST = 32;   // == pilott::KP in common.cuh
thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
int fail_cuda(cudaError_t e, const char* what) {
    g_err = std::string(what) + ": " + cudaGetErrorName(e) + " (" + cudaGetErrorString(e) + ")";
    return e == cudaErrorMemoryAllocation ? PILOTT_ENOMEM : PILOTT_ECUDA;
}
#define CK(expr)                                                   \
    do {                                                           \
        cudaError_t _e = (expr);                                   \
        if (_e != cudaSuccess) return fail_cuda(_e, #expr);        \
    } while (0)

PFN_cuTensorMapEncodeTiled_v12000 g_encode = nullptr;

int get_encode() {
    if (g_encode) return PILOTT_OK;
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || fn == nullptr)
        return fail(PILOTT_ECUDA, "cuTensorMapEncodeTiled is not available from the driver");
    g_encode = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(fn);
    return PILOTT_OK;
}

// 2-D bf16 tensor [rows, ld] (ld contiguous), box = 64 x box_rows, 128-byte swizzle.
int encode_map(CUtensorMap* map, const void* ptr, int64_t rows, int ld, int box_rows) {
    int rc = get_encode();
    if (rc) return rc;
    cuuint64_t gdim[2] = {(cuuint64_t)ld, (cuuint64_t)(rows > 0 ? rows : 1)};
    cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                          box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(PILOTT_ECUDA, "cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r));
    return PILOTT_OK;
}

struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need) {
        if (need <= bytes) return PILOTT_OK;
        if (p) cudaFree(p);
        p = nullptr; bytes = 0;
        size_t want = need + need / 4;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { p = nullptr; return fail_cuda(e, "cudaMalloc(scratch)"); }
        bytes = want;
        return PILOTT_OK;
    }
    void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
};
struct HostBuf {
    void* p = nullptr;
    size_t bytes = 0;
    int ensure(size_t need) {
        if (need <= bytes) return PILOTT_OK;
        if (p) cudaFreeHost(p);
        p = nullptr; bytes = 0;
        cudaError_t e = cudaMallocHost(&p, need + need / 4);
        if (e != cudaSuccess) { p = nullptr; return fail_cuda(e, "cudaMallocHost(staging)"); }
        bytes = need + need / 4;
        return PILOTT_OK;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; bytes = 0; }
};

bool is_pinned(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

}  // namespace

// Scratch for pilott_mem_search_exact_noalloc: lives in the library's own device image, so it
// exists on every GPU that loads the library from load time until the process exits and no
// call ever has to allocate it.  In host code this name is only a handle the runtime
// understands -- it is NOT a device address and must never be handed to a kernel; the real
// address is looked up per device by noalloc_scratch() below.
static __device__ __align__(256) unsigned char g_noalloc_scratch[PILOTT_NOALLOC_SCRATCH_BYTES];

// One in-flight host-buffer search (pilott_mem_search_host_begin / _end).
struct HostSlot {
    DevBuf q_dev, res_dev;
    HostBuf q_host, res_host;
    cudaEvent_t e_h2d = nullptr, e_done = nullptr, e_d2h = nullptr;
    int64_t nq = 0;
    int k = 0;
    int32_t min_prio = 0;
    uint64_t req = 0;
    int64_t now_us = 0;
    bool busy = false, fused = true;
    int32_t n_brute = 0;       // large k: queries that went to the brute-force kernels
    size_t off_s64 = 0, off_s = 0, off_f = 0, off_c = 0, rtot = 0;
};
constexpr int HOST_SLOTS = 4;

struct pilott_mem {
    int dim = 0, ld = 0, device = 0, num_sms = 0;
    int64_t capacity = 0, cap_pad = 0;
    int64_t total = 0;      // rows ever appended (sequence numbering)
    int64_t count = 0;      // rows held
    int64_t head = 0;       // physical slot of the oldest row
    void* vec = nullptr;
    float* inv_norm = nullptr;
    int32_t* prio = nullptr;
    uint64_t* tagmask = nullptr;
    int64_t* expires = nullptr;
    float* scale = nullptr;
    float* tile_max = nullptr;
    unsigned long long* flagged_dev = nullptr;   // queries flagged by the gap check since the last pilott_mem_flagged_total(reset)
    DevBuf qbuf, partials, gthr, exact_scratch, stage_dev, result_dev, misc_dev, refine, result_ids, largek;
    HostBuf stage_host;
    cudaStream_t own_stream = nullptr;
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;   // copy streams of the pipelined host path
    // Ordering between the two families of entry points, which share the handle's scratch and
    // columns: device-pointer calls enqueue on the CALLER's stream, host-buffer calls on own_stream.
    // ev_dev is recorded behind every device-pointer call, ev_own behind every host-buffer call that
    // returns without synchronising; each family waits for the other's last event before it enqueues.
    cudaEvent_t ev_dev = nullptr, ev_own = nullptr;
    bool dev_pending = false, own_pending = false;
    HostSlot hslot[HOST_SLOTS];
    CUtensorMap map_c;
    bool tc_ok = false;
    int opt_kernel = 0;     // 0 auto, 1 generic, 2 tc 1-CTA, 3 tc pair
    int opt_ctas = 0;
    int opt_dbg = 0;
    int opt_list_len = 0;   // 0 auto, 8, 16, 32
    int opt_finish = 0;     // 0 selective re-score (default), 1 the original kernel that re-scores all 32 survivors
    int auto_min_list = 8;  // floor of the automatic list length: raised when searches keep flagging queries
    int64_t searches_since_refine = 0;
    int64_t seq_mul = 1, seq_add = 0;   // reported index = seq * seq_mul + seq_add (row-sharded stores)
    bool profile = false;   // record CUDA events around the scan kernel
    std::vector<cudaEvent_t> ev;        // pairs: [2i] before, [2i+1] after
    size_t ev_used = 0;
    int64_t launches = 0;
    int64_t generation = 0; // bumped by everything that changes which rows are held (append, clear, compact)
    // The scale / tile_max columns of the last search stay valid while no row changed and the filter
    // is the same (and, if any row may carry an expiry, the same `now`): the pre-pass is skipped then.
    bool expiry_possible = false;           // some append since the last clear passed an expires column with a non-zero (or unseen) entry
    bool sk_valid = false;
    int64_t sk_gen = -1, sk_now = 0;
    int32_t sk_prio = 0;
    uint64_t sk_req = 0;
};

// Exchange buffers of one rank of a row-sharded store (xchg.cu).
struct pilott_xchg {
    int device = 0, rank = 0, world = 1;
    int64_t block_bytes = 0, flags_off = 0, total_bytes = 0;
    unsigned char* mine = nullptr;
    std::vector<unsigned char*> peers;       // peer bases as mapped here (peers[rank] == mine)
    unsigned char** peers_dev = nullptr;
    int32_t* err_dev = nullptr;
    uint32_t* done_dev = nullptr;            // CTA arrival counter of the fused finish + publish
    bool connected = false;
    // publish by copy: a stream of its own for the copy engines, the flag values as a constant
    // table in device memory (entry i holds i + 1: a copy engine reads its source when the copy
    // RUNS, possibly many steps after it was enqueued, so the source must never change), one
    // event per caller buffer
    cudaStream_t copy_stream = nullptr;
    uint64_t* flag_vals = nullptr;
    cudaEvent_t ev_ready = nullptr;
    cudaEvent_t ev_buf[8] = {};
};

namespace {

int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// Called at the top of every device-pointer entry point: work this handle still has in flight on
// its own stream (pilott_mem_search_host_begin) must finish before `st` touches the shared scratch.
int order_device_call(pilott_mem* h, cudaStream_t st) {
    if (h->own_pending) CK(cudaStreamWaitEvent(st, h->ev_own, 0));
    return PILOTT_OK;
}
// ... and at its end: remember where the work went.
int note_device_call(pilott_mem* h, cudaStream_t st) {
    CK(cudaEventRecord(h->ev_dev, st));
    h->dev_pending = true;
    return PILOTT_OK;
}
// Called at the top of every host-buffer entry point: own_stream waits for the last device-pointer
// call (an append whose rows the search must see, a search still using qbuf / scale / partials).
int order_host_call(pilott_mem* h) {
    if (h->dev_pending) CK(cudaStreamWaitEvent(h->own_stream, h->ev_dev, 0));
    h->dev_pending = false;          // own_stream is now ordered behind it, and so is everything enqueued there later
    return PILOTT_OK;
}

// true: scale / tile_max already hold exactly what a pre-pass with these filters would write.
bool scale_is_current(const pilott_mem* h, int32_t min_prio, uint64_t req, int64_t now_us) {
    return h->sk_valid && h->sk_gen == h->generation && h->sk_prio == min_prio && h->sk_req == req &&
           h->sk_now == (h->expiry_possible ? now_us : 0);
}
void scale_note(pilott_mem* h, int32_t min_prio, uint64_t req, int64_t now_us) {
    h->sk_valid = true; h->sk_gen = h->generation; h->sk_prio = min_prio; h->sk_req = req;
    h->sk_now = h->expiry_possible ? now_us : 0;
}

void free_columns(pilott_mem* h) {
    cudaFree(h->vec); cudaFree(h->inv_norm); cudaFree(h->prio); cudaFree(h->tagmask);
    cudaFree(h->expires); cudaFree(h->scale); cudaFree(h->tile_max);
    h->vec = nullptr; h->inv_norm = nullptr; h->prio = nullptr; h->tagmask = nullptr;
    h->expires = nullptr; h->scale = nullptr; h->tile_max = nullptr;
}

int alloc_columns(pilott_mem* h) {
    const size_t n = (size_t)h->capacity;
    CK(cudaMalloc(&h->vec, n * h->ld * 2));
    CK(cudaMalloc(&h->inv_norm, n * 4));
    CK(cudaMalloc(&h->prio, n * 4));
    CK(cudaMalloc(&h->tagmask, n * 8));
    CK(cudaMalloc(&h->expires, n * 8));
    CK(cudaMalloc(&h->scale, (size_t)h->cap_pad * 4));
    CK(cudaMalloc(&h->tile_max, (size_t)(h->cap_pad / 256) * 4));
    CK(cudaMemset(h->vec, 0, n * h->ld * 2));
    CK(cudaMemset(h->inv_norm, 0, n * 4));
    CK(cudaMemset(h->prio, 0, n * 4));
    CK(cudaMemset(h->tagmask, 0, n * 8));
    CK(cudaMemset(h->expires, 0, n * 8));
    h->tc_ok = false;
    if (encode_map(&h->map_c, h->vec, h->capacity, h->ld, 256) == PILOTT_OK) {
        // the pair kernel loads 128-row halves; a second map would only differ in the box, so
        // the box is chosen per launch below (see search_core).
        h->tc_ok = true;
    }
    return PILOTT_OK;
}

// When profiling is on: records the "before" event and hands back the "after" event to record.
int profile_begin(pilott_mem* h, cudaStream_t st, cudaEvent_t* after) {
    *after = nullptr;
    if (!h->profile) return PILOTT_OK;
    if (h->ev_used + 2 > h->ev.size()) {
        for (int i = 0; i < 2; ++i) {
            cudaEvent_t e;
            CK(cudaEventCreate(&e));
            h->ev.push_back(e);
        }
    }
    CK(cudaEventRecord(h->ev[h->ev_used], st));
    *after = h->ev[h->ev_used + 1];
    h->ev_used += 2;
    return PILOTT_OK;
}

// Chooses the scan kernel: 1 generic, 2 tcgen05 1-CTA, 3 tcgen05 CTA pair.
int pick_kernel(const pilott_mem* h) {
    if (h->opt_kernel == 1) return 1;
    if (!h->tc_ok) return 1;
    if (h->opt_kernel == 2 || h->opt_kernel == 3) return h->opt_kernel;
    return 3;      // CTA-pair tcgen05 kernel: the fast path
}

int search_core(pilott_mem* h, const float* q_dev, int64_t nq, int k, int32_t min_prio, uint64_t req,
                int64_t now_us, int64_t* out_idx, float* out_score, double* out_score64,
                int32_t* out_inexact, uint8_t* out_flags, cudaStream_t st, const pilott_xchg* x = nullptr,
                int64_t step = 0) {
    if (nq <= 0) return PILOTT_OK;
    if (k < 1 || k > PILOTT_MAX_FUSED_K) return fail(PILOTT_EINVAL, "k must be in [1, 32] for the fused path");
    const int64_t nq_pad = round_up(nq, 256);
    int rc = h->qbuf.ensure((size_t)nq_pad * h->ld * 2);
    if (rc) return rc;
    // per-search words every scan variant may use: [nq] shared thresholds + [nq][32] first-tile maxima
    const size_t nq32 = (size_t)round_up(nq, 32);           // keeps the [nq][32] block 128-byte aligned
    const size_t n_prog = 1024;                               // progress words of the scan's work units (lock-step window)
    const size_t n_words = nq32 + (size_t)nq * 32 + n_prog;
    rc = h->gthr.ensure(n_words * 4);
    if (rc) return rc;
    // one launch: query conversion + scale pre-pass + zeroing of those words and of the inexact counter
    const bool do_scale = !scale_is_current(h, min_prio, req, now_us);
    CK(pilott::launch_prepare(q_dev, nq, h->dim, h->ld, h->qbuf.p, h->inv_norm, h->prio, h->tagmask, h->expires,
                              h->count, h->capacity, h->cap_pad, min_prio, req, now_us, h->scale, h->tile_max,
                              (uint32_t*)h->gthr.p, (int64_t)n_words, out_inexact, do_scale, st));
    scale_note(h, min_prio, req, now_us);
    h->launches += 1;

    int kern = pick_kernel(h);
    const int ctas = h->opt_ctas > 0 ? h->opt_ctas : h->num_sms;
    pilott::ScanArgs a{};
    a.qb = h->qbuf.p; a.vec = h->vec; a.scale = h->scale; a.tile_max = h->tile_max; a.nq = nq;
    a.nslots = h->count; a.capacity = h->capacity; a.head = h->head; a.ld = h->ld;
    a.dbg = h->opt_dbg;
    CUtensorMap map_q, map_c;
    pilott::FinishArgs f{};
    int lists = 0;
    a.list_len = 32;
    if (kern != 1) {
        const int cg = kern == 3 ? 2 : 1;
        rc = encode_map(&map_q, h->qbuf.p, nq, h->ld, 128);
        if (rc) return rc;
        if (cg == 1) map_c = h->map_c;
        else { rc = encode_map(&map_c, h->vec, h->capacity, h->ld, 128); if (rc) return rc; }
        a.slots = pilott::scan_tc_slots(nq, h->count, cg, ctas, &a.slots_rem);
        // 16-entry lists are enough whenever the lists together hold several times k candidates (the
        // gap check in finish_kernel catches the rare query whose top-k crowd into one list).
        const int nlists = (a.slots + (a.slots_rem > 0 ? 1 : 0)) * pilott::scan_tc_lists_per_slot();
        // With 16+ lists and k <= 10, 8-entry lists do (and free shared memory for a fifth stage).
        const bool lists16 = k <= 12 || nlists >= 8;
        const bool lists8 = cg == 2 && k <= 10 && nlists >= 16;
        if (h->opt_list_len == 8 || h->opt_list_len == 16 || h->opt_list_len == 32) a.list_len = h->opt_list_len;
        else a.list_len = lists8 ? 8 : (lists16 ? 16 : 32);
        // Bursty corpora (runs of near-duplicate rows stored side by side) overflow short lists and
        // send queries to the second-chance pass; after such a batch the floor goes up (refine_core),
        // after 512 searches without one it comes back down a step.
        if (h->opt_list_len == 0 && a.list_len < h->auto_min_list) a.list_len = h->auto_min_list;
        if (++h->searches_since_refine > 512 && h->auto_min_list > 8) { h->auto_min_list /= 2; h->searches_since_refine = 0; }
        lists = (a.slots + (a.slots_rem > 0 ? 1 : 0)) * pilott::scan_tc_lists_per_slot();
        f.rows_per_group = pilott::scan_tc_rows_per_group(cg);
        f.lists_base = a.slots * pilott::scan_tc_lists_per_slot();
        f.groups_rem = a.slots_rem;
        f.lists_extra = pilott::scan_tc_lists_per_slot();
        rc = h->partials.ensure((size_t)nq * lists * KP_HOST * 8);
        if (rc) return rc;
        a.partials = (uint64_t*)h->partials.p;
        a.gthr = (uint32_t*)h->gthr.p;
        a.gmax = a.gthr + nq32;
        // Experiment kept behind PILOTT_LOCKSTEP=n: keep the query groups of a slot within n tiles of each
        // other, so that a corpus tile is read from HBM once on long sweeps (see scan_tc.cu).  Needs every
        // unit resident at once (one wave).
        {
            const int groups_n = (int)((nq + 128 * cg - 1) / (128 * cg));
            const int units_n = groups_n * a.slots;
            static const int lockstep = [] { const char* e = getenv("PILOTT_LOCKSTEP"); return e ? atoi(e) : -1; }();
            const bool want = lockstep > 0;      // off unless PILOTT_LOCKSTEP=n asks for it: measured +1 % at 10 M x 1024, -4 % at 5 M, -9 % at 1 M
            if (want && groups_n > 1 && a.slots_rem == 0 && units_n * cg <= h->num_sms && (size_t)units_n <= n_prog) {
                a.progress = a.gthr + nq32 + (size_t)nq * 32;
                a.window = lockstep > 1 ? lockstep : 4;
            }
        }
        a.krank = k + 6 <= 16 ? 16 : 32;
        f.gthr = a.gthr;
        cudaEvent_t e1 = nullptr;
        rc = profile_begin(h, st, &e1);
        if (rc) return rc;
        CK(pilott::launch_scan_tc(a, &map_q, &map_c, cg, ctas, st));
        if (e1) CK(cudaEventRecord(e1, st));
    } else {
        a.slots = pilott::scan_generic_slots(nq, h->count, h->num_sms);
        lists = a.slots;
        f.rows_per_group = 1; f.lists_base = lists; f.groups_rem = 0; f.lists_extra = 0;
        rc = h->partials.ensure((size_t)nq * lists * KP_HOST * 8);
        if (rc) return rc;
        a.partials = (uint64_t*)h->partials.p;
        cudaEvent_t e1 = nullptr;
        rc = profile_begin(h, st, &e1);
        if (rc) return rc;
        CK(pilott::launch_scan_generic(a, h->num_sms, st));
        if (e1) CK(cudaEventRecord(e1, st));
    }
    f.partials = a.partials; f.slots = lists; f.list_len = a.list_len; f.qb = h->qbuf.p; f.vec = h->vec; f.ld = h->ld;
    f.pool_n = 32;
    f.nq = nq; f.capacity = h->capacity; f.head = h->head; f.base_seq = h->total - h->count;
    f.k = k; f.seq_mul = h->seq_mul; f.seq_add = h->seq_add;
    f.err = h->ld * 2.4e-7 > 1e-4 ? h->ld * 2.4e-7 : 1e-4;
    f.out_idx = out_idx; f.out_score = out_score; f.out_score64 = out_score64;
    f.flags = out_flags; f.inexact = out_inexact;
    if (x) {
        f.peers = x->peers_dev; f.world = x->world; f.rank = x->rank; f.block_bytes = x->block_bytes;
        f.flags_off = x->flags_off; f.step = step; f.done_counter = x->done_dev;
    }
    f.flagged_total = h->flagged_dev;
    f.legacy = h->opt_finish == 1;
    CK(pilott::launch_finish(f, st));
    h->launches += 2;
    return PILOTT_OK;
}

int exact_core(pilott_mem* h, const float* q_dev, int64_t nq, const int32_t* qids, int64_t nsel, int k,
               int32_t min_prio, uint64_t req, int64_t now_us, int64_t* out_idx, float* out_score,
               double* out_score64, cudaStream_t st) {
    if (nq <= 0) return PILOTT_OK;
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (!qids) nsel = nq;
    if (nsel <= 0) return PILOTT_OK;
    int rc = h->qbuf.ensure((size_t)round_up(nq, 256) * h->ld * 2);
    if (rc) return rc;
    CK(pilott::launch_convert_queries(q_dev, nq, h->dim, h->ld, h->qbuf.p, st));
    h->launches += 1;
    const int64_t rows = h->count > 0 ? h->count : 1;
    int64_t chunk = (256ll << 20) / (rows * 8);
    if (chunk < 1) chunk = 1;
    if (chunk > nsel) chunk = nsel;
    if (chunk > 65535) chunk = 65535;
    rc = h->exact_scratch.ensure((size_t)chunk * rows * 8);
    if (rc) return rc;
    std::vector<int32_t> ids;
    const int32_t* ids_dev = qids;
    if (!qids) {
        ids.resize(nsel);
        for (int64_t i = 0; i < nsel; ++i) ids[i] = (int32_t)i;
        rc = h->misc_dev.ensure((size_t)nsel * 4);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->misc_dev.p, ids.data(), (size_t)nsel * 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));      // ids is a stack-lifetime pageable buffer
        ids_dev = (const int32_t*)h->misc_dev.p;
    }
    for (int64_t lo = 0; lo < nsel; lo += chunk) {
        const int64_t n = nsel - lo < chunk ? nsel - lo : chunk;
        CK(pilott::launch_exact(h->qbuf.p, h->vec, h->ld, h->prio, h->tagmask, h->expires, h->count,
                                h->capacity, h->head, h->total - h->count, h->seq_mul, h->seq_add, ids_dev + lo, n, k, min_prio,
                                req, now_us, (double*)h->exact_scratch.p, out_idx, out_score,
                                out_score64, st));
        h->launches += 2;
    }
    return PILOTT_OK;
}

// Second chance for the queries a fused search flagged (see finish.cu): dense block of the
// flagged queries -> COLLECT scan with one fixed threshold per query -> exact re-score and rank
// of the collected rows.  out_idx / out_score64 hold the fused results on entry (their k-th exact
// score is the lower bound the thresholds come from) and the exact results on return.  Queries
// with more than REFINE_CAP rows inside the error margin, and every case this pass does not
// cover (k > 32, no fp64 scores, pair kernel not selectable), go to the brute-force path.
// Synchronises the stream (it has to read the overflow counter).
constexpr int REFINE_CAP = 1024;

int refine_core(pilott_mem* h, const float* q_dev, int64_t nq, const int32_t* qids, int64_t nsel, int k,
                int32_t min_prio, uint64_t req, int64_t now_us, int64_t* out_idx, float* out_score,
                double* out_score64, int32_t* n_brute, cudaStream_t st) {
    if (n_brute) *n_brute = 0;
    if (nq <= 0 || (qids && nsel <= 0)) return PILOTT_OK;
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (!qids || !out_score64 || k > PILOTT_MAX_FUSED_K || pick_kernel(h) != 3 || h->count == 0) {
        if (n_brute) *n_brute = (int32_t)(qids ? nsel : nq);
        return exact_core(h, q_dev, nq, qids, nsel, k, min_prio, req, now_us, out_idx, out_score, out_score64, st);
    }
    // a batch that flags more than ~1.5 % of its queries raises the automatic list length
    if (nsel * 64 > nq && h->auto_min_list < 32) h->auto_min_list *= 2;
    h->searches_since_refine = 0;

    const int64_t nsel_pad = round_up(nsel, 256);
    int rc = h->qbuf.ensure((size_t)round_up(nq, 256) * h->ld * 2);
    if (rc) return rc;
    const size_t qb2_b = (size_t)nsel_pad * h->ld * 2, w_b = (size_t)nsel_pad * 4;
    const size_t cand_b = (size_t)nsel * REFINE_CAP * 4, ovf_b = (size_t)round_up(nsel, 8);
    rc = h->refine.ensure(qb2_b + 2 * w_b + cand_b + ovf_b + 8);
    if (rc) return rc;
    char* base = (char*)h->refine.p;
    void* qb2 = base;
    uint32_t* thr_bits = (uint32_t*)(base + qb2_b);
    uint32_t* cnt = (uint32_t*)(base + qb2_b + w_b);
    uint32_t* cand = (uint32_t*)(base + qb2_b + 2 * w_b);
    uint8_t* ovf = (uint8_t*)(base + qb2_b + 2 * w_b + cand_b);
    int32_t* n_ovf = (int32_t*)(base + qb2_b + 2 * w_b + cand_b + ovf_b);

    // query conversion + scale pre-pass for THESE filters (the scan reads h->scale / h->tile_max)
    CK(pilott::launch_prepare(q_dev, nq, h->dim, h->ld, h->qbuf.p, h->inv_norm, h->prio, h->tagmask, h->expires,
                              h->count, h->capacity, h->cap_pad, min_prio, req, now_us, h->scale, h->tile_max,
                              nullptr, 0, nullptr, !scale_is_current(h, min_prio, req, now_us), st));
    scale_note(h, min_prio, req, now_us);
    const double err = h->ld * 2.4e-7 > 1e-4 ? h->ld * 2.4e-7 : 1e-4;       // same bound as the gap check
    CK(pilott::launch_refine_prepare(h->qbuf.p, qids, nsel, nsel_pad, h->ld, out_idx, out_score64, k, err, qb2,
                                     thr_bits, cnt, n_ovf, st));
    pilott::ScanArgs a{};
    a.qb = qb2; a.vec = h->vec; a.scale = h->scale; a.tile_max = h->tile_max; a.nq = nsel;
    a.nslots = h->count; a.capacity = h->capacity; a.head = h->head; a.ld = h->ld;
    a.partials = (uint64_t*)cand; a.gthr = thr_bits; a.gmax = cnt; a.collect_cap = REFINE_CAP;
    a.list_len = 8; a.krank = 16; a.dbg = 0;
    const int ctas = h->opt_ctas > 0 ? h->opt_ctas : h->num_sms;
    a.slots = pilott::scan_tc_slots(nsel, h->count, 2, ctas, &a.slots_rem);
    CUtensorMap map_q, map_c;
    rc = encode_map(&map_q, qb2, nsel_pad, h->ld, 128);
    if (rc) return rc;
    rc = encode_map(&map_c, h->vec, h->capacity, h->ld, 128);
    if (rc) return rc;
    CK(pilott::launch_scan_tc(a, &map_q, &map_c, 2, ctas, st));
    CK(pilott::launch_refine_finish(cand, cnt, REFINE_CAP, qb2, h->vec, h->ld, qids, nsel, h->capacity, h->head,
                                    h->total - h->count, h->seq_mul, h->seq_add, k, out_idx, out_score, out_score64,
                                    ovf, n_ovf, nullptr, err, nullptr, nullptr, st));
    h->launches += 4;
    int32_t novf = 0;
    CK(cudaMemcpyAsync(&novf, n_ovf, 4, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (novf > 0) {
        // more near-ties than the candidate array holds: those queries go through the brute-force path
        std::vector<uint8_t> fl((size_t)nsel);
        std::vector<int32_t> ids((size_t)nsel), over;
        CK(cudaMemcpyAsync(fl.data(), ovf, (size_t)nsel, cudaMemcpyDeviceToHost, st));
        CK(cudaMemcpyAsync(ids.data(), qids, (size_t)nsel * 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        for (int64_t i = 0; i < nsel; ++i) if (fl[i]) over.push_back(ids[i]);
        rc = h->misc_dev.ensure(over.size() * 4);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->misc_dev.p, over.data(), over.size() * 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        if (n_brute) *n_brute = (int32_t)over.size();
        return exact_core(h, q_dev, nq, (const int32_t*)h->misc_dev.p, (int64_t)over.size(), k, min_prio, req, now_us,
                          out_idx, out_score, out_score64, st);
    }
    return PILOTT_OK;
}

// Large k (PILOTT_MAX_FUSED_K < k <= PILOTT_MAX_LARGE_K): fused top-32 first (it supplies s_4 and
// s_32 of every query), then COLLECT passes whose per-query thresholds are guesses at rank 1.5k --
// from a normal-tail model first, refitted from the row count each pass measures afterwards --
// and finally an open threshold.  refine_finish_kernel re-scores what a pass collected, ranks it
// and checks the guess (the k-th exact score must clear every uncollected row); queries that fail
// the check or collected more than LARGEK_CAP rows go into the next pass, whatever is left after
// the last one goes to brute force.  Synchronises the stream between passes.
constexpr int LARGEK_CAP = 4096;
constexpr int64_t LARGEK_CHUNK = 2048;      // queries per round of passes

int largek_core(pilott_mem* h, const float* q_dev, int64_t nq, int k, int32_t min_prio, uint64_t req,
                int64_t now_us, int64_t* out_idx, float* out_score, double* out_score64, int32_t* n_brute,
                cudaStream_t st) {
    if (n_brute) *n_brute = 0;
    if (nq <= 0) return PILOTT_OK;
    if (k <= PILOTT_MAX_FUSED_K || k > PILOTT_MAX_LARGE_K || pick_kernel(h) != 3 || h->count == 0) {
        if (n_brute) *n_brute = (int32_t)nq;
        return exact_core(h, q_dev, nq, nullptr, nq, k, min_prio, req, now_us, out_idx, out_score, out_score64, st);
    }
    if (nq > LARGEK_CHUNK) {             // 16 KB of candidate slots per query: bound the scratch, batch by batch
        for (int64_t lo = 0; lo < nq; lo += LARGEK_CHUNK) {
            const int64_t n = nq - lo < LARGEK_CHUNK ? nq - lo : LARGEK_CHUNK;
            int32_t nb = 0;
            const int rcc = largek_core(h, q_dev + lo * h->dim, n, k, min_prio, req, now_us, out_idx + lo * k,
                                        out_score + lo * k, out_score64 ? out_score64 + lo * k : nullptr, &nb, st);
            if (rcc) return rcc;
            if (n_brute) *n_brute += nb;
        }
        return PILOTT_OK;
    }
    // 1. fused top-32 into scratch (prepare inside it converts the queries and builds the filter mask)
    const size_t t_idx_b = (size_t)nq * 32 * 8, t_s_b = (size_t)nq * 32 * 4;
    const size_t fl_b = (size_t)round_up(nq, 8);
    int rc = h->largek.ensure(2 * t_idx_b + t_s_b + fl_b + 16 + (size_t)nq * 12);
    if (rc) return rc;
    char* tb = (char*)h->largek.p;
    int64_t* top_idx = (int64_t*)tb;
    double* top_s64 = (double*)(tb + t_idx_b);
    float* top_s = (float*)(tb + 2 * t_idx_b);
    uint8_t* top_flags = (uint8_t*)(tb + 2 * t_idx_b + t_s_b);
    int32_t* top_cnt = (int32_t*)(tb + 2 * t_idx_b + t_s_b + fl_b);
    double* hist_t = (double*)(tb + 2 * t_idx_b + t_s_b + fl_b + 16);       // per query: last threshold tried ...
    uint32_t* hist_c = (uint32_t*)(hist_t + nq);                            // ... and how many rows lay above it
    CK(cudaMemsetAsync(hist_t, 0, (size_t)nq * 12, st));
    rc = search_core(h, q_dev, nq, PILOTT_MAX_FUSED_K, min_prio, req, now_us, top_idx, top_s, top_s64, top_cnt, top_flags, st);
    if (rc) return rc;

    // 2. passes over the queries still open
    const double err = h->ld * 2.4e-7 > 1e-4 ? h->ld * 2.4e-7 : 1e-4;
    std::vector<int32_t> open_ids((size_t)nq), brute;
    for (int64_t i = 0; i < nq; ++i) open_ids[i] = (int32_t)i;
    const double factors[4] = {1.5, 1.5, 2.0, 0.0};       // target ranks as multiples of k; 0 = open threshold
    const int ctas = h->opt_ctas > 0 ? h->opt_ctas : h->num_sms;
    CUtensorMap map_q, map_c;
    rc = encode_map(&map_c, h->vec, h->capacity, h->ld, 128);
    if (rc) return rc;
    for (int pass = 0; pass < 4 && !open_ids.empty(); ++pass) {
        const int64_t nsel = (int64_t)open_ids.size(), nsel_pad = round_up(nsel, 256);
        const size_t qb2_b = (size_t)nsel_pad * h->ld * 2, w_b = (size_t)nsel_pad * 4, d_b = (size_t)nsel_pad * 8;
        const size_t cand_b = (size_t)nsel * LARGEK_CAP * 4, ovf_b = (size_t)round_up(nsel, 8), ids_b = (size_t)round_up(nsel * 4, 8);
        rc = h->refine.ensure(qb2_b + 2 * w_b + d_b + cand_b + ovf_b + ids_b + 8);
        if (rc) return rc;
        char* base = (char*)h->refine.p;
        void* qb2 = base;
        uint32_t* thr_bits = (uint32_t*)(base + qb2_b);
        uint32_t* cnt = (uint32_t*)(base + qb2_b + w_b);
        double* tcos = (double*)(base + qb2_b + 2 * w_b);
        uint32_t* cand = (uint32_t*)(base + qb2_b + 2 * w_b + d_b);
        uint8_t* status = (uint8_t*)(base + qb2_b + 2 * w_b + d_b + cand_b);
        int32_t* ids_dev = (int32_t*)(base + qb2_b + 2 * w_b + d_b + cand_b + ovf_b);
        int32_t* n_open = (int32_t*)(base + qb2_b + 2 * w_b + d_b + cand_b + ovf_b + ids_b);
        CK(cudaMemcpyAsync(ids_dev, open_ids.data(), (size_t)nsel * 4, cudaMemcpyHostToDevice, st));
        CK(pilott::launch_largek_prepare(h->qbuf.p, ids_dev, nsel, nsel_pad, h->ld, top_idx, top_s64, k, factors[pass],
                                         8.0 * err * (pass + 1), (double)h->count, hist_t, hist_c, (double)LARGEK_CAP, qb2, thr_bits, tcos,
                                         cnt, n_open, st));
        pilott::ScanArgs a{};
        a.qb = qb2; a.vec = h->vec; a.scale = h->scale; a.tile_max = h->tile_max; a.nq = nsel;
        a.nslots = h->count; a.capacity = h->capacity; a.head = h->head; a.ld = h->ld;
        a.partials = (uint64_t*)cand; a.gthr = thr_bits; a.gmax = cnt; a.collect_cap = LARGEK_CAP;
        a.list_len = 8; a.krank = 16; a.dbg = 0;
        a.slots = pilott::scan_tc_slots(nsel, h->count, 2, ctas, &a.slots_rem);
        rc = encode_map(&map_q, qb2, nsel_pad, h->ld, 128);
        if (rc) return rc;
        CK(pilott::launch_scan_tc(a, &map_q, &map_c, 2, ctas, st));
        CK(pilott::launch_refine_finish(cand, cnt, LARGEK_CAP, qb2, h->vec, h->ld, ids_dev, nsel, h->capacity, h->head,
                                        h->total - h->count, h->seq_mul, h->seq_add, k, out_idx, out_score, out_score64,
                                        status, n_open, tcos, err, hist_t, hist_c, st));
        h->launches += 3;
        int32_t nopen = 0;
        CK(cudaMemcpyAsync(&nopen, n_open, 4, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (nopen == 0) { open_ids.clear(); break; }
        std::vector<uint8_t> stt((size_t)nsel);
        CK(cudaMemcpyAsync(stt.data(), status, (size_t)nsel, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        std::vector<int32_t> next;
        if (getenv("PILOTT_DEBUG_LARGEK")) {
            std::vector<uint32_t> cc((size_t)nsel);
            std::vector<double> tt((size_t)nsel);
            cudaMemcpy(cc.data(), cnt, (size_t)nsel * 4, cudaMemcpyDeviceToHost);
            cudaMemcpy(tt.data(), tcos, (size_t)nsel * 8, cudaMemcpyDeviceToHost);
            int shown = 0;
            for (int64_t i = 0; i < nsel && shown < 6; ++i)
                if (stt[i]) { fprintf(stderr, "[largek pass %d] query %d status %d rows %u threshold %.6f\n", pass, open_ids[i], (int)stt[i], cc[i], tt[i]); ++shown; }
            fprintf(stderr, "[largek pass %d] %d of %lld still open\n", pass, nopen, (long long)nsel);
        }
        for (int64_t i = 0; i < nsel; ++i) {
            if (stt[i] != 0 && pass < 3) next.push_back(open_ids[i]);          // guess too high or too low: refit, retry
            else if (stt[i] != 0) brute.push_back(open_ids[i]);                // out of passes
        }
        open_ids.swap(next);
    }
    if (!brute.empty()) {
        rc = h->misc_dev.ensure(brute.size() * 4);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->misc_dev.p, brute.data(), brute.size() * 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        if (n_brute) *n_brute = (int32_t)brute.size();
        return exact_core(h, q_dev, nq, (const int32_t*)h->misc_dev.p, (int64_t)brute.size(), k, min_prio, req, now_us,
                          out_idx, out_score, out_score64, st);
    }
    return PILOTT_OK;
}

// Device address of g_noalloc_scratch on `device` (which must be the current device).  The
// lookup runs once per device; later calls return the cached pointer.
constexpr int kNoallocMaxDevices = 64;
std::mutex g_noalloc_mu;
void* g_noalloc_ptr[kNoallocMaxDevices] = {};

int noalloc_scratch(int device, unsigned char** out) {
    if (device < 0 || device >= kNoallocMaxDevices) return fail(PILOTT_EINVAL, "device index out of range");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess) return fail_cuda(e, "cudaGetSymbolAddress(g_noalloc_scratch)");
        g_noalloc_ptr[device] = ptr;
    }
    *out = static_cast<unsigned char*>(g_noalloc_ptr[device]);
    return PILOTT_OK;
}
}  // namespace

extern "C" int pilott_kp(void) { return 32; }

extern "C" {

const char* pilott_last_error(void) { return g_err.c_str(); }
int pilott_version(void) { return 100; }

int pilott_mem_create(int dim, int64_t capacity, int device, pilott_mem_t** out) {
    if (!out) return fail(PILOTT_EINVAL, "out is NULL");
    *out = nullptr;
    if (dim < 1 || dim > 16384) return fail(PILOTT_EINVAL, "dim must be in [1, 16384]");
    if (capacity < 1 || capacity > 0x7fffff00ll) return fail(PILOTT_EINVAL, "capacity must be in [1, 2^31)");
    int ndev = 0;
    CK(cudaGetDeviceCount(&ndev));
    if (device < 0 || device >= ndev) return fail(PILOTT_EINVAL, "no such CUDA device");
    DeviceGuard g(device);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(PILOTT_EUNSUPPORTED, "pilott_b200 kernels are built for sm_100a only");
    pilott_mem* h = new (std::nothrow) pilott_mem();
    if (!h) return fail(PILOTT_ENOMEM, "host allocation failed");
    h->dim = dim; h->ld = (int)round_up(dim, 64); h->device = device; h->num_sms = prop.multiProcessorCount;
    h->capacity = capacity; h->cap_pad = round_up(capacity, 256) + 256;
    int rc = alloc_columns(h);
    if (rc == PILOTT_OK) {
        cudaError_t e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
        if (e != cudaSuccess) rc = fail_cuda(e, "cudaStreamCreate");
    }
    if (rc == PILOTT_OK) {
        cudaError_t e = cudaMalloc(&h->flagged_dev, 8);
        if (e == cudaSuccess) e = cudaMemset(h->flagged_dev, 0, 8);
        if (e != cudaSuccess) rc = fail_cuda(e, "cudaMalloc(flagged counter)");
    }
    if (rc == PILOTT_OK) {
        cudaError_t e = cudaEventCreateWithFlags(&h->ev_dev, cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&h->ev_own, cudaEventDisableTiming);
        if (e != cudaSuccess) rc = fail_cuda(e, "cudaEventCreate");
    }
    if (rc == PILOTT_OK && h->tc_ok) {
        cudaError_t e = pilott::scan_tc_init();
        if (e != cudaSuccess) { h->tc_ok = false; cudaGetLastError(); }
    }
    if (rc == PILOTT_OK) {
        // Set-up of pilott_mem_search_exact_noalloc for this device, done HERE so that the entry
        // point itself never triggers it: with CUDA's lazy module loading the library's device
        // image -- and the 16 MiB backing g_noalloc_scratch -- is only loaded on a device when one of
        // its symbols is first used there, and that load allocates.  create() allocates anyway.
        unsigned char* unused = nullptr;
        rc = noalloc_scratch(device, &unused);
    }
    if (rc != PILOTT_OK) {
        free_columns(h);
        cudaFree(h->flagged_dev);
        if (h->own_stream) cudaStreamDestroy(h->own_stream);
        if (h->ev_dev) cudaEventDestroy(h->ev_dev);
        if (h->ev_own) cudaEventDestroy(h->ev_own);
        delete h;
        return rc;
    }
    CK(cudaDeviceSynchronize());
    if (const char* f = getenv("PILOTT_FINISH_FORM")) h->opt_finish = atoi(f) == 1 ? 1 : 0;   // developer switch (option 9)
    *out = h;
    return PILOTT_OK;
}

int pilott_mem_destroy(pilott_mem_t* h) {
    if (!h) return PILOTT_OK;
    DeviceGuard g(h->device);
    cudaDeviceSynchronize();
    free_columns(h);
    cudaFree(h->flagged_dev);
    h->qbuf.release(); h->partials.release(); h->gthr.release(); h->exact_scratch.release(); h->stage_dev.release();
    h->result_dev.release(); h->misc_dev.release(); h->refine.release(); h->result_ids.release(); h->largek.release();
    h->stage_host.release();
    for (cudaEvent_t e : h->ev) cudaEventDestroy(e);
    for (auto& sl : h->hslot) {
        sl.q_dev.release(); sl.res_dev.release(); sl.q_host.release(); sl.res_host.release();
        if (sl.e_h2d) cudaEventDestroy(sl.e_h2d);
        if (sl.e_done) cudaEventDestroy(sl.e_done);
        if (sl.e_d2h) cudaEventDestroy(sl.e_d2h);
    }
    if (h->h2d_stream) cudaStreamDestroy(h->h2d_stream);
    if (h->d2h_stream) cudaStreamDestroy(h->d2h_stream);
    if (h->own_stream) cudaStreamDestroy(h->own_stream);
    if (h->ev_dev) cudaEventDestroy(h->ev_dev);
    if (h->ev_own) cudaEventDestroy(h->ev_own);
    delete h;
    return PILOTT_OK;
}

int pilott_mem_clear(pilott_mem_t* h, void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    (void)stream;
    h->total = 0; h->count = 0; h->head = 0;
    h->expiry_possible = false;
    ++h->generation;
    return PILOTT_OK;
}

int pilott_mem_info(const pilott_mem_t* h, int64_t* total, int64_t* count, int64_t* capacity, int* dim,
                    int* row_stride) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (total) *total = h->total;
    if (count) *count = h->count;
    if (capacity) *capacity = h->capacity;
    if (dim) *dim = h->dim;
    if (row_stride) *row_stride = h->ld;
    return PILOTT_OK;
}

int pilott_mem_set_option(pilott_mem_t* h, int which, int64_t value) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (which == 0) {
        if (value < 0 || value > 3) return fail(PILOTT_EINVAL, "scan kernel must be 0..3");
        if (value >= 2 && !h->tc_ok) return fail(PILOTT_EUNSUPPORTED, "tensor-map encoding unavailable: tcgen05 path disabled");
        h->opt_kernel = (int)value;
    } else if (which == 1) {
        if (value < 0 || value > 4096) return fail(PILOTT_EINVAL, "CTA count out of range");
        h->opt_ctas = (int)value;
    } else if (which == 7) {
        if (value < h->count) return fail(PILOTT_EINVAL, "total cannot be smaller than the rows held");
        h->total = value;        // restoring a snapshot: continue the sequence numbering where it stopped
    } else if (which == 4) {
        h->profile = value != 0;
    } else if (which == 5) {
        if (value < 1) return fail(PILOTT_EINVAL, "sequence stride must be >= 1");
        h->seq_mul = value;
    } else if (which == 6) {
        if (value < 0) return fail(PILOTT_EINVAL, "sequence offset must be >= 0");
        h->seq_add = value;
    } else if (which == 3) {
        if (value != 0 && value != 8 && value != 16 && value != 32) return fail(PILOTT_EINVAL, "list length must be 0, 8, 16 or 32");
        h->opt_list_len = (int)value;
    } else if (which == 9) {
        if (value != 0 && value != 1) return fail(PILOTT_EINVAL, "finish form must be 0 or 1");
        h->opt_finish = (int)value;
    } else if (which == 2) {
#ifdef PILOTT_DEBUG
        h->opt_dbg = (int)value;      // developer switches (profiling experiments); 0 = off
#else
        if (value != 0)
            return fail(PILOTT_EUNSUPPORTED, "developer switches exist only in -DPILOTT_DEBUG builds of the library");
        h->opt_dbg = 0;
#endif
    } else {
        return fail(PILOTT_EINVAL, "unknown option");
    }
    return PILOTT_OK;
}

int64_t pilott_mem_launch_count(const pilott_mem_t* h) { return h ? h->launches : 0; }

int pilott_mem_flagged_total(pilott_mem_t* h, int64_t* total, int reset) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    DeviceGuard g(h->device);
    CK(cudaDeviceSynchronize());
    unsigned long long v = 0;
    CK(cudaMemcpy(&v, h->flagged_dev, 8, cudaMemcpyDeviceToHost));
    if (reset) CK(cudaMemset(h->flagged_dev, 0, 8));
    if (total) *total = (int64_t)v;
    return PILOTT_OK;
}

int pilott_mem_profile_read(pilott_mem_t* h, double* scan_ms_total, int64_t* n_scans) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    DeviceGuard g(h->device);
    double tot = 0.0;
    for (size_t i = 0; i + 1 < h->ev_used; i += 2) {
        CK(cudaEventSynchronize(h->ev[i + 1]));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, h->ev[i], h->ev[i + 1]));
        tot += ms;
    }
    if (scan_ms_total) *scan_ms_total = tot;
    if (n_scans) *n_scans = (int64_t)(h->ev_used / 2);
    h->ev_used = 0;
    return PILOTT_OK;
}

static int append_core(pilott_mem_t* h, const float* x_dev, int64_t n, const int32_t* prio_dev,
                       const uint64_t* tagmask_dev, const int64_t* expires_us_dev, int64_t* first_seq,
                       void* stream) {
    if (first_seq) *first_seq = h->total;
    if (expires_us_dev) h->expiry_possible = true;      // device column: its values are not visible here
    int64_t first_slot = h->head + h->count;
    if (first_slot >= h->capacity) first_slot -= h->capacity;
    CK(pilott::launch_append(x_dev, n, h->dim, h->ld, h->vec, h->inv_norm, h->prio, h->tagmask, h->expires,
                             prio_dev, tagmask_dev, expires_us_dev, first_slot, h->capacity,
                             (cudaStream_t)stream));
    h->launches += 1;
    // ring bookkeeping: deque(maxlen) drops the oldest rows (enhanced_memory.py:27,86)
    const int64_t over = h->count + n - h->capacity;
    if (over > 0) { h->head = (h->head + over) % h->capacity; h->count = h->capacity; }
    else h->count += n;
    h->total += n;
    ++h->generation;
    return PILOTT_OK;
}

int pilott_mem_append(pilott_mem_t* h, const float* x_dev, int64_t n, const int32_t* prio_dev,
                      const uint64_t* tagmask_dev, const int64_t* expires_us_dev, int64_t* first_seq,
                      void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (n < 0 || n > h->capacity) return fail(PILOTT_EINVAL, "append batch must hold between 0 and capacity rows");
    if (n == 0) { if (first_seq) *first_seq = h->total; return PILOTT_OK; }
    if (!x_dev) return fail(PILOTT_EINVAL, "x is NULL");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = append_core(h, x_dev, n, prio_dev, tagmask_dev, expires_us_dev, first_seq, stream);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_mem_append_host(pilott_mem_t* h, const float* x_host, int64_t n, const int32_t* prio_host,
                           const uint64_t* tagmask_host, const int64_t* expires_us_host,
                           int64_t* first_seq) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (n < 0 || n > h->capacity) return fail(PILOTT_EINVAL, "append batch must hold between 0 and capacity rows");
    if (n == 0) { if (first_seq) *first_seq = h->total; return PILOTT_OK; }
    if (!x_host) return fail(PILOTT_EINVAL, "x is NULL");
    DeviceGuard g(h->device);
    const size_t xb = (size_t)n * h->dim * 4, pb = (size_t)n * 4, tb = (size_t)n * 8, eb = (size_t)n * 8;
    const size_t off_p = round_up(xb, 256), off_t = off_p + round_up(pb, 256), off_e = off_t + round_up(tb, 256);
    const size_t tot = off_e + round_up(eb, 256);
    int rc = h->stage_host.ensure(tot);
    if (rc) return rc;
    rc = h->stage_dev.ensure(tot);
    if (rc) return rc;
    char* hs = (char*)h->stage_host.p;
    char* ds = (char*)h->stage_dev.p;
    memcpy(hs, x_host, xb);
    if (prio_host) memcpy(hs + off_p, prio_host, pb);
    if (tagmask_host) memcpy(hs + off_t, tagmask_host, tb);
    bool any_exp = false;              // an all-zero column ("never expires") is the same as no column
    if (expires_us_host) {
        for (int64_t i = 0; i < n && !any_exp; ++i) any_exp = expires_us_host[i] != 0;
        if (any_exp) memcpy(hs + off_e, expires_us_host, eb);
    }
    cudaStream_t st = h->own_stream;
    rc = order_host_call(h);
    if (rc) return rc;
    CK(cudaMemcpyAsync(ds, hs, tot, cudaMemcpyHostToDevice, st));
    rc = append_core(h, (const float*)ds, n, prio_host ? (const int32_t*)(ds + off_p) : nullptr,
                           tagmask_host ? (const uint64_t*)(ds + off_t) : nullptr,
                           any_exp ? (const int64_t*)(ds + off_e) : nullptr, first_seq, st);
    if (rc) return rc;
    CK(cudaStreamSynchronize(st));
    h->own_pending = false;
    return PILOTT_OK;
}

int pilott_mem_search(pilott_mem_t* h, const float* q_dev, int64_t nq, int k, int32_t min_prio,
                      uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev, float* out_score_dev,
                      double* out_score64_dev, int32_t* out_inexact_dev, uint8_t* out_flags_dev,
                      void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (nq > 0 && (!q_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = search_core(h, q_dev, nq, k, min_prio, req_tagmask, now_us, out_idx_dev, out_score_dev,
                     out_score64_dev, out_inexact_dev, out_flags_dev, (cudaStream_t)stream);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_mem_search_exact(pilott_mem_t* h, const float* q_dev, int64_t nq, const int32_t* query_ids_dev,
                            int64_t nsel, int k, int32_t min_prio, uint64_t req_tagmask, int64_t now_us,
                            int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev,
                            void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (nq > 0 && (!q_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = exact_core(h, q_dev, nq, query_ids_dev, nsel, k, min_prio, req_tagmask, now_us, out_idx_dev,
                    out_score_dev, out_score64_dev, (cudaStream_t)stream);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_mem_search_refine(pilott_mem_t* h, const float* q_dev, int64_t nq, const int32_t* query_ids_dev,
                             int64_t nsel, int k, int32_t min_prio, uint64_t req_tagmask, int64_t now_us,
                             int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev,
                             int32_t* n_bruteforce, void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (nq > 0 && (!q_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = refine_core(h, q_dev, nq, query_ids_dev, nsel, k, min_prio, req_tagmask, now_us, out_idx_dev,
                     out_score_dev, out_score64_dev, n_bruteforce, (cudaStream_t)stream);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_mem_search_large(pilott_mem_t* h, const float* q_dev, int64_t nq, int k, int32_t min_prio,
                            uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev, float* out_score_dev,
                            double* out_score64_dev, int32_t* n_bruteforce, void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (nq > 0 && (!q_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = largek_core(h, q_dev, nq, k, min_prio, req_tagmask, now_us, out_idx_dev, out_score_dev, out_score64_dev,
                     n_bruteforce, (cudaStream_t)stream);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_mem_search_host(pilott_mem_t* h, const float* q_host, int64_t nq, int k, int32_t min_prio,
                           uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_host,
                           float* out_score_host, double* out_score64_host, int32_t* n_fallback) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (n_fallback) *n_fallback = 0;
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (nq == 0) return PILOTT_OK;
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (!q_host || !out_idx_host || !out_score_host) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    cudaStream_t st = h->own_stream;
    const size_t qb = (size_t)nq * h->dim * 4;
    // result block: idx | score64 | score | flags | inexact
    const size_t ib = (size_t)nq * k * 8, s64b = ib, sb = (size_t)nq * k * 4, fb = round_up(nq, 8);
    const size_t off_s64 = ib, off_s = off_s64 + s64b, off_f = off_s + round_up(sb, 8), off_c = off_f + fb;
    const size_t rtot = off_c + 8;
    int rc = h->stage_host.ensure(qb > rtot ? qb : rtot);
    if (rc) return rc;
    rc = h->stage_dev.ensure(qb);
    if (rc) return rc;
    rc = h->result_dev.ensure(rtot);
    if (rc) return rc;
    rc = order_host_call(h);
    if (rc) return rc;
    if (is_pinned(q_host)) {          // caller's buffer is page-locked: DMA straight from it
        CK(cudaMemcpyAsync(h->stage_dev.p, q_host, qb, cudaMemcpyHostToDevice, st));
    } else {
        memcpy(h->stage_host.p, q_host, qb);
        CK(cudaMemcpyAsync(h->stage_dev.p, h->stage_host.p, qb, cudaMemcpyHostToDevice, st));
    }
    char* rd = (char*)h->result_dev.p;
    const float* qd = (const float*)h->stage_dev.p;
    int64_t* d_idx = (int64_t*)rd;
    double* d_s64 = (double*)(rd + off_s64);
    float* d_s = (float*)(rd + off_s);
    uint8_t* d_f = (uint8_t*)(rd + off_f);
    int32_t* d_c = (int32_t*)(rd + off_c);
    const bool fused = k <= PILOTT_MAX_FUSED_K;
    if (fused) {
        rc = search_core(h, qd, nq, k, min_prio, req_tagmask, now_us, d_idx, d_s, d_s64, d_c, d_f, st);
    } else {
        int32_t nb = 0;
        rc = largek_core(h, qd, nq, k, min_prio, req_tagmask, now_us, d_idx, d_s, d_s64, &nb, st);
        if (n_fallback) *n_fallback = nb;      // queries that needed the brute-force kernels
    }
    if (rc) return rc;
    char* rh = (char*)h->stage_host.p;
    CK(cudaMemcpyAsync(rh, rd, rtot, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (fused && *(int32_t*)(rh + off_c) > 0) {
        // second-chance pass (brute force behind it) for the queries whose gap check failed
        std::vector<int32_t> ids;
        for (int64_t i = 0; i < nq; ++i) if (((uint8_t*)(rh + off_f))[i]) ids.push_back((int32_t)i);
        // (own buffer: refine_core uses misc_dev for the ids it hands on to the brute-force path)
        rc = h->result_ids.ensure(ids.size() * 4);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->result_ids.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        rc = refine_core(h, qd, nq, (const int32_t*)h->result_ids.p, (int64_t)ids.size(), k, min_prio, req_tagmask,
                         now_us, d_idx, d_s, d_s64, nullptr, st);
        if (rc) return rc;
        CK(cudaMemcpyAsync(rh, rd, rtot, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (n_fallback) *n_fallback = (int32_t)ids.size();
    }
    memcpy(out_idx_host, rh, ib);
    memcpy(out_score_host, rh + off_s, sb);
    if (out_score64_host) memcpy(out_score64_host, rh + off_s64, s64b);
    h->own_pending = false;           // own_stream was synchronised above
    return PILOTT_OK;
}

int pilott_mem_search_host_begin(pilott_mem_t* h, int slot, const float* q_host, int64_t nq, int k, int32_t min_prio,
                                 uint64_t req_tagmask, int64_t now_us) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (slot < 0 || slot >= HOST_SLOTS) return fail(PILOTT_EINVAL, "slot must be in [0, 4)");
    if (nq < 1) return fail(PILOTT_EINVAL, "nq must be >= 1");
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (!q_host) return fail(PILOTT_EINVAL, "NULL query pointer");
    HostSlot& sl = h->hslot[slot];
    if (sl.busy) return fail(PILOTT_EINVAL, "slot is in flight: call pilott_mem_search_host_end first");
    DeviceGuard g(h->device);
    if (!h->h2d_stream) CK(cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking));
    if (!h->d2h_stream) CK(cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
    if (!sl.e_h2d) {
        CK(cudaEventCreateWithFlags(&sl.e_h2d, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.e_done, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&sl.e_d2h, cudaEventDisableTiming));
    }
    const size_t qb = (size_t)nq * h->dim * 4;
    const size_t ib = (size_t)nq * k * 8, sb = (size_t)nq * k * 4, fb = (size_t)round_up(nq, 8);
    sl.off_s64 = ib; sl.off_s = 2 * ib; sl.off_f = sl.off_s + (size_t)round_up((int64_t)sb, 8); sl.off_c = sl.off_f + fb;
    sl.rtot = sl.off_c + 8;
    int rc = sl.q_dev.ensure(qb);
    if (rc) return rc;
    rc = sl.res_dev.ensure(sl.rtot);
    if (rc) return rc;
    rc = sl.res_host.ensure(sl.rtot);
    if (rc) return rc;
    const void* src = q_host;
    if (!is_pinned(q_host)) {                       // pageable: stage through the slot's pinned buffer
        rc = sl.q_host.ensure(qb);
        if (rc) return rc;
        memcpy(sl.q_host.p, q_host, qb);
        src = sl.q_host.p;
    }
    CK(cudaMemcpyAsync(sl.q_dev.p, src, qb, cudaMemcpyHostToDevice, h->h2d_stream));
    CK(cudaEventRecord(sl.e_h2d, h->h2d_stream));
    cudaStream_t st = h->own_stream;
    rc = order_host_call(h);
    if (rc) return rc;
    CK(cudaStreamWaitEvent(st, sl.e_h2d, 0));
    char* rd = (char*)sl.res_dev.p;
    sl.nq = nq; sl.k = k; sl.min_prio = min_prio; sl.req = req_tagmask; sl.now_us = now_us;
    sl.fused = k <= PILOTT_MAX_FUSED_K;
    if (sl.fused)
        rc = search_core(h, (const float*)sl.q_dev.p, nq, k, min_prio, req_tagmask, now_us, (int64_t*)rd,
                         (float*)(rd + sl.off_s), (double*)(rd + sl.off_s64), (int32_t*)(rd + sl.off_c),
                         (uint8_t*)(rd + sl.off_f), st);
    else
        rc = largek_core(h, (const float*)sl.q_dev.p, nq, k, min_prio, req_tagmask, now_us, (int64_t*)rd,
                         (float*)(rd + sl.off_s), (double*)(rd + sl.off_s64), &sl.n_brute, st);
    if (rc) return rc;
    CK(cudaEventRecord(sl.e_done, st));
    CK(cudaEventRecord(h->ev_own, st));
    h->own_pending = true;            // a device-pointer call that follows waits for this search
    CK(cudaStreamWaitEvent(h->d2h_stream, sl.e_done, 0));
    CK(cudaMemcpyAsync(sl.res_host.p, rd, sl.rtot, cudaMemcpyDeviceToHost, h->d2h_stream));
    CK(cudaEventRecord(sl.e_d2h, h->d2h_stream));
    sl.busy = true;
    return PILOTT_OK;
}

int pilott_mem_search_host_end(pilott_mem_t* h, int slot, int64_t* out_idx_host, float* out_score_host,
                               double* out_score64_host, int32_t* n_fallback) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (slot < 0 || slot >= HOST_SLOTS) return fail(PILOTT_EINVAL, "slot must be in [0, 4)");
    if (n_fallback) *n_fallback = 0;
    HostSlot& sl = h->hslot[slot];
    if (!sl.busy) return fail(PILOTT_EINVAL, "slot is not in flight");
    if (!out_idx_host || !out_score_host) return fail(PILOTT_EINVAL, "NULL output pointer");
    DeviceGuard g(h->device);
    sl.busy = false;
    CK(cudaEventSynchronize(sl.e_d2h));
    char* rh = (char*)sl.res_host.p;
    char* rd = (char*)sl.res_dev.p;
    const int64_t nq = sl.nq;
    const int k = sl.k;
    if (!sl.fused) {
        if (n_fallback) *n_fallback = sl.n_brute;
    } else if (*(int32_t*)(rh + sl.off_c) > 0) {
        // second-chance pass for the flagged queries (brute force behind it), then the block again
        cudaStream_t st = h->own_stream;
        std::vector<int32_t> ids;
        for (int64_t i = 0; i < nq; ++i) if (((uint8_t*)(rh + sl.off_f))[i]) ids.push_back((int32_t)i);
        int rc = h->result_ids.ensure(ids.size() * 4);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->result_ids.p, ids.data(), ids.size() * 4, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));
        rc = refine_core(h, (const float*)sl.q_dev.p, nq, (const int32_t*)h->result_ids.p, (int64_t)ids.size(), k,
                         sl.min_prio, sl.req, sl.now_us, (int64_t*)rd, (float*)(rd + sl.off_s),
                         (double*)(rd + sl.off_s64), nullptr, st);
        if (rc) return rc;
        CK(cudaMemcpyAsync(rh, rd, sl.rtot, cudaMemcpyDeviceToHost, st));
        CK(cudaStreamSynchronize(st));
        if (n_fallback) *n_fallback = (int32_t)ids.size();
    }
    memcpy(out_idx_host, rh, (size_t)nq * k * 8);
    memcpy(out_score_host, rh + sl.off_s, (size_t)nq * k * 4);
    if (out_score64_host) memcpy(out_score64_host, rh + sl.off_s64, (size_t)nq * k * 8);
    return PILOTT_OK;
}

int pilott_mem_compact(pilott_mem_t* h, int64_t now_us, int64_t* kept, uint8_t* keep_host) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    DeviceGuard g(h->device);
    cudaStream_t st = h->own_stream;
    const int64_t n = h->count;
    if (kept) *kept = n;
    if (n == 0) return PILOTT_OK;
    CK(cudaDeviceSynchronize());
    int rc = h->misc_dev.ensure((size_t)n);
    if (rc) return rc;
    CK(pilott::launch_compact_flags(h->expires, n, h->head, h->capacity, now_us, (uint8_t*)h->misc_dev.p, st));
    std::vector<uint8_t> keep((size_t)n);
    CK(cudaMemcpyAsync(keep.data(), h->misc_dev.p, (size_t)n, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    std::vector<int64_t> src;
    src.reserve((size_t)n);
    for (int64_t i = 0; i < n; ++i) if (keep[i]) src.push_back((h->head + i) % h->capacity);
    if (keep_host) memcpy(keep_host, keep.data(), (size_t)n);
    const int64_t m = (int64_t)src.size();
    if (kept) *kept = m;
    h->launches += 1;
    if (m == n) return PILOTT_OK;
    // Gather the survivors into fresh columns, then swap.  (Needs room for a second copy of the
    // columns while it runs; a store that fills more than half of HBM cannot be compacted this way.)
    // Everything that can fail is done before the handle is touched, and any failure after the
    // new columns exist frees them and puts the old ones back, so an error leaves the store as it was.
    if (m > 0) {
        rc = h->stage_dev.ensure((size_t)m * 8);
        if (rc) return rc;
        CK(cudaMemcpyAsync(h->stage_dev.p, src.data(), (size_t)m * 8, cudaMemcpyHostToDevice, st));
        CK(cudaStreamSynchronize(st));          // src is a host vector
    }
    struct Columns {
        void* vec; float* inv_norm; int32_t* prio; uint64_t* tagmask; int64_t* expires; float* scale; float* tile_max;
        CUtensorMap map_c; bool tc_ok;
    };
    const Columns old{h->vec, h->inv_norm, h->prio, h->tagmask, h->expires, h->scale, h->tile_max, h->map_c, h->tc_ok};
    auto restore = [&]() {
        free_columns(h);
        h->vec = old.vec; h->inv_norm = old.inv_norm; h->prio = old.prio; h->tagmask = old.tagmask;
        h->expires = old.expires; h->scale = old.scale; h->tile_max = old.tile_max; h->map_c = old.map_c; h->tc_ok = old.tc_ok;
    };
    h->vec = nullptr; h->inv_norm = nullptr; h->prio = nullptr; h->tagmask = nullptr;
    h->expires = nullptr; h->scale = nullptr; h->tile_max = nullptr;
    rc = alloc_columns(h);
    if (rc) { restore(); return rc; }
    if (m > 0) {
        cudaError_t e = pilott::launch_gather_rows(old.vec, old.inv_norm, old.prio, old.tagmask, old.expires, h->ld,
                                                   (const int64_t*)h->stage_dev.p, m, h->vec, h->inv_norm, h->prio,
                                                   h->tagmask, h->expires, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { restore(); return fail_cuda(e, "pilott_mem_compact: gather of the surviving rows"); }
        h->launches += 1;
    }
    cudaFree(old.vec); cudaFree(old.inv_norm); cudaFree(old.prio); cudaFree(old.tagmask);
    cudaFree(old.expires); cudaFree(old.scale); cudaFree(old.tile_max);
    h->head = 0; h->count = m;       // total is unchanged: survivors are re-based to total-m+age
    ++h->generation;
    return PILOTT_OK;
}

int pilott_mem_export(pilott_mem_t* h, uint16_t* vec_bits_host, float* inv_norm_host, int32_t* prio_host,
                      uint64_t* tagmask_host, int64_t* expires_us_host) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    DeviceGuard g(h->device);
    CK(cudaDeviceSynchronize());
    const int64_t n = h->count;
    // two contiguous physical runs in age order: [head, min(head+n, cap)) then [0, rest)
    const int64_t run1 = (h->head + n <= h->capacity) ? n : h->capacity - h->head;
    const int64_t run2 = n - run1;
    if (vec_bits_host) {
        CK(cudaMemcpy2D(vec_bits_host, (size_t)h->dim * 2, (const char*)h->vec + (size_t)h->head * h->ld * 2,
                        (size_t)h->ld * 2, (size_t)h->dim * 2, (size_t)run1, cudaMemcpyDeviceToHost));
        if (run2 > 0)
            CK(cudaMemcpy2D(vec_bits_host + (size_t)run1 * h->dim, (size_t)h->dim * 2, h->vec, (size_t)h->ld * 2,
                            (size_t)h->dim * 2, (size_t)run2, cudaMemcpyDeviceToHost));
    }
#define EXPORT_COL(dst, src, T)                                                                       \
    if (dst) {                                                                                        \
        CK(cudaMemcpy(dst, (const T*)src + h->head, (size_t)run1 * sizeof(T), cudaMemcpyDeviceToHost)); \
        if (run2 > 0) CK(cudaMemcpy(dst + run1, src, (size_t)run2 * sizeof(T), cudaMemcpyDeviceToHost)); \
    }
    EXPORT_COL(inv_norm_host, h->inv_norm, float)
    EXPORT_COL(prio_host, h->prio, int32_t)
    EXPORT_COL(tagmask_host, h->tagmask, uint64_t)
    EXPORT_COL(expires_us_host, h->expires, int64_t)
#undef EXPORT_COL
    return PILOTT_OK;
}

int pilott_topk_merge(const int64_t* idx_dev, const double* score64_dev, int n_shards, int64_t nq, int k,
                      int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev, void* stream) {
    if (nq < 0 || n_shards < 1 || k < 1) return fail(PILOTT_EINVAL, "bad merge shape");
    if ((int64_t)n_shards * k > 2048) return fail(PILOTT_EINVAL, "n_shards * k must be <= 2048");
    if (nq > 0 && (!idx_dev || !score64_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL pointer");
    CK(pilott::launch_topk_merge(idx_dev, score64_dev, n_shards, nq, k, nq * k, out_idx_dev, out_score_dev,
                                 out_score64_dev, (cudaStream_t)stream));
    return PILOTT_OK;
}

int pilott_topk_merge_packed(const int64_t* packed_dev, int n_shards, int64_t nq, int k,
                             int64_t* out_idx_dev, float* out_score_dev, double* out_score64_dev, void* stream) {
    if (nq < 0 || n_shards < 1 || k < 1) return fail(PILOTT_EINVAL, "bad merge shape");
    if ((int64_t)n_shards * k > 2048) return fail(PILOTT_EINVAL, "n_shards * k must be <= 2048");
    if (nq > 0 && (!packed_dev || !out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL pointer");
    CK(pilott::launch_topk_merge(packed_dev, reinterpret_cast<const double*>(packed_dev + nq * k), n_shards, nq, k,
                                 2 * nq * k, out_idx_dev, out_score_dev, out_score64_dev, (cudaStream_t)stream));
    return PILOTT_OK;
}

/* ---- exchange over NVLink peer memory (row-sharded store) ---- */
int pilott_xchg_create(int device, int rank, int world, int64_t block_bytes, pilott_xchg_t** out, void* handle_out) {
    if (!out || !handle_out) return fail(PILOTT_EINVAL, "NULL output pointer");
    *out = nullptr;
    if (world < 1 || world > 64 || rank < 0 || rank >= world) return fail(PILOTT_EINVAL, "bad rank/world");
    if (block_bytes < 16 || block_bytes % 16) return fail(PILOTT_EINVAL, "block_bytes must be a positive multiple of 16");
    static_assert(sizeof(cudaIpcMemHandle_t) == PILOTT_IPC_HANDLE_BYTES, "IPC handle size");
    DeviceGuard g(device);
    auto* x = new pilott_xchg();
    x->device = device; x->rank = rank; x->world = world; x->block_bytes = block_bytes;
    x->flags_off = pilott::XCHG_SLOTS * (int64_t)world * block_bytes;     // see xchg.cu
    x->total_bytes = x->flags_off + pilott::XCHG_SLOTS * (int64_t)world * 8;
    cudaError_t e = cudaMalloc(&x->mine, (size_t)x->total_bytes);
    if (e == cudaSuccess) e = cudaMemset(x->mine, 0, (size_t)x->total_bytes);
    if (e == cudaSuccess) e = cudaMalloc(&x->peers_dev, (size_t)world * sizeof(void*));
    if (e == cudaSuccess) e = cudaMalloc(&x->err_dev, 4);
    if (e == cudaSuccess) e = cudaMemset(x->err_dev, 0, 4);
    if (e == cudaSuccess) e = cudaMalloc(&x->done_dev, 4);
    if (e == cudaSuccess) e = cudaMemset(x->done_dev, 0, 4);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&x->copy_stream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaMalloc((void**)&x->flag_vals, pilott::XCHG_FLAG_RING * sizeof(uint64_t));
    if (e == cudaSuccess) {
        std::vector<uint64_t> vals((size_t)pilott::XCHG_FLAG_RING);
        for (int i = 0; i < pilott::XCHG_FLAG_RING; ++i) vals[(size_t)i] = (uint64_t)i + 1;
        e = cudaMemcpy(x->flag_vals, vals.data(), vals.size() * sizeof(uint64_t), cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&x->ev_ready, cudaEventDisableTiming);
    for (int i = 0; i < 8 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&x->ev_buf[i], cudaEventDisableTiming);
    cudaIpcMemHandle_t hnd;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&hnd, x->mine);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        cudaFree(x->mine); cudaFree(x->peers_dev); cudaFree(x->err_dev); cudaFree(x->done_dev); cudaFree(x->flag_vals);
        if (x->copy_stream) cudaStreamDestroy(x->copy_stream);
        if (x->ev_ready) cudaEventDestroy(x->ev_ready);
        for (cudaEvent_t ev : x->ev_buf) if (ev) cudaEventDestroy(ev);
        delete x;
        return fail_cuda(e, "pilott_xchg_create");
    }
    memcpy(handle_out, &hnd, sizeof(hnd));
    *out = x;
    return PILOTT_OK;
}

int pilott_xchg_connect(pilott_xchg_t* x, const void* handles) {
    if (!x || !handles) return fail(PILOTT_EINVAL, "NULL pointer");
    if (x->connected) return fail(PILOTT_EINVAL, "already connected");
    DeviceGuard g(x->device);
    x->peers.assign((size_t)x->world, nullptr);
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank) { x->peers[r] = x->mine; continue; }
        cudaIpcMemHandle_t hnd;
        memcpy(&hnd, (const char*)handles + (size_t)r * sizeof(hnd), sizeof(hnd));
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, hnd, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            for (int j = 0; j < r; ++j) if (j != x->rank && x->peers[j]) cudaIpcCloseMemHandle(x->peers[j]);
            x->peers.clear();
            return fail_cuda(e, "cudaIpcOpenMemHandle (peer buffer)");
        }
        x->peers[r] = (unsigned char*)p;
    }
    CK(cudaMemcpy(x->peers_dev, x->peers.data(), (size_t)x->world * sizeof(void*), cudaMemcpyHostToDevice));
    x->connected = true;
    return PILOTT_OK;
}

int pilott_mem_search_publish(pilott_mem_t* h, pilott_xchg_t* x, int64_t step, const float* q_dev, int64_t nq, int k,
                              int32_t min_prio, uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev,
                              float* out_score_dev, double* out_score64_dev, int32_t* out_inexact_dev,
                              uint8_t* out_flags_dev, void* stream) {
    if (!h || !x) return fail(PILOTT_EINVAL, "handle is NULL");
    if (!x->connected) return fail(PILOTT_EINVAL, "exchange is not connected");
    if (x->device != h->device) return fail(PILOTT_EINVAL, "store and exchange live on different devices");
    if (nq < 1 || step < 0) return fail(PILOTT_EINVAL, "nq must be >= 1 and step >= 0 (every rank publishes every step)");
    if (2 * nq * k * 8 > x->block_bytes) return fail(PILOTT_EINVAL, "[2, nq, k] block exceeds block_bytes");
    if (!q_dev || !out_idx_dev || !out_score_dev) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    DeviceGuard g(h->device);
    int rc = order_device_call(h, (cudaStream_t)stream);
    if (rc) return rc;
    rc = search_core(h, q_dev, nq, k, min_prio, req_tagmask, now_us, out_idx_dev, out_score_dev,
                     out_score64_dev, out_inexact_dev, out_flags_dev, (cudaStream_t)stream, x, step);
    if (rc) return rc;
    return note_device_call(h, (cudaStream_t)stream);
}

int pilott_xchg_publish(pilott_xchg_t* x, const void* block_dev, int64_t bytes, int64_t step, void* stream) {
    if (!x || !block_dev) return fail(PILOTT_EINVAL, "NULL pointer");
    if (!x->connected) return fail(PILOTT_EINVAL, "exchange is not connected");
    if (bytes < 16 || bytes % 16 || bytes > x->block_bytes) return fail(PILOTT_EINVAL, "bytes must be a multiple of 16 within block_bytes");
    if (step < 0) return fail(PILOTT_EINVAL, "step is negative");
    DeviceGuard g(x->device);
    CK(pilott::launch_xchg_publish(x->peers_dev, x->world, x->rank, x->block_bytes, x->flags_off, block_dev, bytes,
                                   step, (cudaStream_t)stream));
    return PILOTT_OK;
}

int pilott_xchg_publish_copy(pilott_xchg_t* x, const void* block_dev, int64_t bytes, int64_t step, int buffer_id,
                             void* stream) {
    if (!x || !block_dev) return fail(PILOTT_EINVAL, "NULL pointer");
    if (!x->connected) return fail(PILOTT_EINVAL, "exchange is not connected");
    if (bytes < 16 || bytes % 16 || bytes > x->block_bytes) return fail(PILOTT_EINVAL, "bytes must be a multiple of 16 within block_bytes");
    if (step < 0) return fail(PILOTT_EINVAL, "step is negative");
    if (buffer_id < 0 || buffer_id >= 8) return fail(PILOTT_EINVAL, "buffer_id must be in [0, 8)");
    DeviceGuard g(x->device);
    const int slot = (int)(step & (pilott::XCHG_SLOTS - 1));
    // the block is complete when `stream` gets here; the copies run beside whatever `stream` does next
    CK(cudaEventRecord(x->ev_ready, (cudaStream_t)stream));
    CK(cudaStreamWaitEvent(x->copy_stream, x->ev_ready, 0));
    for (int i = 0; i < x->world; ++i) {
        const int r = (x->rank + i) % x->world;           // every rank starts with a different destination
        unsigned char* base = x->peers[r];
        CK(cudaMemcpyAsync(base + ((int64_t)slot * x->world + x->rank) * x->block_bytes, block_dev, (size_t)bytes,
                           cudaMemcpyDeviceToDevice, x->copy_stream));
        // same stream: the flag lands after the block it announces
        CK(cudaMemcpyAsync(base + x->flags_off + ((int64_t)slot * x->world + x->rank) * 8, x->flag_vals + (step & (pilott::XCHG_FLAG_RING - 1)), 8,
                           cudaMemcpyDeviceToDevice, x->copy_stream));
    }
    CK(cudaEventRecord(x->ev_buf[buffer_id], x->copy_stream));
    return PILOTT_OK;
}

int pilott_xchg_wait_buffer(pilott_xchg_t* x, int buffer_id, void* stream) {
    if (!x) return fail(PILOTT_EINVAL, "NULL pointer");
    if (buffer_id < 0 || buffer_id >= 8) return fail(PILOTT_EINVAL, "buffer_id must be in [0, 8)");
    DeviceGuard g(x->device);
    // (a never-recorded event is complete; one that has completed needs no wait in the stream, and a
    //  wait between two kernels would cost them their programmatic overlap)
    const cudaError_t q = cudaEventQuery(x->ev_buf[buffer_id]);
    if (q == cudaErrorNotReady) CK(cudaStreamWaitEvent((cudaStream_t)stream, x->ev_buf[buffer_id], 0));
    else if (q != cudaSuccess) return fail_cuda(q, "cudaEventQuery(copy done)");
    return PILOTT_OK;
}

int pilott_xchg_merge(pilott_xchg_t* x, int64_t step, int64_t nq, int k, int64_t* out_idx_dev, float* out_score_dev,
                      double* out_score64_dev, void* stream) {
    if (!x) return fail(PILOTT_EINVAL, "NULL pointer");
    if (!x->connected) return fail(PILOTT_EINVAL, "exchange is not connected");
    if (nq < 0 || k < 1 || step < 0) return fail(PILOTT_EINVAL, "bad merge shape");
    if ((int64_t)x->world * k > 2048) return fail(PILOTT_EINVAL, "world * k must be <= 2048");
    if (2 * nq * k * 8 > x->block_bytes) return fail(PILOTT_EINVAL, "[2, nq, k] block exceeds block_bytes");
    if (nq > 0 && (!out_idx_dev || !out_score_dev)) return fail(PILOTT_EINVAL, "NULL output pointer");
    DeviceGuard g(x->device);
    CK(pilott::launch_xchg_merge(x->mine, x->world, x->block_bytes, x->flags_off, step, nq, k, out_idx_dev,
                                 out_score_dev, out_score64_dev, x->err_dev, (cudaStream_t)stream));
    return PILOTT_OK;
}

int pilott_mem_search_sharded(pilott_mem_t* h, pilott_xchg_t* x, int64_t step, const float* q_dev, int64_t nq, int k,
                              int32_t min_prio, uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev,
                              float* out_score_dev, double* out_score64_dev, int32_t* out_inexact_dev,
                              uint8_t* out_flags_dev, void* block_dev, void* stream) {
    if (!h || !x) return fail(PILOTT_EINVAL, "handle is NULL");
    if (!x->connected) return fail(PILOTT_EINVAL, "exchange is not connected");
    if (x->device != h->device) return fail(PILOTT_EINVAL, "store and exchange live on different devices");
    if (nq < 1 || step < 0) return fail(PILOTT_EINVAL, "nq must be >= 1 and step >= 0 (every rank takes part in every step)");
    if (2 * nq * k * 8 > x->block_bytes) return fail(PILOTT_EINVAL, "[2, nq, k] block exceeds block_bytes");
    if (!q_dev || !out_idx_dev || !out_score_dev || !block_dev) return fail(PILOTT_EINVAL, "NULL query, output or block pointer");
    DeviceGuard g(h->device);
    cudaStream_t st = (cudaStream_t)stream;
    int rc = order_device_call(h, st);
    if (rc) return rc;
    // local top-k of this rank's rows (global indices) into the caller's block: [0] indices, [1] fp64 score bits
    int64_t* blk_idx = (int64_t*)block_dev;
    double* blk_sc = (double*)block_dev + nq * k;
    rc = pilott_xchg_wait_buffer(x, (int)(step & 7), stream);          // the copies of step - 8 have left the block
    if (rc) return rc;
    rc = search_core(h, q_dev, nq, k, min_prio, req_tagmask, now_us, blk_idx, out_score_dev, blk_sc, out_inexact_dev,
                     out_flags_dev, st);
    if (rc) return rc;
    rc = note_device_call(h, st);
    if (rc) return rc;
    rc = pilott_xchg_publish_copy(x, block_dev, 2 * nq * k * 8, step, (int)(step & 7), stream);
    if (rc) return rc;
    return pilott_xchg_merge(x, step, nq, k, out_idx_dev, out_score_dev, out_score64_dev, stream);
}

int pilott_xchg_check(pilott_xchg_t* x) {
    if (!x) return fail(PILOTT_EINVAL, "NULL pointer");
    DeviceGuard g(x->device);
    int32_t err = 0;
    CK(cudaMemcpy(&err, x->err_dev, 4, cudaMemcpyDeviceToHost));      // synchronises with prior work
    if (err) return fail(PILOTT_ECUDA, "a merge waited 2 s for a peer's block and gave up (peer stalled or died)");
    return PILOTT_OK;
}

int pilott_xchg_destroy(pilott_xchg_t* x) {
    if (!x) return PILOTT_OK;
    DeviceGuard g(x->device);
    cudaDeviceSynchronize();
    for (int r = 0; r < (int)x->peers.size(); ++r)
        if (r != x->rank && x->peers[r]) cudaIpcCloseMemHandle(x->peers[r]);
    cudaFree(x->mine); cudaFree(x->peers_dev); cudaFree(x->err_dev); cudaFree(x->done_dev);
    if (x->copy_stream) cudaStreamDestroy(x->copy_stream);
    cudaFree(x->flag_vals);
    if (x->ev_ready) cudaEventDestroy(x->ev_ready);
    for (cudaEvent_t ev : x->ev_buf) if (ev) cudaEventDestroy(ev);
    delete x;
    return PILOTT_OK;
}

int pilott_route_score(const uint64_t* task_req_dev, const uint64_t* task_tags_dev, const int32_t* task_type_dev,
                       const uint8_t* task_flags_dev, int64_t nt, const uint64_t* agent_caps_dev,
                       const uint64_t* agent_spec_dev, const uint8_t* agent_flags_dev,
                       const double* agent_s12_dev, const double* agent_t4_dev, int na, int words,
                       const double* spec_u_dev, int max_tags, int mode, int32_t* out_agent_dev,
                       double* out_score_dev, void* stream) {
    if (nt < 0 || na < 0) return fail(PILOTT_EINVAL, "negative size");
    if (words < 1 || words > 64) return fail(PILOTT_EINVAL, "words must be in [1, 64]");
    if (max_tags < 64 * words) return fail(PILOTT_EINVAL, "max_tags must be >= 64*words");
    if (mode < 0 || mode > 2) return fail(PILOTT_EINVAL, "mode must be 0, 1 or 2");
    if (nt == 0) return PILOTT_OK;
    if (!task_req_dev || !task_tags_dev || !task_type_dev || !task_flags_dev || !out_agent_dev || !out_score_dev)
        return fail(PILOTT_EINVAL, "NULL task or output pointer");
    if (na > 0 && (!agent_caps_dev || !agent_spec_dev || !agent_flags_dev || !agent_s12_dev || !agent_t4_dev || !spec_u_dev))
        return fail(PILOTT_EINVAL, "NULL agent table pointer");
    CK(pilott::launch_route(task_req_dev, task_tags_dev, task_type_dev, task_flags_dev, nt, agent_caps_dev,
                            agent_spec_dev, agent_flags_dev, agent_s12_dev, agent_t4_dev, na, words, spec_u_dev,
                            max_tags, mode, out_agent_dev, out_score_dev, (cudaStream_t)stream));
    return PILOTT_OK;
}

int64_t pilott_route_tc_workspace(int64_t nt, int na, int words) {
    if (nt < 0 || na < 0 || words < 1) return 0;
    return pilott::route_tc_workspace_bytes(nt, na, words);
}

int pilott_route_score_tc(const uint64_t* task_req_dev, const uint64_t* task_tags_dev, const int32_t* task_type_dev,
                          const uint8_t* task_flags_dev, int64_t nt, const uint64_t* agent_caps_dev,
                          const uint64_t* agent_spec_dev, const uint8_t* agent_flags_dev,
                          const double* agent_s12_dev, const double* agent_t4_dev, int na, int words,
                          const double* spec_u_dev, int max_tags, int mode, int32_t* out_agent_dev,
                          double* out_score_dev, void* workspace_dev, int64_t workspace_bytes, void* stream) {
    if (nt < 0 || na < 0) return fail(PILOTT_EINVAL, "negative size");
    if (max_tags < 64 * words) return fail(PILOTT_EINVAL, "max_tags must be >= 64*words");
    if (mode < 0 || mode > 2) return fail(PILOTT_EINVAL, "mode must be 0, 1 or 2");
    if (nt == 0) return PILOTT_OK;
    if (!pilott::route_tc_supported(na, words, max_tags))
        return fail(PILOTT_EUNSUPPORTED, "tensor-core routing takes 1..3 mask words and an agent table that fits shared memory "
                                         "(about 1500 agents): use pilott_route_score");
    if (!task_req_dev || !task_tags_dev || !task_type_dev || !task_flags_dev || !out_agent_dev || !out_score_dev ||
        !agent_caps_dev || !agent_spec_dev || !agent_flags_dev || !agent_s12_dev || !agent_t4_dev || !spec_u_dev)
        return fail(PILOTT_EINVAL, "NULL pointer");
    if (!workspace_dev || workspace_bytes < pilott::route_tc_workspace_bytes(nt, na, words))
        return fail(PILOTT_EINVAL, "workspace is smaller than pilott_route_tc_workspace(nt, na, words)");
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    int rc = get_encode();
    if (rc) return rc;
    CK(pilott::launch_route_tc(task_req_dev, task_tags_dev, task_type_dev, task_flags_dev, nt, agent_caps_dev, agent_spec_dev,
                               agent_flags_dev, agent_s12_dev, agent_t4_dev, na, words, spec_u_dev, max_tags, mode,
                               out_agent_dev, out_score_dev, workspace_dev, sms, &encode_map, (cudaStream_t)stream));
    return PILOTT_OK;
}

int pilott_task_priority(const uint8_t* urgent_dev, const int32_t* complexity_dev, const int32_t* n_deps_dev,
                         int64_t nt, uint8_t* out_dev, void* stream) {
    if (nt < 0) return fail(PILOTT_EINVAL, "negative size");
    if (nt == 0) return PILOTT_OK;
    if (!urgent_dev || !complexity_dev || !n_deps_dev || !out_dev) return fail(PILOTT_EINVAL, "NULL pointer");
    CK(pilott::launch_task_priority(urgent_dev, complexity_dev, n_deps_dev, nt, out_dev, (cudaStream_t)stream));
    return PILOTT_OK;
}

int pilott_mem_search_exact_noalloc(pilott_mem_t* h, const float* q_dev, int64_t nq,
                                    const int32_t* query_ids_dev, int64_t nsel, int k, int32_t min_prio,
                                    uint64_t req_tagmask, int64_t now_us, int64_t* out_idx_dev,
                                    float* out_score_dev, double* out_score64_dev, void* stream) {
    if (!h) return fail(PILOTT_EINVAL, "handle is NULL");
    if (nq < 0) return fail(PILOTT_EINVAL, "nq is negative");
    if (k < 1) return fail(PILOTT_EINVAL, "k must be >= 1");
    if (!query_ids_dev) nsel = nq;
    if (nq == 0 || nsel <= 0) return PILOTT_OK;
    if (!q_dev || !out_idx_dev || !out_score_dev) return fail(PILOTT_EINVAL, "NULL query or output pointer");
    if (nsel > 65535) return fail(PILOTT_EINVAL, "pilott_mem_search_exact_noalloc: at most 65535 selected queries per call");
    // Layout inside the fixed buffer: [bf16 queries, nq x row_stride][fp64 scores, nsel x rows held]
    const uint64_t q_bytes = (uint64_t)round_up(nq * (int64_t)h->ld * 2, 256);
    const uint64_t s_bytes = (uint64_t)nsel * (uint64_t)(h->count > 0 ? h->count : 1) * 8u;
    if (q_bytes + s_bytes > (uint64_t)PILOTT_NOALLOC_SCRATCH_BYTES)
        return fail(PILOTT_EINVAL,
                    "pilott_mem_search_exact_noalloc: this call needs " + std::to_string(q_bytes + s_bytes) +
                    " bytes of scratch but the built-in buffer holds " + std::to_string(PILOTT_NOALLOC_SCRATCH_BYTES) +
                    " (8 bytes per selected query per stored row): select fewer queries per call, or use "
                    "pilott_mem_search_exact, which allocates");
    DeviceGuard g(h->device);
    unsigned char* scratch = nullptr;
    int rc = noalloc_scratch(h->device, &scratch);      // cached by pilott_mem_create: no lookup, no module load here
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    rc = order_device_call(h, st);
    if (rc) return rc;
    // Nothing in the buffer has to start at zero: the query block and every score the selection
    // reads are rewritten by the two kernels below before they are read.
    CK(pilott::launch_convert_queries(q_dev, nq, h->dim, h->ld, scratch, st));
    CK(pilott::launch_exact(scratch, h->vec, h->ld, h->prio, h->tagmask, h->expires, h->count, h->capacity,
                            h->head, h->total - h->count, h->seq_mul, h->seq_add, query_ids_dev, nsel, k,
                            min_prio, req_tagmask, now_us, reinterpret_cast<double*>(scratch + q_bytes),
                            out_idx_dev, out_score_dev, out_score64_dev, st));
    h->launches += h->count > 0 ? 3 : 2;                // convert, [scores,] select
    return note_device_call(h, st);
}

}  // extern "C"
