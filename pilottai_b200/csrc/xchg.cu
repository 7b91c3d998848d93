// xchg.cu -- exchange step of the row-sharded store over NVLink peer memory, without NCCL.
//
// Every rank owns one buffer  [8 slots][world][block_bytes]  plus  flags [8 slots][world]  and
// has every peer's buffer mapped (CUDA IPC).  Step s uses slot s & 7.
//   publish: one CTA per destination rank stores this rank's (index, score) block straight into
//            the destination's buffer at [slot][my rank] with 16-byte stores over NVLink, fences
//            at system scope and then releases flag[slot][my rank] = (s mod 65536) + 1 there.
//   merge:   a one-CTA kernel first acquires the `world` flags of the slot (spins until each reads
//            that value: all blocks have landed), then the merge kernel ranks the world*k candidates of its query by
//            (score desc, index asc) exactly like topk_merge_kernel.
//   publish by copy (abi.cu, pilott_xchg_publish_copy): the same block and flag written by the copy
//            engines instead (cudaMemcpyAsync into each peer's buffer, then an 8-byte copy of the
//            flag value, same stream, so in order): no SM is taken from the scan of the next batch.
// Ordering rule per rank: up to three publishes may be enqueued ahead of the oldest outstanding
// merge -- publish(s+1..s+3) before merge(s), so the exchange of a batch hides behind the local
// searches that follow and ranks may drift apart -- but publish(s+4) must follow merge(s).  Eight
// slots are then enough: a peer overwrites my slot s & 7 with step s + 8 only after its own
// merge(s + 4), which needed my publish(s + 4), which my stream ordered after my merge(s).
// The spin is bounded (2 s of %globaltimer): a missing peer sets an error word and the merge
// writes "no result" rows instead of hanging the GPU.
#include "common.cuh"
#include "internal.h"

namespace pilott {

__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// peers[r] = base of rank r's buffer as mapped in THIS process.
__global__ void __launch_bounds__(512)
xchg_publish_kernel(unsigned char* const* __restrict__ peers, int world, int rank, int64_t block_bytes,
                    int64_t flags_off, const uint4* __restrict__ src, int64_t bytes, int64_t step) {
    const int dst = blockIdx.x;
    const int slot = (int)(step & (XCHG_SLOTS - 1));
    unsigned char* base = peers[dst];
    uint4* out = reinterpret_cast<uint4*>(base + ((int64_t)slot * world + rank) * block_bytes);
    const int64_t n16 = bytes / 16;
    for (int64_t i = threadIdx.x; i < n16; i += blockDim.x) out[i] = src[i];
    __syncthreads();                                   // the CTA's stores happen before thread 0's fence
    if (threadIdx.x == 0) {
        __threadfence_system();                        // one cumulative fence, then the release
        uint64_t* flag = reinterpret_cast<uint64_t*>(base + flags_off) + (int64_t)slot * world + rank;
        st_release_sys(flag, PILOTT_XCHG_FLAG(step));
    }
}

// The wait itself: ONE small CTA spins until every rank's flag of the slot reads this step's value
// (bounded: 2 s, then the error word).  It is a kernel of its own so that a rank that runs ahead of
// its peers parks a single CTA -- a thousand spinning merge CTAs would take every thread slot of the
// GPU, and with them the slots the rank's own outgoing copies and searches need: two ranks waiting
// for each other that way never finish.
__global__ void __launch_bounds__(64)
xchg_wait_kernel(const unsigned char* __restrict__ mine, int world, int64_t flags_off, int64_t step,
                 int32_t* __restrict__ err_word) {
    const int slot = (int)(step & (XCHG_SLOTS - 1));
    for (int r = threadIdx.x; r < world; r += blockDim.x) {
        const uint64_t* flag = reinterpret_cast<const uint64_t*>(mine + flags_off) + (int64_t)slot * world + r;
        const uint64_t want = PILOTT_XCHG_FLAG(step);
        const uint64_t t0 = globaltimer_ns();
        while (ld_acquire_sys(flag) != want) {
            __nanosleep(200);
            if (globaltimer_ns() - t0 > 2000000000ull) { atomicExch(err_word, 1); break; }
        }
    }
}

__global__ void __launch_bounds__(256)
xchg_merge_kernel(const unsigned char* __restrict__ mine, int world, int64_t block_bytes, int64_t flags_off,
                  int64_t step, int64_t nq, int k, int64_t* __restrict__ out_idx, float* __restrict__ out_score,
                  double* __restrict__ out_score64, int32_t* __restrict__ err_word) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    __shared__ int s_ok;
    const int total = world * k;
    double* s_sc = reinterpret_cast<double*>(sm_raw);
    int64_t* s_idx = reinterpret_cast<int64_t*>(s_sc + total);
    const int slot = (int)(step & (XCHG_SLOTS - 1));
    const int64_t q = blockIdx.x;
    if (threadIdx.x == 0) s_ok = 1;
    __syncthreads();
    if (threadIdx.x < world) {                          // one thread per peer flag; xchg_wait_kernel has already waited
        const uint64_t* flag = reinterpret_cast<const uint64_t*>(mine + flags_off) + (int64_t)slot * world + threadIdx.x;
        if (ld_acquire_sys(flag) != PILOTT_XCHG_FLAG(step)) { s_ok = 0; atomicExch(err_word, 1); }
    }
    __syncthreads();
    const bool ok = s_ok != 0;
    const unsigned char* sl = mine + (int64_t)slot * world * block_bytes;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int sh = i / k, j = i % k;
        const int64_t* blk = reinterpret_cast<const int64_t*>(sl + (int64_t)sh * block_bytes);   // [2][nq][k]
        // .cg: the slot was read four steps ago, a stale L1 line must not serve these loads
        s_idx[i] = ok ? __ldcg(blk + q * k + j) : -1;
        s_sc[i] = ok ? __longlong_as_double(__ldcg(blk + nq * k + q * k + j)) : -INFINITY;
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
        for (int j = 0; j < total; ++j) {
            const int64_t oj = s_idx[j];
            const double sj = s_sc[j];
            if (oj >= 0 && (sj > ms || (sj == ms && oj < me))) ++rank;
        }
        if (rank < k) {
            out_idx[q * k + rank] = me;
            out_score[q * k + rank] = (float)ms;
            if (out_score64) out_score64[q * k + rank] = ms;
        }
    }
}

cudaError_t launch_xchg_publish(unsigned char* const* peers_dev, int world, int rank, int64_t block_bytes,
                                int64_t flags_off, const void* src, int64_t bytes, int64_t step, cudaStream_t st) {
    xchg_publish_kernel<<<(unsigned)world, 512, 0, st>>>(peers_dev, world, rank, block_bytes, flags_off,
                                                         (const uint4*)src, bytes, step);
    return cudaGetLastError();
}

cudaError_t launch_xchg_merge(const void* mine, int world, int64_t block_bytes, int64_t flags_off, int64_t step,
                              int64_t nq, int k, int64_t* out_idx, float* out_score, double* out_score64,
                              int32_t* err_word, cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    xchg_wait_kernel<<<1, 64, 0, st>>>((const unsigned char*)mine, world, flags_off, step, err_word);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    xchg_merge_kernel<<<(unsigned)nq, 256, (size_t)world * k * 16, st>>>((const unsigned char*)mine, world, block_bytes,
                                                                        flags_off, step, nq, k, out_idx, out_score,
                                                                        out_score64, err_word);
    return cudaGetLastError();
}

}  // namespace pilott
