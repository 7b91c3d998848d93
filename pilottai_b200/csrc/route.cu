// route.cu -- K5: task x agent suitability scoring + masked first-max argmax.
//
// One thread per task; the agent table is staged through shared memory in chunks and read
// by all lanes at the same address (broadcast).  Matching is set arithmetic on 64-bit
// words (required capabilities subset test, type membership bit, tag intersection
// popcount); the score is assembled from host-precomputed per-agent fp64 terms with at
// most two IEEE fp64 additions (one rounding each, no FMA contraction), which reproduces
// Python's left-to-right evaluation of
//   base*0.4 + (1-load)*0.3 + spec*0.2 + perf*0.1       (pilott/core/router.py:83-88)
// bit for bit.  Strict `>` while scanning agents in table order keeps the FIRST maximum,
// as max()/`>` do in the reference (router.py:62, load_balancer.py:285-287,
// task_delegator.py:82).
#include "common.cuh"
#include "internal.h"

namespace pilott {

// Shared-memory agent record, in 8-byte words (all lanes read the same record = broadcast):
//   [0, nw) capability words | [nw, 2nw) specialization words | t4 | flags | s12[0..2] | pad to even
// so that for one-word vocabularies (caps, spec) and (t4, flags) each come with one 16-byte load.
__host__ __device__ inline int route_record_words(int nw) { return (2 * nw + 5 + 1) & ~1; }

template <int W, int MODE>
__global__ void __launch_bounds__(256)
route_kernel(const uint64_t* __restrict__ task_req, const uint64_t* __restrict__ task_tags,
             const int32_t* __restrict__ task_type, const uint8_t* __restrict__ task_flags, int64_t nt,
             const uint64_t* __restrict__ agent_caps, const uint64_t* __restrict__ agent_spec,
             const uint8_t* __restrict__ agent_flags, const double* __restrict__ agent_s12,
             const double* __restrict__ agent_t4, int na, int words, const double* __restrict__ spec_u,
             int max_tags, int chunk, int32_t* __restrict__ out_agent,
             double* __restrict__ out_score) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int nw = W > 0 ? W : words;
    const int RW = route_record_words(nw);
    uint64_t* s_rec = reinterpret_cast<uint64_t*>(sm_raw);             // [chunk][RW]
    double* s_u = reinterpret_cast<double*>(s_rec + (size_t)chunk * RW);   // [2*(max_tags+1)]

    for (int i = threadIdx.x; i < 2 * (max_tags + 1); i += blockDim.x) s_u[i] = spec_u[i];

    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = t < nt;
    uint64_t req[W > 0 ? W : 1], tags[W > 0 ? W : 1];
    int type = -1;
    uint8_t tf = 0;
    if (live) {
        type = task_type[t];
        tf = task_flags[t];
        if (W > 0) {
#pragma unroll
            for (int w = 0; w < (W > 0 ? W : 1); ++w) { req[w] = task_req[t * nw + w]; tags[w] = task_tags[t * nw + w]; }
        }
    }
    const bool has_req_key = tf & 1, has_type_key = tf & 2, req_unknown = tf & 4;
    // type membership as one AND: word index + single-bit mask (mask 0 = task names no type)
    const int tw = type >= 0 ? (type >> 6) : 0;
    const uint64_t tm = type >= 0 ? (1ull << (type & 63)) : 0ull;
    // agent.py:551-556 as one mask test: the base score is 0 when the task carries the key and the
    // agent lacks the attribute (flag bit 1 clear), lacks a required capability, or the task names
    // a capability no agent knows (bit 2, which no agent flag ever has).  Without the key nothing
    // is required at all.
    const uint32_t need = has_req_key ? (2u | (req_unknown ? 4u : 0u)) : 0u;
    if (W > 0 && !has_req_key) {
#pragma unroll
        for (int w = 0; w < (W > 0 ? W : 1); ++w) req[w] = 0ull;
    }
    const int u_hi = max_tags + 1;
    double best = 0.0;
    int best_a = -1;

    for (int a0 = 0; a0 < na; a0 += chunk) {
        const int nc = min(chunk, na - a0);
        __syncthreads();
        for (int i = threadIdx.x; i < nc; i += blockDim.x) {
            uint64_t* r = s_rec + (size_t)i * RW;
            for (int w = 0; w < nw; ++w) {
                r[w] = agent_caps[(size_t)(a0 + i) * nw + w];
                r[nw + w] = agent_spec[(size_t)(a0 + i) * nw + w];
            }
            r[2 * nw] = (uint64_t)__double_as_longlong(agent_t4[a0 + i]);
            r[2 * nw + 1] = agent_flags[a0 + i] & 3u;
            r[2 * nw + 2] = (uint64_t)__double_as_longlong(agent_s12[(size_t)(a0 + i) * 3]);
            r[2 * nw + 3] = (uint64_t)__double_as_longlong(agent_s12[(size_t)(a0 + i) * 3 + 1]);
            r[2 * nw + 4] = (uint64_t)__double_as_longlong(agent_s12[(size_t)(a0 + i) * 3 + 2]);
        }
        __syncthreads();
        if (!live) continue;
        // score of agent a of this chunk; false = agent skipped for this task.  Branch-free.
        auto eval = [&](const int a, double& sc) -> bool {
            const uint64_t* r = s_rec + (size_t)a * RW;
            uint64_t miss = 0, tsel = 0;
            int tagc = 0;
            if (W == 1) {
                const ulonglong2 cs = *reinterpret_cast<const ulonglong2*>(r);          // caps, spec
                miss = req[0] & ~cs.x;
                tagc = __popcll(tags[0] & cs.y);
                tsel = cs.y & tm;
            } else if (W == 2) {
                const ulonglong2 c = *reinterpret_cast<const ulonglong2*>(r);
                const ulonglong2 sp = *reinterpret_cast<const ulonglong2*>(r + 2);
                miss = (req[0] & ~c.x) | (req[1] & ~c.y);
                tagc = __popcll(tags[0] & sp.x) + __popcll(tags[1] & sp.y);
                tsel = (tw ? sp.y : sp.x) & tm;
            } else {
                for (int w = 0; w < nw; ++w) {
                    miss |= task_req[t * nw + w] & ~r[w];
                    tagc += __popcll(task_tags[t * nw + w] & r[nw + w]);
                }
                tsel = r[nw + tw] & tm;
                if (!has_req_key) miss = 0;
            }
            const ulonglong2 tfl = *reinterpret_cast<const ulonglong2*>(r + 2 * nw);    // t4, flags
            const uint32_t f = (uint32_t)tfl.y;
            const bool tbit = tsel != 0;
            const bool zero_base = ((uint32_t)miss | (uint32_t)(miss >> 32) | (need & ~f)) != 0u;
            const int sel = zero_base ? 0 : ((has_type_key & tbit) ? 2 : 1);           // agent.py:562-564
            sc = __longlong_as_double((long long)r[2 * nw + 2 + sel]);
            if (MODE == 0) {
                const double u = s_u[(tbit ? u_hi : 0) + tagc];                        // router.py:116-126
                sc = __dadd_rn(__dadd_rn(sc, u), __longlong_as_double((long long)tfl.x));
            }
            bool ok = f & 1;                                           // not viable / skipped
            if (MODE == 2) ok &= sc > 0.0;                             // task_delegator.py:75
            return ok;
        };
        // Four agents per trip: their shared-memory loads and fp64 additions are independent, so the
        // chains overlap; the first-max update then runs in table order.
#define ROUTE_TAKE(K, S, A)                                              \
        {                                                                \
            const bool upd = (K) & ((best_a < 0) | ((S) > best));        \
            best = upd ? (S) : best;                                     \
            best_a = upd ? (A) : best_a;                                 \
        }
        int a = 0;
        for (; a + 4 <= nc; a += 4) {
            double s0, s1, s2, s3;
            const bool k0 = eval(a, s0), k1 = eval(a + 1, s1), k2 = eval(a + 2, s2), k3 = eval(a + 3, s3);
            ROUTE_TAKE(k0, s0, a0 + a)
            ROUTE_TAKE(k1, s1, a0 + a + 1)
            ROUTE_TAKE(k2, s2, a0 + a + 2)
            ROUTE_TAKE(k3, s3, a0 + a + 3)
        }
        for (; a < nc; ++a) {
            double s0;
            const bool k0 = eval(a, s0);
            ROUTE_TAKE(k0, s0, a0 + a)
        }
#undef ROUTE_TAKE
    }
    if (live) {
        out_agent[t] = best_a;
        out_score[t] = best_a >= 0 ? best : __longlong_as_double(0x7ff8000000000000ll);
    }
}

cudaError_t launch_route(const uint64_t* task_req, const uint64_t* task_tags,
                         const int32_t* task_type, const uint8_t* task_flags, int64_t nt,
                         const uint64_t* agent_caps, const uint64_t* agent_spec,
                         const uint8_t* agent_flags, const double* agent_s12,
                         const double* agent_t4, int na, int words, const double* spec_u,
                         int max_tags, int mode, int32_t* out_agent, double* out_score,
                         cudaStream_t st) {
    if (nt <= 0) return cudaSuccess;
    const size_t per_agent = 8 * (size_t)route_record_words(words);
    const size_t fixed = 2 * (size_t)(max_tags + 1) * 8 + 64;
    // small chunks: 12 KB per CTA lets 5+ CTAs share an SM, and the kernel lives on hiding the
    // load -> select -> load -> add chains behind other warps
    int chunk = (int)((12 * 1024 - fixed) / per_agent);
    if (chunk < 64) chunk = (int)((96 * 1024 - fixed) / per_agent);      // very wide vocabularies
    chunk &= ~7;
    if (chunk < 8) return cudaErrorInvalidValue;            // vocabulary too wide for shared memory
    if (chunk > ((na + 7) & ~7)) chunk = (na + 7) & ~7;
    if (chunk < 8) chunk = 8;                               // na == 0: every task gets "no agent"
    const size_t smem = fixed + (size_t)chunk * per_agent;
    const unsigned grid = (unsigned)((nt + 255) / 256);
#define ROUTE_LAUNCH(WT, MD)                                                                     \
    do {                                                                                         \
        cudaError_t e = cudaFuncSetAttribute(route_kernel<WT, MD>,                               \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                             (int)smem);                                         \
        if (e != cudaSuccess) return e;                                                          \
        route_kernel<WT, MD><<<grid, 256, smem, st>>>(task_req, task_tags, task_type, task_flags, \
                                                      nt, agent_caps, agent_spec, agent_flags,   \
                                                      agent_s12, agent_t4, na, words, spec_u,    \
                                                      max_tags, chunk, out_agent, out_score);    \
    } while (0)
#define ROUTE_MODES(WT)                                                                          \
    do {                                                                                         \
        if (mode == 0) ROUTE_LAUNCH(WT, 0);                                                      \
        else if (mode == 1) ROUTE_LAUNCH(WT, 1);                                                 \
        else ROUTE_LAUNCH(WT, 2);                                                                \
    } while (0)
    if (words == 1) ROUTE_MODES(1);
    else if (words == 2) ROUTE_MODES(2);
    else ROUTE_MODES(0);
#undef ROUTE_MODES
#undef ROUTE_LAUNCH
    return cudaGetLastError();
}

// TaskRouter.get_task_priority, pilott/core/router.py:135-145
__global__ void __launch_bounds__(256)
task_priority_kernel(const uint8_t* __restrict__ urgent, const int32_t* __restrict__ complexity,
                     const int32_t* __restrict__ ndeps, int64_t nt, uint8_t* __restrict__ out) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nt;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int c = complexity[i], d = ndeps[i];
        out[i] = urgent[i] ? 3 : (c > 8 || d > 5) ? 2 : (c > 5 || d > 3) ? 1 : 0;
    }
}

cudaError_t launch_task_priority(const uint8_t* urgent, const int32_t* complexity,
                                 const int32_t* ndeps, int64_t nt, uint8_t* out, cudaStream_t st) {
    if (nt <= 0) return cudaSuccess;
    int64_t g = (nt + 255) / 256;
    if (g > 148 * 8) g = 148 * 8;
    task_priority_kernel<<<(unsigned)g, 256, 0, st>>>(urgent, complexity, ndeps, nt, out);
    return cudaGetLastError();
}

}  // namespace pilott
