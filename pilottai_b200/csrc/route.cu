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

template <int W>
__global__ void __launch_bounds__(256)
route_kernel(const uint64_t* __restrict__ task_req, const uint64_t* __restrict__ task_tags,
             const int32_t* __restrict__ task_type, const uint8_t* __restrict__ task_flags, int64_t nt,
             const uint64_t* __restrict__ agent_caps, const uint64_t* __restrict__ agent_spec,
             const uint8_t* __restrict__ agent_flags, const double* __restrict__ agent_s12,
             const double* __restrict__ agent_t4, int na, int words, const double* __restrict__ spec_u,
             int max_tags, int mode, int chunk, int32_t* __restrict__ out_agent,
             double* __restrict__ out_score) {
    extern __shared__ __align__(16) unsigned char sm_raw[];
    const int nw = W > 0 ? W : words;
    double* s_s12 = reinterpret_cast<double*>(sm_raw);                 // [chunk][3]
    double* s_t4 = s_s12 + (size_t)chunk * 3;                          // [chunk]
    double* s_u = s_t4 + chunk;                                        // [2*(max_tags+1)]
    uint64_t* s_caps = reinterpret_cast<uint64_t*>(s_u + 2 * (max_tags + 1));   // [chunk][nw]
    uint64_t* s_spec = s_caps + (size_t)chunk * nw;                    // [chunk][nw]
    uint8_t* s_flags = reinterpret_cast<uint8_t*>(s_spec + (size_t)chunk * nw);  // [chunk]

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
    double best = 0.0;
    int best_a = -1;

    for (int a0 = 0; a0 < na; a0 += chunk) {
        const int nc = min(chunk, na - a0);
        __syncthreads();
        for (int i = threadIdx.x; i < nc * 3; i += blockDim.x) s_s12[i] = agent_s12[(size_t)a0 * 3 + i];
        for (int i = threadIdx.x; i < nc; i += blockDim.x) { s_t4[i] = agent_t4[a0 + i]; s_flags[i] = agent_flags[a0 + i]; }
        for (int i = threadIdx.x; i < nc * nw; i += blockDim.x) {
            s_caps[i] = agent_caps[(size_t)a0 * nw + i];
            s_spec[i] = agent_spec[(size_t)a0 * nw + i];
        }
        __syncthreads();
        if (!live) continue;
        for (int a = 0; a < nc; ++a) {
            const uint8_t f = s_flags[a];
            if (!(f & 1)) continue;                                    // not viable / skipped
            bool missing;
            int tagc = 0;
            bool tbit = false;
            if (W > 0) {
                uint64_t miss = 0;
#pragma unroll
                for (int w = 0; w < (W > 0 ? W : 1); ++w) {
                    miss |= req[w] & ~s_caps[a * nw + w];
                    tagc += __popcll(tags[w] & s_spec[a * nw + w]);
                }
                missing = miss != 0;
                if (type >= 0) tbit = (s_spec[a * nw + (type >> 6)] >> (type & 63)) & 1;
            } else {
                uint64_t miss = 0;
                for (int w = 0; w < nw; ++w) {
                    miss |= task_req[t * nw + w] & ~s_caps[a * nw + w];
                    tagc += __popcll(task_tags[t * nw + w] & s_spec[a * nw + w]);
                }
                missing = miss != 0;
                if (type >= 0) tbit = (s_spec[a * nw + (type >> 6)] >> (type & 63)) & 1;
            }
            // agent.py:551-556: key present and (attribute missing -> exception -> 0.0, or caps missing)
            const bool zero_base = has_req_key && (!(f & 2) || missing || req_unknown);
            const int sel = zero_base ? 0 : ((has_type_key && tbit) ? 2 : 1);   // agent.py:562-564
            double sc = s_s12[a * 3 + sel];
            if (mode == 0) {
                const double u = s_u[(tbit ? (max_tags + 1) : 0) + tagc];      // router.py:116-126
                sc = __dadd_rn(__dadd_rn(sc, u), s_t4[a]);
            }
            if (mode == 2 && !(sc > 0.0)) continue;                    // task_delegator.py:75
            if (best_a < 0 || sc > best) { best = sc; best_a = a0 + a; }
        }
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
    const size_t per_agent = 3 * 8 + 8 + 2 * 8 * (size_t)words + 1;
    const size_t fixed = 2 * (size_t)(max_tags + 1) * 8 + 64;
    int chunk = (int)((96 * 1024 - fixed) / per_agent);
    chunk &= ~7;                                   // keeps the uint64 sub-arrays 8-byte aligned
    if (chunk < 8) return cudaErrorInvalidValue;            // vocabulary too wide for shared memory
    if (chunk > ((na + 7) & ~7)) chunk = (na + 7) & ~7;
    if (chunk < 8) chunk = 8;                               // na == 0: every task gets "no agent"
    const size_t smem = fixed + (size_t)chunk * per_agent;
    const unsigned grid = (unsigned)((nt + 255) / 256);
#define ROUTE_LAUNCH(WT)                                                                         \
    do {                                                                                         \
        cudaError_t e = cudaFuncSetAttribute(route_kernel<WT>,                                   \
                                             cudaFuncAttributeMaxDynamicSharedMemorySize,        \
                                             (int)smem);                                         \
        if (e != cudaSuccess) return e;                                                          \
        route_kernel<WT><<<grid, 256, smem, st>>>(task_req, task_tags, task_type, task_flags, nt, \
                                                  agent_caps, agent_spec, agent_flags, agent_s12, \
                                                  agent_t4, na, words, spec_u, max_tags, mode,   \
                                                  chunk, out_agent, out_score);                  \
    } while (0)
    if (words == 1) ROUTE_LAUNCH(1);
    else if (words == 2) ROUTE_LAUNCH(2);
    else ROUTE_LAUNCH(0);
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
