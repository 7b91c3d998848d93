// route_tc.cu -- K5 in the form the north star names: task x agent matching as a GEMM on the
// tcgen05 tensor cores, score assembly + first-max argmax in the epilogue that reads TMEM.
//
// route.cu does the matching as set arithmetic on 64-bit words, one thread per task, every agent
// visited by every thread (about 33 instructions per pair).  Here the three set sizes the scorers
// need come out of ONE bf16 GEMM whose fp32 result is an exact small integer:
//   D[task, agent] = sum_f A[task, f] * B[agent, f]
//   f in [0, C)        A = 1    if the task requires capability f,   B = 1 if the agent has it
//   f in [C, C+S)      A = 512  if the task's type is f - C,         B = 1 if the agent specialises in it
//   f in [C+S, C+2S)   A = 1024 if the task carries tag f - C - S,   B = 1 if the agent specialises in it
//   f  = C+2S          A = 256  if the task has the 'required_capabilities' key,
//                                                                    B = 1 if the agent LACKS that attribute
// (C = S = 64 * words).  All values are powers of two, exact in bf16, and the sum stays below 2^18,
// exact in fp32 whatever the order:  bits [0,8) = required capabilities the agent has, bit 8 = the
// attribute is missing, bit 9 = type match, bits [10,18) = tag matches.  The epilogue thread that
// owns a task row (TMEM lane) unpacks them and does exactly what route.cu's eval() does:
// base selector (pilott/core/agent.py:551-564), per-agent fp64 terms precomputed on the host in the
// reference's operation order, at most two IEEE fp64 additions (pilott/core/router.py:83-88), first
// maximum in table order (router.py:62, load_balancer.py:285-287, task_delegator.py:82).
// Results are bit-identical to route.cu's (tests/test_gpu_routing.py runs both).
//
// Structure (one CTA per SM, persistent over 128-task tiles; 384 threads as in scan_tc.cu):
// warp 0 TMA producer, warp 1 MMA issuer, warp 2 TMEM allocator, warps 4..11 epilogue in two sets
// (set s owns columns [128 s, 128 s + 128) of every 256-agent tile); smem ring of (A k-block,
// B k-block) stages, two TMEM accumulator stages.  Agents live in TMEM columns, so "first maximum"
// is a strict > while walking columns upward, and the two sets are combined at the end of a task
// tile by (score desc, agent index asc).
#include "common.cuh"
#include "internal.h"
#include "tc_ptx.cuh"

namespace pilott {
namespace rtc {

constexpr int BM = 128, BN = 256, BK = 64, UK = 16;
constexpr int ACC_STAGES = 2, STAGES = 3, SETS = 2, THREADS = 128 + SETS * 128;
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE_BYTES = A_BYTES + B_BYTES;
constexpr uint16_t BF_1 = 0x3F80, BF_256 = 0x4380, BF_512 = 0x4400, BF_1024 = 0x4480;

__host__ __device__ inline int feat_cols(int words) { return (192 * words + 1 + 63) & ~63; }

__device__ __forceinline__ bool bit_of(const uint64_t* __restrict__ w, int b) { return (w[b >> 6] >> (b & 63)) & 1ull; }

// One warp per row, 16-byte stores.  Rows past `n` are zeros (tile padding).
__global__ void __launch_bounds__(256)
expand_tasks_kernel(const uint64_t* __restrict__ req, const uint64_t* __restrict__ tags, const int32_t* __restrict__ type,
                    const uint8_t* __restrict__ flags, int64_t nt, int64_t nt_pad, int words, int kp,
                    uint16_t* __restrict__ A) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= nt_pad) return;
    const int C = 64 * words;
    const bool live = row < nt;
    const uint8_t tf = live ? flags[row] : 0;
    const bool has_req = tf & 1;
    const int ty = live ? type[row] : -1;
    for (int c = lane * 8; c < kp; c += 256) {
        uint16_t v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int e = c + j;
            uint16_t x = 0;
            if (live) {
                if (e < C) x = (has_req && bit_of(req + row * words, e)) ? BF_1 : 0;
                else if (e < 2 * C) x = (ty == e - C) ? BF_512 : 0;
                else if (e < 3 * C) x = bit_of(tags + row * words, e - 2 * C) ? BF_1024 : 0;
                else if (e == 3 * C) x = has_req ? BF_256 : 0;
            }
            v[j] = x;
        }
        *reinterpret_cast<uint4*>(A + row * kp + c) =
            make_uint4(v[0] | (uint32_t)v[1] << 16, v[2] | (uint32_t)v[3] << 16, v[4] | (uint32_t)v[5] << 16, v[6] | (uint32_t)v[7] << 16);
    }
}

__global__ void __launch_bounds__(256)
expand_agents_kernel(const uint64_t* __restrict__ caps, const uint64_t* __restrict__ spec, const uint8_t* __restrict__ flags,
                     int na, int na_pad, int words, int kp, uint16_t* __restrict__ B) {
    const int lane = threadIdx.x & 31;
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= na_pad) return;
    const int C = 64 * words;
    const bool live = row < na;
    const bool lacks_attr = live && !(flags[row] & 2);
    for (int c = lane * 8; c < kp; c += 256) {
        uint16_t v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int e = c + j;
            uint16_t x = 0;
            if (live) {
                if (e < C) x = bit_of(caps + (int64_t)row * words, e) ? BF_1 : 0;
                else if (e < 2 * C) x = bit_of(spec + (int64_t)row * words, e - C) ? BF_1 : 0;
                else if (e < 3 * C) x = bit_of(spec + (int64_t)row * words, e - 2 * C) ? BF_1 : 0;
                else if (e == 3 * C) x = lacks_attr ? BF_1 : 0;
            }
            v[j] = x;
        }
        *reinterpret_cast<uint4*>(B + (int64_t)row * kp + c) =
            make_uint4(v[0] | (uint32_t)v[1] << 16, v[2] | (uint32_t)v[3] << 16, v[4] | (uint32_t)v[5] << 16, v[6] | (uint32_t)v[7] << 16);
    }
}


// Per-agent terms in shared memory: s12[0..2], t4 (32 bytes).  Agents that cannot win (not viable,
// tile padding) carry s12 = -inf, so their score can never pass the `>` tests below.
struct AgentTerms { double s12[3]; double t4; };

template <int MODE>
__global__ void __launch_bounds__(THREADS, 1)
route_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                const uint64_t* __restrict__ task_req, const uint8_t* __restrict__ task_flags, int64_t nt,
                const uint8_t* __restrict__ agent_flags, const double* __restrict__ agent_s12,
                const double* __restrict__ agent_t4, int na, int na_pad, int words, int kblocks,
                const double* __restrict__ spec_u, int max_tags, int32_t* __restrict__ out_agent,
                double* __restrict__ out_score) {
    extern __shared__ unsigned char smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;        // 1024-B aligned (swizzle atom)
    unsigned char* base_ptr = smem_dyn + (base - smem_u32(smem_dyn));
    AgentTerms* s_lut = reinterpret_cast<AgentTerms*>(base_ptr + STAGES * STAGE_BYTES);            // [na_pad]
    double* s_u = reinterpret_cast<double*>(s_lut + na_pad);                                       // [2][max_tags + 1]
    double* s_xs = s_u + 2 * (max_tags + 1);                                                       // [BM] set 1's best score
    int* s_xa = reinterpret_cast<int*>(s_xs + BM);                                                 // [BM] ... and agent
    const size_t bar_off = (size_t)STAGES * STAGE_BYTES + (size_t)na_pad * sizeof(AgentTerms) +
                           (size_t)(2 * (max_tags + 1) + BM) * 8 + BM * 4;
    const uint32_t bar_base = base + (uint32_t)((bar_off + 15) & ~(size_t)15);
    const uint32_t bar_full = bar_base, bar_empty = bar_base + 8 * STAGES;
    const uint32_t bar_tfull = bar_base + 16 * STAGES, bar_tempty = bar_tfull + 8 * ACC_STAGES;
    const uint32_t tmem_slot = bar_tempty + 8 * ACC_STAGES;
    volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(base_ptr + (tmem_slot - base));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t ntt = (nt + BM - 1) / BM;                              // task tiles
    const int nat = na_pad / BN;                                         // agent tiles

    for (int a = threadIdx.x; a < na_pad; a += THREADS) {
        const bool ok = a < na && (agent_flags[a] & 1);
        AgentTerms t;
        for (int j = 0; j < 3; ++j) t.s12[j] = ok ? agent_s12[(size_t)a * 3 + j] : -INFINITY;
        t.t4 = ok ? agent_t4[a] : 0.0;
        s_lut[a] = t;
    }
    for (int i = threadIdx.x; i < 2 * (max_tags + 1); i += THREADS) s_u[i] = spec_u[i];
    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(bar_full + 8 * i, 1); mbar_init(bar_empty + 8 * i, 1); }
        for (int i = 0; i < ACC_STAGES; ++i) { mbar_init(bar_tfull + 8 * i, 1); mbar_init(bar_tempty + 8 * i, 4 * SETS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) tmem_alloc<1>(tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot_ptr;

    if (warp == 0) {
        // ===== TMA producer: (A k-block of the task tile, B k-block of the agent tile) per stage =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int64_t tt = blockIdx.x; tt < ntt; tt += gridDim.x)
                for (int at = 0; at < nat; ++at)
                    for (int kb = 0; kb < kblocks; ++kb) {
                        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                        const uint32_t sa = base + stage * STAGE_BYTES;
                        mbar_arrive_expect_tx(bar_full + 8 * stage, STAGE_BYTES);
                        tma_load_2d(sa, &map_a, kb * BK, (int)(tt * BM), bar_full + 8 * stage);
                        tma_load_2d(sa + A_BYTES, &map_b, kb * BK, at * BN, bar_full + 8 * stage);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
        }
        __syncwarp();
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            int as = 0; uint32_t aphase = 0;
            for (int64_t tt = blockIdx.x; tt < ntt; tt += gridDim.x)
                for (int at = 0; at < nat; ++at) {
                    mbar_wait(bar_tempty + 8 * as, aphase ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + as * BN;
                    for (int kb = 0; kb < kblocks; ++kb) {
                        mbar_wait(bar_full + 8 * stage, phase);
                        tc_fence_after();
                        const uint32_t sa = base + stage * STAGE_BYTES;
                        const uint64_t adesc = make_desc(sa), bdesc = make_desc(sa + A_BYTES);
#pragma unroll
                        for (int k = 0; k < BK / UK; ++k)
                            umma_bf16<1>(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        umma_commit<1>(bar_empty + 8 * stage);
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                    }
                    umma_commit<1>(bar_tfull + 8 * as);
                    if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
                }
        }
        __syncwarp();
    } else if (warp >= 4) {
        // ===== epilogue: unpack the counts, assemble the score, keep the first maximum =====
        const int set = (warp - 4) >> 2, quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const int u_hi = max_tags + 1;
        int as = 0; uint32_t aphase = 0;
        for (int64_t tt = blockIdx.x; tt < ntt; tt += gridDim.x) {
            const int64_t t = tt * BM + row;
            const bool live = t < nt;
            const uint8_t tf = live ? task_flags[t] : 0;
            const bool has_req = tf & 1, has_type = tf & 2, unknown = tf & 4;
            int nreq = 0;
            if (live && has_req)
                for (int w = 0; w < words; ++w) nreq += __popcll(task_req[t * words + w]);
            double best = 0.0;
            int best_a = -1;
            for (int at = 0; at < nat; ++at) {
                mbar_wait(bar_tfull + 8 * as, aphase);
                tc_fence_after();
                const uint32_t taddr0 = tmem_base + ((uint32_t)(quarter * 32) << 16) + as * BN + set * (BN / SETS);
                const int a0 = at * BN + set * (BN / SETS);
                uint32_t va[16], vb[16];
                tmem_ld16_issue(taddr0, va);
                auto process = [&](const uint32_t (&v)[16], const int c) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int x = __float2int_rn(__uint_as_float(v[j]));              // exact small integer
                        const bool ty = (x >> 9) & 1;
                        const bool zero_base = ((x & 0xff) != nreq) | (((x >> 8) & 1) != 0) | unknown;
                        const int sel = zero_base ? 0 : ((has_type & ty) ? 2 : 1);        // agent.py:551-564
                        const AgentTerms* ag = s_lut + a0 + c + j;
                        double sc = ag->s12[sel];
                        if (MODE == 0) sc = __dadd_rn(__dadd_rn(sc, s_u[(ty ? u_hi : 0) + ((x >> 10) & 0xff)]), ag->t4);   // router.py:83-88
                        const bool ok = MODE == 2 ? sc > 0.0 : sc > -INFINITY;            // -inf: agent not viable / padding
                        const bool upd = ok & ((best_a < 0) | (sc > best));               // strict >: the first maximum stays
                        best = upd ? sc : best;
                        best_a = upd ? a0 + c + j : best_a;
                    }
                };
#pragma unroll 1
                for (int c = 0; c < BN / SETS; c += 32) {
                    tmem_ld_wait(va);
                    tmem_ld16_issue(taddr0 + c + 16, vb);
                    process(va, c);
                    tmem_ld_wait(vb);
                    if (c + 32 < BN / SETS) tmem_ld16_issue(taddr0 + c + 32, va);
                    process(vb, c + 16);
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
                if (++as == ACC_STAGES) { as = 0; aphase ^= 1; }
            }
            // combine the two column halves: higher score wins, equal scores go to the lower agent index
            if (set == 1) { s_xs[row] = best; s_xa[row] = best_a; }
            named_bar_sync(1, SETS * 128);
            if (set == 0) {
                const double os = s_xs[row];
                const int oa = s_xa[row];
                if (oa >= 0 && (best_a < 0 || os > best || (os == best && oa < best_a))) { best = os; best_a = oa; }
                if (live) {
                    out_agent[t] = best_a;
                    out_score[t] = best_a >= 0 ? best : __longlong_as_double(0x7ff8000000000000ll);
                }
            }
            named_bar_sync(1, SETS * 128);                               // s_xs / s_xa free for the next task tile
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        tmem_dealloc<1>(tmem_base, 512);
    }
}

static size_t route_tc_smem(int na_pad, int max_tags) {
    return 1024 + (size_t)STAGES * STAGE_BYTES + (size_t)na_pad * sizeof(AgentTerms) + (size_t)(2 * (max_tags + 1) + BM) * 8 + BM * 4 +
           16 + 8 * (2 * STAGES + 2 * ACC_STAGES) + 16;
}

}  // namespace rtc

int64_t route_tc_workspace_bytes(int64_t nt, int na, int words) {
    const int kp = rtc::feat_cols(words);
    const int64_t nt_pad = (nt + rtc::BM - 1) / rtc::BM * rtc::BM;
    const int64_t na_pad = ((int64_t)na + rtc::BN - 1) / rtc::BN * rtc::BN;
    return (nt_pad + (na_pad > 0 ? na_pad : rtc::BN)) * kp * 2 + 256;
}

bool route_tc_supported(int na, int words, int max_tags) {
    if (words < 1 || words > 3 || na < 1) return false;
    const int na_pad = (na + rtc::BN - 1) / rtc::BN * rtc::BN;
    return rtc::route_tc_smem(na_pad, max_tags) <= 227 * 1024;
}

// Expands the bit sets into the two bf16 feature matrices (workspace), then runs the GEMM + epilogue.
// encode(map, ptr, rows, row_elems, box_rows) is the caller's tensor-map encoder (abi.cu).
cudaError_t launch_route_tc(const uint64_t* task_req, const uint64_t* task_tags, const int32_t* task_type,
                            const uint8_t* task_flags, int64_t nt, const uint64_t* agent_caps, const uint64_t* agent_spec,
                            const uint8_t* agent_flags, const double* agent_s12, const double* agent_t4, int na, int words,
                            const double* spec_u, int max_tags, int mode, int32_t* out_agent, double* out_score,
                            void* workspace, int num_sms, int (*encode)(CUtensorMap*, const void*, int64_t, int, int),
                            cudaStream_t st) {
    using namespace rtc;
    if (nt <= 0) return cudaSuccess;
    const int kp = feat_cols(words);
    const int64_t nt_pad = (nt + BM - 1) / BM * BM;
    const int na_pad = (na + BN - 1) / BN * BN;
    uint16_t* A = reinterpret_cast<uint16_t*>(workspace);
    uint16_t* B = A + nt_pad * kp;
    expand_tasks_kernel<<<(unsigned)((nt_pad + 7) / 8), 256, 0, st>>>(task_req, task_tags, task_type, task_flags, nt, nt_pad, words, kp, A);
    expand_agents_kernel<<<(unsigned)((na_pad + 7) / 8), 256, 0, st>>>(agent_caps, agent_spec, agent_flags, na, na_pad, words, kp, B);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    CUtensorMap map_a, map_b;
    if (encode(&map_a, A, nt_pad, kp, BM) != 0 || encode(&map_b, B, na_pad, kp, BN) != 0) return cudaErrorInvalidValue;
    const size_t smem = route_tc_smem(na_pad, max_tags);
    const int64_t ntt = nt_pad / BM;
    const unsigned grid = (unsigned)(ntt < num_sms ? ntt : num_sms);
#define RTC_LAUNCH(MD)                                                                                                    \
    do {                                                                                                                  \
        e = cudaFuncSetAttribute(route_tc_kernel<MD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);             \
        if (e != cudaSuccess) return e;                                                                                   \
        route_tc_kernel<MD><<<grid, THREADS, smem, st>>>(map_a, map_b, task_req, task_flags, nt, agent_flags, agent_s12,  \
                                                         agent_t4, na, na_pad, words, kp / BK, spec_u, max_tags, out_agent, \
                                                         out_score);                                                      \
    } while (0)
    if (mode == 0) RTC_LAUNCH(0);
    else if (mode == 1) RTC_LAUNCH(1);
    else RTC_LAUNCH(2);
#undef RTC_LAUNCH
    return cudaGetLastError();
}

}  // namespace pilott
