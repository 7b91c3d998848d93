// store_kernels.cu -- K1 append/quantise (+inverse norm), query conversion, per-search
// scale/mask pre-pass, compaction helpers.  All HBM-bound: coalesced 16-byte accesses,
// one warp per row, warp-shuffle reduction for the sum of squares.
#include "common.cuh"
#include "internal.h"

namespace pilott {

// One warp converts one fp32 row to bf16 (RNE) with 16-byte stores, and returns the fp32
// sum of squares of the ROUNDED values (what the similarity kernel will actually see).
// Columns [dim, ld) are written as zeros.
__device__ __forceinline__ float convert_row(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                             int dim, int ld, int lane) {
    float ss = 0.f;
    const bool vec_ok = ((dim & 7) == 0) && ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
    for (int c = lane * 8; c < ld; c += 256) {
        float v[8];
        if (vec_ok && c + 8 <= dim) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(src + c));
            const float4 b = __ldg(reinterpret_cast<const float4*>(src + c + 4));
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
            v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = (c + j < dim) ? __ldg(src + c + j) : 0.f;
        }
        uint32_t packed[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat16 lo = __float2bfloat16_rn(v[2 * j]);
            const __nv_bfloat16 hi = __float2bfloat16_rn(v[2 * j + 1]);
            const float flo = __bfloat162float(lo), fhi = __bfloat162float(hi);
            ss = fmaf(flo, flo, ss);
            ss = fmaf(fhi, fhi, ss);
            packed[j] = (uint32_t)__bfloat16_as_ushort(lo) | ((uint32_t)__bfloat16_as_ushort(hi) << 16);
        }
        *reinterpret_cast<uint4*>(dst + c) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
    }
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) ss += __shfl_xor_sync(FULL, ss, j);
    return ss;
}

// K1: append n rows starting at physical slot first_slot (wrapping at capacity).
__global__ void __launch_bounds__(256)
append_kernel(const float* __restrict__ x, int64_t n, int dim, int ld, __nv_bfloat16* __restrict__ vec,
              float* __restrict__ inv_norm, int32_t* __restrict__ prio, uint64_t* __restrict__ tagmask,
              int64_t* __restrict__ expires, const int32_t* __restrict__ in_prio,
              const uint64_t* __restrict__ in_tag, const int64_t* __restrict__ in_exp,
              int64_t first_slot, int64_t capacity) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t i = warp; i < n; i += nwarps) {
        int64_t slot = first_slot + i;
        if (slot >= capacity) slot -= capacity;
        const float ss = convert_row(x + i * dim, vec + slot * ld, dim, ld, lane);
        if (lane == 0) {
            inv_norm[slot] = ss > 0.f ? 1.0f / sqrtf(ss) : 0.f;
            prio[slot] = in_prio ? in_prio[i] : 0;
            tagmask[slot] = in_tag ? in_tag[i] : 0ull;
            expires[slot] = in_exp ? in_exp[i] : 0ll;
        }
    }
}

__global__ void __launch_bounds__(256)
convert_queries_kernel(const float* __restrict__ q, int64_t nq, int dim, int ld,
                       __nv_bfloat16* __restrict__ qb) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t i = warp; i < nq; i += nwarps) convert_row(q + i * dim, qb + i * ld, dim, ld, lane);
}

// Per-search pre-pass: scale[n] = inverse norm of a candidate row, NaN for rows the
// filter rejects (enhanced_memory.py:118-131 priority/tag filter, :20-21 expiry) and for
// slots past the end.  NaN makes `score >= threshold` false in the scan for free.
// One block = one 256-row tile per iteration; tile_max[t] = largest valid scale of tile t
// (-inf when the whole tile is filtered): the scan's fast path bounds a whole chunk of
// scores with it and only loads per-row scales on the rare hit.
__global__ void __launch_bounds__(256)
scale_kernel(const float* __restrict__ inv_norm, const int32_t* __restrict__ prio,
             const uint64_t* __restrict__ tagmask, const int64_t* __restrict__ expires,
             int64_t nslots, int64_t cap_pad, int32_t min_prio, uint64_t req, int64_t now_us,
             float* __restrict__ scale, float* __restrict__ tile_max) {
    __shared__ float s_w[8];
    const float qnan = __int_as_float(0x7fc00000);
    const int64_t ntiles = cap_pad / 256;
    for (int64_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const int64_t n = t * 256 + threadIdx.x;
        float s = qnan;
        if (n < nslots) {
            const int64_t e = expires[n];
            const bool ok = prio[n] >= min_prio && (tagmask[n] & req) == req && !(e != 0 && now_us > e);
            if (ok) s = inv_norm[n];
        }
        scale[n] = s;
        float m = fmaxf(-INFINITY, s);                    // fmaxf ignores NaN
#pragma unroll
        for (int j = 16; j > 0; j >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, j));
        if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = m;
        __syncthreads();
        if (threadIdx.x == 0) {
            float mm = s_w[0];
#pragma unroll
            for (int w = 1; w < 8; ++w) mm = fmaxf(mm, s_w[w]);
            tile_max[t] = mm;
        }
        __syncthreads();
    }
}

// One 256-row tile of the scale pre-pass by ONE warp (8 rows per lane, 16-byte loads when the
// whole tile lies inside the columns), same results as scale_kernel.
__device__ __forceinline__ void scale_tile_warp(const float* __restrict__ inv_norm, const int32_t* __restrict__ prio,
                                                const uint64_t* __restrict__ tagmask, const int64_t* __restrict__ expires,
                                                int64_t nslots, int64_t capacity, int32_t min_prio, uint64_t req,
                                                int64_t now_us, float* __restrict__ scale, float* __restrict__ tile_max,
                                                int64_t t, int lane) {
    const float qnan = __int_as_float(0x7fc00000);
    const int64_t n0 = t * 256 + lane * 8;
    float s[8];
    if (t * 256 + 256 <= capacity) {
        const float4 i0 = __ldg(reinterpret_cast<const float4*>(inv_norm + n0)), i1 = __ldg(reinterpret_cast<const float4*>(inv_norm + n0 + 4));
        const int4 p0 = __ldg(reinterpret_cast<const int4*>(prio + n0)), p1 = __ldg(reinterpret_cast<const int4*>(prio + n0 + 4));
        const float iv[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
        const int pv[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const ulonglong2 tg = __ldg(reinterpret_cast<const ulonglong2*>(tagmask + n0 + 2 * j));
            const longlong2 ex = __ldg(reinterpret_cast<const longlong2*>(expires + n0 + 2 * j));
            const uint64_t tv[2] = {tg.x, tg.y};
            const int64_t ev[2] = {ex.x, ex.y};
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const int e = 2 * j + u;
                const bool ok = n0 + e < nslots && pv[e] >= min_prio && (tv[u] & req) == req && !(ev[u] != 0 && now_us > ev[u]);
                s[e] = ok ? iv[e] : qnan;
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int64_t n = n0 + e;
            s[e] = qnan;
            if (n < nslots) {
                const int64_t ex = expires[n];
                if (prio[n] >= min_prio && (tagmask[n] & req) == req && !(ex != 0 && now_us > ex)) s[e] = inv_norm[n];
            }
        }
    }
    *reinterpret_cast<float4*>(scale + n0) = make_float4(s[0], s[1], s[2], s[3]);
    *reinterpret_cast<float4*>(scale + n0 + 4) = make_float4(s[4], s[5], s[6], s[7]);
    float m = -INFINITY;
#pragma unroll
    for (int e = 0; e < 8; ++e) m = fmaxf(m, s[e]);                 // fmaxf ignores NaN
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) m = fmaxf(m, __shfl_xor_sync(FULL, m, j));
    if (lane == 0) tile_max[t] = m;
}

// Everything a fused search needs before the scan, in ONE launch: query conversion (one warp per
// query row), the scale pre-pass (one warp per 256-row tile) and the zeroing of the per-search
// words (shared thresholds, first-tile list maxima, the inexact counter).
__global__ void __launch_bounds__(256)
prepare_kernel(const float* __restrict__ q, int64_t nq, int dim, int ld, __nv_bfloat16* __restrict__ qb,
               const float* __restrict__ inv_norm, const int32_t* __restrict__ prio,
               const uint64_t* __restrict__ tagmask, const int64_t* __restrict__ expires, int64_t nslots,
               int64_t capacity, int64_t cap_pad, int32_t min_prio, uint64_t req, int64_t now_us,
               float* __restrict__ scale, float* __restrict__ tile_max, uint32_t* __restrict__ zero_words,
               int64_t n_zero_words, int32_t* __restrict__ inexact, int do_scale) {
    pdl_trigger();                                                  // the scan may start its prologue now ...
    pdl_wait();                                                     // ... and this grid waits for whatever still reads the buffers it rewrites
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    const int64_t ntiles = do_scale ? cap_pad / 256 : 0;            // 0: scale / tile_max of these filters are still valid
    const int64_t nzero = (n_zero_words + 127) / 128;               // 128 words (512 B) per item
    const int64_t items = nq + ntiles + nzero;
    if (warp == 0 && lane == 0 && inexact) *inexact = 0;
    for (int64_t it = warp; it < items; it += nwarps) {
        if (it < nq) {
            convert_row(q + it * dim, qb + it * ld, dim, ld, lane);
        } else if (it < nq + ntiles) {
            scale_tile_warp(inv_norm, prio, tagmask, expires, nslots, capacity, min_prio, req, now_us, scale, tile_max,
                            it - nq, lane);
        } else {
            const int64_t w0 = (it - nq - ntiles) * 128 + lane * 4;
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (w0 + j < n_zero_words) zero_words[w0 + j] = 0u;
        }
    }
}

// keep[i] for the i-th oldest held row (age order): not expired at now_us.
__global__ void __launch_bounds__(256)
compact_flags_kernel(const int64_t* __restrict__ expires, int64_t count, int64_t head,
                     int64_t capacity, int64_t now_us, uint8_t* __restrict__ keep) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count;
         i += (int64_t)gridDim.x * blockDim.x) {
        int64_t slot = head + i;
        if (slot >= capacity) slot -= capacity;
        const int64_t e = expires[slot];
        keep[i] = !(e != 0 && now_us > e);
    }
}

// Gather rows src_slot[i] -> dense destination row i (one warp per row, 16-byte copies).
__global__ void __launch_bounds__(256)
gather_rows_kernel(const __nv_bfloat16* __restrict__ vec, const float* __restrict__ inv_norm,
                   const int32_t* __restrict__ prio, const uint64_t* __restrict__ tagmask,
                   const int64_t* __restrict__ expires, int ld, const int64_t* __restrict__ src_slot,
                   int64_t n, __nv_bfloat16* __restrict__ dvec, float* __restrict__ dinv,
                   int32_t* __restrict__ dprio, uint64_t* __restrict__ dtag, int64_t* __restrict__ dexp) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t i = warp; i < n; i += nwarps) {
        const int64_t s = src_slot[i];
        const uint4* src = reinterpret_cast<const uint4*>(vec + s * ld);
        uint4* dst = reinterpret_cast<uint4*>(dvec + i * ld);
        for (int c = lane; c < ld / 8; c += 32) dst[c] = src[c];
        if (lane == 0) {
            dinv[i] = inv_norm[s]; dprio[i] = prio[s]; dtag[i] = tagmask[s]; dexp[i] = expires[s];
        }
    }
}

static inline int grid_for(int64_t work_items, int per_block, int cap) {
    int64_t g = (work_items + per_block - 1) / per_block;
    if (g < 1) g = 1;
    if (g > cap) g = cap;
    return (int)g;
}

cudaError_t launch_append(const float* x, int64_t n, int dim, int ld, void* vec, float* inv_norm,
                          int32_t* prio, uint64_t* tagmask, int64_t* expires,
                          const int32_t* in_prio, const uint64_t* in_tag, const int64_t* in_exp,
                          int64_t first_slot, int64_t capacity, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    append_kernel<<<grid_for(n, 8, 148 * 8), 256, 0, st>>>(
        x, n, dim, ld, (__nv_bfloat16*)vec, inv_norm, prio, tagmask, expires, in_prio, in_tag,
        in_exp, first_slot, capacity);
    return cudaGetLastError();
}

cudaError_t launch_convert_queries(const float* q, int64_t nq, int dim, int ld, void* qb,
                                   cudaStream_t st) {
    if (nq <= 0) return cudaSuccess;
    convert_queries_kernel<<<grid_for(nq, 8, 148 * 8), 256, 0, st>>>(q, nq, dim, ld,
                                                                    (__nv_bfloat16*)qb);
    return cudaGetLastError();
}

cudaError_t launch_scale(const float* inv_norm, const int32_t* prio, const uint64_t* tagmask,
                         const int64_t* expires, int64_t nslots, int64_t cap_pad, int32_t min_prio,
                         uint64_t req, int64_t now_us, float* scale, float* tile_max, cudaStream_t st) {
    scale_kernel<<<grid_for(cap_pad, 256, 148 * 8), 256, 0, st>>>(
        inv_norm, prio, tagmask, expires, nslots, cap_pad, min_prio, req, now_us, scale, tile_max);
    return cudaGetLastError();
}

cudaError_t launch_prepare(const float* q, int64_t nq, int dim, int ld, void* qb, const float* inv_norm,
                           const int32_t* prio, const uint64_t* tagmask, const int64_t* expires, int64_t nslots,
                           int64_t capacity, int64_t cap_pad, int32_t min_prio, uint64_t req, int64_t now_us,
                           float* scale, float* tile_max, uint32_t* zero_words, int64_t n_zero_words,
                           int32_t* inexact, bool do_scale, cudaStream_t st) {
    const int64_t items = nq + (do_scale ? cap_pad / 256 : 0) + (n_zero_words + 127) / 128;
    return launch_pdl(prepare_kernel, dim3(grid_for(items, 8, 148 * 8)), dim3(256), 0, st,
                      q, nq, dim, ld, (__nv_bfloat16*)qb, inv_norm, prio, tagmask, expires, nslots, capacity, cap_pad, min_prio, req,
                      now_us, scale, tile_max, zero_words, n_zero_words, inexact, do_scale ? 1 : 0);
}

cudaError_t launch_compact_flags(const int64_t* expires, int64_t count, int64_t head,
                                 int64_t capacity, int64_t now_us, uint8_t* keep, cudaStream_t st) {
    if (count <= 0) return cudaSuccess;
    compact_flags_kernel<<<grid_for(count, 256, 148 * 8), 256, 0, st>>>(expires, count, head,
                                                                        capacity, now_us, keep);
    return cudaGetLastError();
}

cudaError_t launch_gather_rows(const void* vec, const float* inv_norm, const int32_t* prio,
                               const uint64_t* tagmask, const int64_t* expires, int ld,
                               const int64_t* src_slot, int64_t n, void* dvec, float* dinv,
                               int32_t* dprio, uint64_t* dtag, int64_t* dexp, cudaStream_t st) {
    if (n <= 0) return cudaSuccess;
    gather_rows_kernel<<<grid_for(n, 8, 148 * 8), 256, 0, st>>>(
        (const __nv_bfloat16*)vec, inv_norm, prio, tagmask, expires, ld, src_slot, n,
        (__nv_bfloat16*)dvec, dinv, dprio, dtag, dexp);
    return cudaGetLastError();
}

}  // namespace pilott
