// common.cuh -- shared device helpers for the pilott_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pilott {

constexpr int KP = 32;    // candidates kept per (query, slot) by the approximate pass
constexpr int CAP = 64;   // per-row append buffer (KP survivors + pending appends)
constexpr unsigned FULL = 0xffffffffu;

// ---- programmatic dependent launch ----------------------------------------------------------
// The kernels of one search step (prepare -> scan -> finish [-> merge]) are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization: each lets its successor start launching at
// once (pdl_trigger) and waits for its predecessor to have completed, memory flushed, before it
// touches anything the predecessor wrote or still reads (pdl_wait).  Launch latency and the
// prologue of kernel i+1 then overlap the tail of kernel i.  Both instructions are no-ops in a
// kernel launched without the attribute.  Every kernel launched with it MUST call pdl_wait in every
// CTA before its first global access: completion is transitive only through the waits.
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// ---- order-preserving float <-> uint32, and the 64-bit candidate key -------------------
// key = ord(score) << 32 | (0xFFFFFFFF - age):  larger key == better candidate
// (higher score first, then OLDER row first: the reference returns older items first
//  among equals, pilott/memory/enhanced_memory.py:114).  key 0 == empty.
__device__ __forceinline__ uint32_t f2ord(float f) {
    uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}
__device__ __forceinline__ uint64_t make_key(float s, uint32_t age) {
    return ((uint64_t)f2ord(s) << 32) | (uint64_t)(0xFFFFFFFFu - age);
}
__device__ __forceinline__ uint32_t key_age(uint64_t k) { return 0xFFFFFFFFu - (uint32_t)k; }
__device__ __forceinline__ float key_score(uint64_t k) { return ord2f((uint32_t)(k >> 32)); }

__device__ __forceinline__ uint64_t shfl_xor64(uint64_t v, int m) {
    uint32_t lo = __shfl_xor_sync(FULL, (uint32_t)v, m);
    uint32_t hi = __shfl_xor_sync(FULL, (uint32_t)(v >> 32), m);
    return ((uint64_t)hi << 32) | lo;
}
__device__ __forceinline__ uint64_t shfl64(uint64_t v, int src) {
    uint32_t lo = __shfl_sync(FULL, (uint32_t)v, src);
    uint32_t hi = __shfl_sync(FULL, (uint32_t)(v >> 32), src);
    return ((uint64_t)hi << 32) | lo;
}

// Bitonic sort of one 64-bit key per lane across the warp.  DESC: lane 0 ends largest.
template <bool DESC>
__device__ __forceinline__ uint64_t warp_sort64(uint64_t v, int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const uint64_t o = shfl_xor64(v, j);
            const bool up = (((lane & k) == 0) != DESC);
            const bool lower = (lane & j) == 0;
            const bool keep_min = (lower == up);
            const uint64_t mn = v < o ? v : o, mx = v < o ? o : v;
            v = keep_min ? mn : mx;
        }
    }
    return v;
}
__device__ __forceinline__ uint64_t warp_min64(uint64_t v) {
#pragma unroll
    for (int j = 16; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor64(v, j);
        v = v < o ? v : o;
    }
    return v;
}

// Raw candidate entry (written by the tcgen05 epilogue's append): low 32 bits = physical
// ring slot.  Converted key: low 32 bits = 0xFFFFFFFF - age.  The conversion is deferred to
// prune time so the hot loop does not pay for it.
__device__ __forceinline__ uint64_t raw_to_key(uint64_t raw, uint32_t head, uint32_t cap) {
    int32_t age = (int32_t)((uint32_t)raw - head);
    if (age < 0) age += (int32_t)cap;
    return (raw & 0xFFFFFFFF00000000ull) | (uint64_t)(0xFFFFFFFFu - (uint32_t)age);
}

// Keep the KPT (8, 16 or 32) best of a row buffer holding cnt (<= CAPT <= 64) entries.  Whole
// warp cooperates.  Entries at index >= conv are still raw.  Leaves the survivors (unsorted,
// empty = 0) in buf[0..KPT) and returns them per lane (lanes >= KPT return 0).
template <int CAPT, int KPT>
__device__ __forceinline__ uint64_t warp_prune(uint64_t* buf, int cnt, int conv, uint32_t head,
                                               uint32_t cap, int lane, int& new_cnt, float& new_thr) {
    uint64_t a = lane < cnt ? buf[lane] : 0ull;
    uint64_t b = (CAPT > 32 && (lane + 32) < cnt) ? buf[lane + 32] : 0ull;
    if (lane < cnt && lane >= conv) a = raw_to_key(a, head, cap);
    if ((lane + 32) < cnt && (lane + 32) >= conv) b = raw_to_key(b, head, cap);
    a = warp_sort64<true>(a, lane);
    uint64_t m = a;
    if (CAPT > 32) {                           // second half of the buffer exists
        b = warp_sort64<false>(b, lane);
        m = a > b ? a : b;                     // half-cleaner of a bitonic sequence: the top 32
    }
    if (KPT <= 16) {                           // one more half-cleaner step: top 16 in lanes 0..15
        const uint64_t o = shfl_xor64(m, 16);
        m = (lane < 16) ? (m > o ? m : o) : 0ull;
    }
    if (KPT == 8) {                            // and another: top 8 in lanes 0..7
        const uint64_t o = shfl_xor64(m, 8);
        m = (lane < 8) ? (m > o ? m : o) : 0ull;
    }
    uint64_t mn = m;
#pragma unroll
    for (int j = KPT / 2; j > 0; j >>= 1) {
        const uint64_t o = shfl_xor64(mn, j);
        mn = mn < o ? mn : o;
    }
    mn = shfl64(mn, 0);
    if (lane < KPT) buf[lane] = m;
    new_cnt = cnt < KPT ? cnt : KPT;
    new_thr = cnt >= KPT ? key_score(mn) : -INFINITY;
    return m;
}

// Prune every row of this warp's 32-row group whose bit is set in `mask`.
// rows = base of the warp's 32 row buffers (row r at rows + r*CAPT).
// B independent bitonic sorts, step loop outermost so the B dependency chains interleave.
template <bool DESC, int B>
__device__ __forceinline__ void warp_sort64_multi(uint64_t (&v)[B], int lane) {
#pragma unroll
    for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
        for (int j = k >> 1; j > 0; j >>= 1) {
            const bool up = (((lane & k) == 0) != DESC);
            const bool keep_min = (((lane & j) == 0) == up);
#pragma unroll
            for (int i = 0; i < B; ++i) {
                const uint64_t o = shfl_xor64(v[i], j);
                const uint64_t mn = v[i] < o ? v[i] : o, mx = v[i] < o ? o : v[i];
                v[i] = keep_min ? mn : mx;
            }
        }
    }
}

// warp_prune for B rows at once (same results; ~B times the work in about the latency of one).
template <int CAPT, int KPT, int B>
__device__ __forceinline__ void warp_prune_multi(uint64_t* const (&buf)[B], const int (&cnt)[B], const int (&conv)[B],
                                                 uint32_t head, uint32_t cap, int lane, uint64_t (&m)[B],
                                                 int (&new_cnt)[B], float (&new_thr)[B]) {
    uint64_t a[B], b[B];
#pragma unroll
    for (int i = 0; i < B; ++i) {
        a[i] = lane < cnt[i] ? buf[i][lane] : 0ull;
        b[i] = (CAPT > 32 && (lane + 32) < cnt[i]) ? buf[i][lane + 32] : 0ull;
        if (lane < cnt[i] && lane >= conv[i]) a[i] = raw_to_key(a[i], head, cap);
        if ((lane + 32) < cnt[i] && (lane + 32) >= conv[i]) b[i] = raw_to_key(b[i], head, cap);
    }
    warp_sort64_multi<true, B>(a, lane);
    if (CAPT > 32) warp_sort64_multi<false, B>(b, lane);
    uint64_t mn[B];
#pragma unroll
    for (int i = 0; i < B; ++i) {
        m[i] = (CAPT > 32 && b[i] > a[i]) ? b[i] : a[i];
        if (KPT <= 16) {
            const uint64_t o = shfl_xor64(m[i], 16);
            m[i] = (lane < 16) ? (m[i] > o ? m[i] : o) : 0ull;
        }
        if (KPT == 8) {
            const uint64_t o = shfl_xor64(m[i], 8);
            m[i] = (lane < 8) ? (m[i] > o ? m[i] : o) : 0ull;
        }
        mn[i] = m[i];
    }
#pragma unroll
    for (int j = KPT / 2; j > 0; j >>= 1) {
#pragma unroll
        for (int i = 0; i < B; ++i) {
            const uint64_t o = shfl_xor64(mn[i], j);
            mn[i] = mn[i] < o ? mn[i] : o;
        }
    }
#pragma unroll
    for (int i = 0; i < B; ++i) {
        mn[i] = shfl64(mn[i], 0);
        if (lane < KPT) buf[i][lane] = m[i];
        new_cnt[i] = cnt[i] < KPT ? cnt[i] : KPT;
        new_thr[i] = cnt[i] >= KPT ? key_score(mn[i]) : -INFINITY;
    }
}

// Prune every row of this warp's 32-row group whose bit is set in `mask`.
// rows = base of the warp's 32 row buffers (row r at rows + r*CAPT).
template <int CAPT, int KPT>
__device__ __forceinline__ void warp_prune_rows(unsigned mask, uint64_t* rows, int lane, int& cnt,
                                                int& conv, float& thr, uint32_t head, uint32_t cap) {
    __syncwarp();
    while (mask) {
        const int r = __ffs(mask) - 1;
        mask &= mask - 1;
        const int c = __shfl_sync(FULL, cnt, r);
        const int cv = __shfl_sync(FULL, conv, r);
        int nc;
        float nt;
        warp_prune<CAPT, KPT>(rows + r * CAPT, c, cv, head, cap, lane, nc, nt);
        if (lane == r) { cnt = nc; conv = nc; thr = nt; }
    }
    __syncwarp();
}

// Prune ALL 32 rows, four at a time (the end-of-sweep flush, where every row needs it and the
// MMA pipe is already idle: latency is all that matters).  Kept out of the hot loop on purpose:
// inlining it there costs more in instruction-cache misses than it saves.
template <int CAPT, int KPT>
__device__ __forceinline__ void warp_prune_all_rows(uint64_t* rows, int lane, int& cnt, int& conv, float& thr,
                                                    uint32_t head, uint32_t cap) {
    __syncwarp();
#pragma unroll 1
    for (int r0 = 0; r0 < 32; r0 += 4) {
        int c[4], cv[4], nc[4];
        float nt[4];
        uint64_t* bp[4];
        uint64_t m[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            c[i] = __shfl_sync(FULL, cnt, r0 + i);
            cv[i] = __shfl_sync(FULL, conv, r0 + i);
            bp[i] = rows + (r0 + i) * CAPT;
        }
        warp_prune_multi<CAPT, KPT, 4>(bp, c, cv, head, cap, lane, m, nc, nt);
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (lane == r0 + i) { cnt = nc[i]; conv = nc[i]; thr = nt[i]; }
    }
    __syncwarp();
}

__device__ __forceinline__ float bf16_bits_to_float(uint16_t b) {
    return __uint_as_float(((uint32_t)b) << 16);
}

}  // namespace pilott
