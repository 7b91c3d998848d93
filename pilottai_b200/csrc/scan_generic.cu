// scan_generic.cu -- generic (CUDA-core) similarity scan + streaming top-KP select.
// Handles any shape; it is the validation twin of the tcgen05 kernel (scan_tc.cu) and
// shares its selection code (common.cuh).  Not the fast path.
//
// CTA = 256 threads, query tile 64 x corpus tile 64, fp32 accumulation of bf16 inputs.
// A CTA owns one (query tile, slot) pair for its whole life and sweeps the slot's range
// of corpus tiles, so each query row's running threshold / candidate buffer stays in
// shared memory from the first tile to the last.
#include "common.cuh"
#include "internal.h"

namespace pilott {

constexpr int GM = 64, GN = 64, GK = 32;

struct GenericSmem {
    float a[GK][GM + 4];
    float b[GK][GN + 4];
    float s[GM][GN + 1];
    uint64_t buf[GM][CAP];
};

__global__ void __launch_bounds__(256)
scan_generic_kernel(const __nv_bfloat16* __restrict__ qb, const __nv_bfloat16* __restrict__ vec,
                    const float* __restrict__ scale, uint64_t* __restrict__ partials, int64_t nq,
                    int64_t nslots, int64_t capacity, int64_t head, int ld, int slots, int groups) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GenericSmem& sm = *reinterpret_cast<GenericSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = blockIdx.x % groups, slot = blockIdx.x / groups;
    const int64_t q0 = (int64_t)g * GM;
    const int64_t ntiles = (nslots + GN - 1) / GN;
    const int64_t t_lo = ntiles * slot / slots, t_hi = ntiles * (slot + 1) / slots;

    float thr = -INFINITY;      // per query row (threads 0..63 own one row each)
    int cnt = 0, conv = 0;
    const uint32_t head32 = (uint32_t)head, cap32 = (uint32_t)capacity;
    const int ty = tid >> 4, tx = tid & 15;
    const int lrow = tid >> 2, lk = (tid & 3) * 8;     // loader mapping: 64 rows x 4 chunks of 8

    for (int64_t t = t_lo; t < t_hi; ++t) {
        const int64_t n0 = t * GN;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

        for (int k0 = 0; k0 < ld; k0 += GK) {
            uint4 av = make_uint4(0, 0, 0, 0), bv = make_uint4(0, 0, 0, 0);
            if (q0 + lrow < nq) av = *reinterpret_cast<const uint4*>(qb + (q0 + lrow) * ld + k0 + lk);
            if (n0 + lrow < nslots) bv = *reinterpret_cast<const uint4*>(vec + (n0 + lrow) * ld + k0 + lk);
            const uint32_t aw[4] = {av.x, av.y, av.z, av.w}, bw[4] = {bv.x, bv.y, bv.z, bv.w};
            __syncthreads();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                sm.a[lk + 2 * j][lrow] = __uint_as_float(aw[j] << 16);
                sm.a[lk + 2 * j + 1][lrow] = __uint_as_float(aw[j] & 0xffff0000u);
                sm.b[lk + 2 * j][lrow] = __uint_as_float(bw[j] << 16);
                sm.b[lk + 2 * j + 1][lrow] = __uint_as_float(bw[j] & 0xffff0000u);
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < GK; ++k) {
                const float4 a4 = *reinterpret_cast<const float4*>(&sm.a[k][ty * 4]);
                const float4 b4 = *reinterpret_cast<const float4*>(&sm.b[k][tx * 4]);
                const float a[4] = {a4.x, a4.y, a4.z, a4.w}, b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
        }
        // scaled scores -> shared tile
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float sc = scale[n0 + tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i) sm.s[ty * 4 + i][tx * 4 + j] = acc[i][j] * sc;
        }
        __syncthreads();
        // streaming select: warps 0,1; thread r owns query row r
        if (warp < 2) {
            uint64_t* rows = &sm.buf[warp * 32][0];
            uint64_t* mine = &sm.buf[tid][0];
            for (int c0 = 0; c0 < GN; c0 += 8) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float s = sm.s[tid][c0 + j];
                    if (s >= thr)      // raw entry: physical slot in the low word (converted at prune time)
                        mine[cnt++] = ((uint64_t)f2ord(s) << 32) | (uint32_t)(n0 + c0 + j);
                }
                const unsigned full = __ballot_sync(FULL, cnt > CAP - 8);
                if (full) warp_prune_rows<CAP, KP>(full, rows, lane, cnt, conv, thr, head32, cap32);
            }
        }
        // the next tile's first __syncthreads orders the reuse of the s tile
    }
    __syncthreads();
    if (warp < 2) {
        uint64_t* rows = &sm.buf[warp * 32][0];
        __syncwarp();
        for (int r = 0; r < 32; ++r) {
            const int c = __shfl_sync(FULL, cnt, r);
            const int cv = __shfl_sync(FULL, conv, r);
            int nc; float nt;
            const uint64_t m = warp_prune<CAP, KP>(rows + r * CAP, c, cv, head32, cap32, lane, nc, nt);
            const int64_t q = q0 + warp * 32 + r;
            if (q < nq) partials[(q * slots + slot) * KP + lane] = m;
        }
    }
}

int scan_generic_slots(int64_t nq, int64_t nslots, int num_sms) {
    const int64_t groups = (nq + GM - 1) / GM;
    const int64_t ntiles = (nslots + GN - 1) / GN;
    int64_t s = (3LL * num_sms + groups - 1) / groups;
    if (s > ntiles) s = ntiles;
    if (s > 64) s = 64;
    if (s < 1) s = 1;
    return (int)s;
}

cudaError_t launch_scan_generic(const ScanArgs& a, int num_sms, cudaStream_t st) {
    (void)num_sms;
    cudaError_t e = cudaFuncSetAttribute(scan_generic_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)sizeof(GenericSmem));
    if (e != cudaSuccess) return e;
    const int groups = (int)((a.nq + GM - 1) / GM);
    scan_generic_kernel<<<groups * a.slots, 256, sizeof(GenericSmem), st>>>(
        (const __nv_bfloat16*)a.qb, (const __nv_bfloat16*)a.vec, a.scale, a.partials, a.nq,
        a.nslots, a.capacity, a.head, a.ld, a.slots, groups);
    return cudaGetLastError();
}

}  // namespace pilott
