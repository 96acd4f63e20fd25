// Order-preserving stream compaction for the reference's function surface:
//   lf_filter_valid   the boolean gate                                 cvs_erosion.py:359
//   lf_car_counts     how many points every mask selects               cvs_erosion.py:156-157
//   lf_car_compact    car_points = points_valid[mask_indices] for ALL masks in one launch (CSR:
//                     car c's points start at sum of the counts before it)   cvs_erosion.py:156-160
// replacing M boolean-index launches (+ M device syncs) per frame by three launches and two copies.
// A block owns LF_CP_BLOCK consecutive points, a warp 32 * LF_CP_GROUPS consecutive ones, so ranks
// inside a car's list follow the point order: (block, warp, group, lane).
#include "common.cuh"

namespace {

constexpr int CP_WARPS = 8, CP_GROUPS = 8;
constexpr int CP_BLOCK = CP_WARPS * CP_GROUPS * 32;       // 2048 points per block

__global__ void __launch_bounds__(256)
lf_filter_valid_kernel(const long long* __restrict__ u, const long long* __restrict__ v, const double* __restrict__ d,
                       int64_t n, long long W, long long H, double max_depth, uint8_t* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const long long x = u[i], y = v[i];
        const double z = d[i];
        out[i] = (x >= 0) & (x < W) & (y >= 0) & (y < H) & (z > 0.0) & (z < max_depth);
    }
}

// per (block, car) counts; blockcnt is [nblk][m]
__global__ void __launch_bounds__(32 * CP_WARPS)
lf_car_count_kernel(const unsigned long long* __restrict__ bits, int64_t n, int m, int32_t* __restrict__ blockcnt) {
    __shared__ int32_t s_cnt[LF_MAX_CARS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = threadIdx.x; c < m; c += blockDim.x) s_cnt[c] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * CP_BLOCK + warp * (CP_GROUPS * 32);
    for (int g = 0; g < CP_GROUPS; ++g) {
        const int64_t i = base + g * 32 + lane;
        const unsigned long long b = i < n ? bits[i] : 0ULL;
        unsigned lo = __reduce_or_sync(0xFFFFFFFFu, (unsigned)b), hi = __reduce_or_sync(0xFFFFFFFFu, (unsigned)(b >> 32));
        unsigned long long un = ((unsigned long long)hi << 32) | lo;
        while (un) {
            const int c = __ffsll((long long)un) - 1;
            un &= un - 1;
            const unsigned bal = __ballot_sync(0xFFFFFFFFu, (b >> c) & 1ULL);
            if (lane == 0 && c < m) atomicAdd(&s_cnt[c], __popc(bal));
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < m; c += blockDim.x) blockcnt[(int64_t)blockIdx.x * m + c] = s_cnt[c];
}

// exclusive scan over the blocks, per car (one thread per car: nblk is small -- 2048 points per block);
// blockcnt becomes the offset of the block inside the car's list, counts[c] the car's total
__global__ void lf_car_scan_kernel(int32_t* __restrict__ blockcnt, int64_t nblk, int m, int32_t* __restrict__ counts) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= m) return;
    int32_t run = 0;
    for (int64_t b = 0; b < nblk; ++b) {
        const int32_t v = blockcnt[b * m + c];
        blockcnt[b * m + c] = run;
        run += v;
    }
    counts[c] = run;
}

template <typename T>
__global__ void __launch_bounds__(32 * CP_WARPS)
lf_car_compact_kernel(const unsigned long long* __restrict__ bits, int64_t n, int m, const T* __restrict__ xyz,
                      int stride_elems, const int32_t* __restrict__ blockoff, const int64_t* __restrict__ car_start,
                      T* __restrict__ out_xyz, int32_t* __restrict__ out_idx) {
    __shared__ int32_t s_w[CP_WARPS][LF_MAX_CARS];          // per-warp counts, then per-warp offsets inside the block
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = lane; c < m; c += 32) s_w[warp][c] = 0;
    __syncwarp();
    const int64_t base = (int64_t)blockIdx.x * CP_BLOCK + warp * (CP_GROUPS * 32);
    unsigned long long b[CP_GROUPS];
#pragma unroll
    for (int g = 0; g < CP_GROUPS; ++g) {
        const int64_t i = base + g * 32 + lane;
        b[g] = i < n ? bits[i] : 0ULL;
        unsigned lo = __reduce_or_sync(0xFFFFFFFFu, (unsigned)b[g]), hi = __reduce_or_sync(0xFFFFFFFFu, (unsigned)(b[g] >> 32));
        unsigned long long un = ((unsigned long long)hi << 32) | lo;
        while (un) {
            const int c = __ffsll((long long)un) - 1;
            un &= un - 1;
            const unsigned bal = __ballot_sync(0xFFFFFFFFu, (b[g] >> c) & 1ULL);
            if (lane == 0 && c < m) s_w[warp][c] += __popc(bal);
        }
        __syncwarp();
    }
    __syncthreads();
    for (int c = threadIdx.x; c < m; c += blockDim.x) {      // exclusive scan over the warps + the block's offset
        int32_t run = blockoff[(int64_t)blockIdx.x * m + c];
        for (int w = 0; w < CP_WARPS; ++w) {
            const int32_t v = s_w[w][c];
            s_w[w][c] = run;
            run += v;
        }
    }
    __syncthreads();
    const unsigned lt = (1u << lane) - 1u;
#pragma unroll
    for (int g = 0; g < CP_GROUPS; ++g) {
        const int64_t i = base + g * 32 + lane;
        T p[3] = {T(0), T(0), T(0)};
        if (b[g] != 0ULL) {
            p[0] = xyz[i * stride_elems]; p[1] = xyz[i * stride_elems + 1]; p[2] = xyz[i * stride_elems + 2];
        }
        unsigned lo = __reduce_or_sync(0xFFFFFFFFu, (unsigned)b[g]), hi = __reduce_or_sync(0xFFFFFFFFu, (unsigned)(b[g] >> 32));
        unsigned long long un = ((unsigned long long)hi << 32) | lo;
        while (un) {
            const int c = __ffsll((long long)un) - 1;
            un &= un - 1;
            const bool mine = (b[g] >> c) & 1ULL;
            const unsigned bal = __ballot_sync(0xFFFFFFFFu, mine);
            if (c >= m) continue;
            const int32_t at = s_w[warp][c];
            if (mine) {
                const int64_t pos = car_start[c] + at + __popc(bal & lt);
                out_xyz[3 * pos] = p[0]; out_xyz[3 * pos + 1] = p[1]; out_xyz[3 * pos + 2] = p[2];
                if (out_idx) out_idx[pos] = (int32_t)i;
            }
            __syncwarp();
            if (lane == 0) s_w[warp][c] = at + __popc(bal);
            __syncwarp();
        }
    }
}

__global__ void lf_car_starts_kernel(const int32_t* __restrict__ counts, int m, int64_t* __restrict__ car_start) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        int64_t run = 0;
        for (int c = 0; c < m; ++c) { car_start[c] = run; run += counts[c]; }
        car_start[m] = run;
    }
}

}  // namespace

extern "C" int lf_filter_valid(lf_ctx* ctx, const int64_t* u, const int64_t* v, const double* depth, int64_t n, int W,
                               int H, double max_depth, uint8_t* valid, void* stream) {
    if (!ctx || n < 0 || W <= 0 || H <= 0 || (n > 0 && (!u || !v || !depth || !valid))) return LF_E_ARG;
    if (n == 0) return LF_OK;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    int64_t blocks = (n + 255) / 256;
    if (blocks > (int64_t)ctx->sm_count * 32) blocks = (int64_t)ctx->sm_count * 32;
    lf_filter_valid_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        (const long long*)u, (const long long*)v, depth, n, W, H, max_depth, valid);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int64_t lf_car_scratch_elems(int64_t n, int m) {
    if (n < 0 || m < 0) return 0;
    return ((n + CP_BLOCK - 1) / CP_BLOCK) * (int64_t)m + 2 * ((int64_t)m + 1) + 2;   // int32 elements
}

extern "C" int lf_car_counts(lf_ctx* ctx, const uint64_t* bits, int64_t n, int m, int32_t* scratch, int32_t* counts,
                             void* stream) {
    if (!ctx || n < 0 || m < 0 || m > LF_MAX_CARS) return LF_E_ARG;
    if (m == 0) return LF_OK;
    if (!counts || !scratch || (n > 0 && !bits)) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t nblk = (n + CP_BLOCK - 1) / CP_BLOCK;
    if (nblk > 0x7FFFFFFF) return LF_E_UNSUPPORTED;
    if (nblk > 0) lf_car_count_kernel<<<(unsigned)nblk, 32 * CP_WARPS, 0, s>>>((const unsigned long long*)bits, n, m, scratch);
    lf_car_scan_kernel<<<1, LF_MAX_CARS, 0, s>>>(scratch, nblk, m, counts);
    // car starts (int64 [m+1]) live behind the block offsets, 8-byte aligned
    int64_t* starts = (int64_t*)(scratch + ((nblk * m + 1) & ~(int64_t)1));
    lf_car_starts_kernel<<<1, 32, 0, s>>>(counts, m, starts);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_car_compact(lf_ctx* ctx, const uint64_t* bits, int64_t n, int m, const void* xyz, int is_f64,
                              int stride_elems, const int32_t* scratch, void* out_xyz, int32_t* out_idx, void* stream) {
    if (!ctx || n < 0 || m < 0 || m > LF_MAX_CARS || stride_elems < 3) return LF_E_ARG;
    if (m == 0 || n == 0) return LF_OK;
    if (!bits || !xyz || !scratch || !out_xyz) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t nblk = (n + CP_BLOCK - 1) / CP_BLOCK;
    const int64_t* starts = (const int64_t*)(scratch + ((nblk * m + 1) & ~(int64_t)1));
    if (is_f64)
        lf_car_compact_kernel<double><<<(unsigned)nblk, 32 * CP_WARPS, 0, (cudaStream_t)stream>>>(
            (const unsigned long long*)bits, n, m, (const double*)xyz, stride_elems, scratch, starts, (double*)out_xyz, out_idx);
    else
        lf_car_compact_kernel<float><<<(unsigned)nblk, 32 * CP_WARPS, 0, (cudaStream_t)stream>>>(
            (const unsigned long long*)bits, n, m, (const float*)xyz, stride_elems, scratch, starts, (float*)out_xyz, out_idx);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
