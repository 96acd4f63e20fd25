// Stand-alone pieces of the reference's function surface that work on points
// already gated / compacted by the caller:
//   lf_mask_gather       extract_car_points_by_mask      cvs_erosion.py:148-162
//   lf_points_in_box     (oriented_)point_in_bbox        cvs_erosion.py:114-145, V1:118-139
//   lf_count_in_boxes    counting loop of calculate_car_point_statistics  cvs_erosion.py:175-195
// The batch path does all of this inside lf_fuse_batch; these exist so that the
// V1-V5 function signatures can be served one call at a time.
#include "common.cuh"

namespace {

constexpr int LF_COUNT_SMEM_BOXES = 512;              // 96 KB of box records

__device__ __forceinline__ bool lf_slab(double rx, double ry, double rz, const double* e, double ee) {
    double d = __dmul_rn(rx, e[0]);
    d = __fma_rn(ry, e[1], d);
    d = __fma_rn(rz, e[2], d);
    if (!(ee > 0.0)) return false;
    if (d >= 0.0) return d <= ee;
    return __ddiv_rn(d, ee) >= 0.0;
}

__device__ __forceinline__ bool lf_in_box(const lf_box& bx, double X, double Y, double Z, int oriented) {
    if (oriented) {
        const double rx = X - bx.c0[0], ry = Y - bx.c0[1], rz = Z - bx.c0[2];
        return lf_slab(rx, ry, rz, bx.e1, bx.ee[0]) && lf_slab(rx, ry, rz, bx.e2, bx.ee[1]) &&
               lf_slab(rx, ry, rz, bx.e3, bx.ee[2]);
    }
    return X >= bx.amin[0] && X <= bx.amax[0] && Y >= bx.amin[1] && Y <= bx.amax[1] && Z >= bx.amin[2] &&
           Z <= bx.amax[2];
}

__global__ void __launch_bounds__(256)
lf_mask_gather_kernel(const uint32_t* __restrict__ packed, int m, int H, int W, int wpr,
                      const long long* __restrict__ u, const long long* __restrict__ v, int64_t n,
                      unsigned long long* __restrict__ bits_out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t plane = (int64_t)H * wpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const long long x = u[i], y = v[i];
        unsigned long long bits = 0ULL;
        if (x >= 0 && x < W && y >= 0 && y < H) {
            const uint32_t* w = packed + y * wpr + (x >> 5);
            const int sh = (int)(x & 31);
            for (int c = 0; c < m; ++c) bits |= (unsigned long long)((__ldg(w + c * plane) >> sh) & 1u) << c;
        }
        bits_out[i] = bits;
    }
}

// The same with 32-bit arithmetic after one unsigned range test per coordinate (negative values wrap to huge ones),
// the words of eight masks requested together, and each bit moved into place by one funnel shift
// (acc = acc >> 1 | bit << 31).  Needs m * H * wpr < 2^31 words (the host checks).
// What bounds it on a sweep (ncu, profiles/r02c_next_ncu_summary.txt): the L1 sector rate -- the 32 pixels of a warp
// lie on several image rows, so one gather instruction touches ~9.5 sectors, sixteen times per point.
__global__ void __launch_bounds__(256)
lf_mask_gather_lean_kernel(const uint32_t* __restrict__ packed, int m, int H, int W, int wpr,
                           const long long* __restrict__ u, const long long* __restrict__ v, int64_t n,
                           unsigned long long* __restrict__ bits_out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const uint32_t plane = (uint32_t)H * (uint32_t)wpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const unsigned long long x = (unsigned long long)__ldg(u + i), y = (unsigned long long)__ldg(v + i);
        uint32_t lo = 0u, hi = 0u;
        if (x < (unsigned long long)W && y < (unsigned long long)H) {
            const uint32_t off = (uint32_t)y * (uint32_t)wpr + ((uint32_t)x >> 5);   // word index inside one mask
            const uint32_t sh = (uint32_t)x & 31u;
            for (int h = 0; h < m; h += 32) {                 // masks 0..31 into lo, 32..63 into hi
                const int mh = min(32, m - h);
                uint32_t acc = 0u;
                for (int c0 = 0; c0 < mh; c0 += 8) {
                    uint32_t t[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) t[j] = c0 + j < mh ? __ldg(packed + (off + (uint32_t)(h + c0 + j) * plane)) : 0u;
#pragma unroll
                    for (int j = 0; j < 8; ++j) acc = __funnelshift_r(acc, t[j] >> sh, 1);
                }
                // mh rounded up to 8 shifts were made: mask h + c sits at bit 32 - 8 * ceil(mh / 8) + c
                acc >>= (32 - ((mh + 7) & ~7)) & 31;
                if (h == 0) lo = acc; else hi = acc;
            }
        }
        bits_out[i] = ((unsigned long long)hi << 32) | lo;
    }
}

// float / double coordinates as conservative float brackets [lo, hi] of the value (equal for float input)
template <typename T> __device__ __forceinline__ float bracket_lo(T v);
template <typename T> __device__ __forceinline__ float bracket_hi(T v);
template <> __device__ __forceinline__ float bracket_lo<float>(float v) { return v; }
template <> __device__ __forceinline__ float bracket_hi<float>(float v) { return v; }
template <> __device__ __forceinline__ float bracket_lo<double>(double v) { return __double2float_rd(v); }
template <> __device__ __forceinline__ float bracket_hi<double>(double v) { return __double2float_ru(v); }

// x, y, z of point i: one 128-bit load when the rows are float4 records
template <typename T>
__device__ __forceinline__ void load_xyz(const T* xyz, int64_t i, int stride_elems, bool vec4, T& x, T& y, T& z) {
    const T* p = xyz + i * stride_elems;
    x = p[0]; y = p[1]; z = p[2];
}
template <>
__device__ __forceinline__ void load_xyz<float>(const float* xyz, int64_t i, int stride_elems, bool vec4, float& x, float& y,
                                                float& z) {
    if (vec4) {
        const float4 q = ld_stream_f4((const float4*)xyz + i);
        x = q.x; y = q.y; z = q.z;
    } else {
        const float* p = xyz + i * stride_elems;
        x = p[0]; y = p[1]; z = p[2];
    }
}

// certainly outside the box's float32 cull bounds (lo / hi of the record bracket the box; NaN is never "outside" here,
// the exact test then rejects it)
__device__ __forceinline__ bool outside_cull(const float* lo, const float* hi, float xl, float xh, float yl, float yh,
                                             float zl, float zh) {
    return xh < lo[0] || xl > hi[0] || yh < lo[1] || yl > hi[1] || zh < lo[2] || zl > hi[2];
}

template <typename T>
__global__ void __launch_bounds__(256)
lf_points_in_box_kernel(const T* __restrict__ xyz, int64_t n, int stride_elems, int vec4, const lf_box* __restrict__ box,
                        int oriented, uint8_t* __restrict__ out) {
    __shared__ lf_box s_box;
    if (threadIdx.x < sizeof(lf_box) / 8) ((double*)&s_box)[threadIdx.x] = ((const double*)box)[threadIdx.x];
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    // the next iteration's point is requested before this one is tested (two loads per thread in flight)
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    T nx = 0, ny = 0, nz = 0;
    if (i < n) load_xyz<T>(xyz, i, stride_elems, vec4 != 0, nx, ny, nz);
    for (; i < n; i += stride) {
        const T x = nx, y = ny, z = nz;
        if (i + stride < n) load_xyz<T>(xyz, i + stride, stride_elems, vec4 != 0, nx, ny, nz);
        // the float64 test only for points inside the record's float32 cull bounds
        bool in = false;
        if (!outside_cull(s_box.lo, s_box.hi, bracket_lo<T>(x), bracket_hi<T>(x), bracket_lo<T>(y), bracket_hi<T>(y),
                          bracket_lo<T>(z), bracket_hi<T>(z)))
            in = lf_in_box(s_box, (double)x, (double)y, (double)z, oriented);
        out[i] = in ? 1 : 0;
    }
}

__device__ __forceinline__ float wmin_f32(float v) {
    float r;
    asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ float wmax_f32(float v) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}

// Counting loop of calculate_car_point_statistics (cvs_erosion.py:175-195) for one call's points.  A warp takes 128
// consecutive points per iteration (neighbours of a sweep or of one car's point set: a small bounding box), four per
// lane with their loads in flight together, tests the boxes' cull bounds against the bounding box of the 128 points
// -- one box per lane -- and runs the per-point tests only for the boxes that are left (none, for nine iterations in
// ten of a sweep); set ids are read when the first hit of an iteration needs them, hits are added with one atomic
// per (point set, box) present in a group of 32.
template <typename T>
__global__ void __launch_bounds__(256, 4)
lf_count_in_boxes_kernel(const T* __restrict__ xyz, int64_t n, int stride_elems, int vec4,
                         const int32_t* __restrict__ set_id, const lf_box* __restrict__ boxes, int B,
                         int oriented, int32_t* __restrict__ inside) {
    // boxes staged in shared memory when they fit (they do for every frame of the sample: <= 314), their cull
    // bounds a second time as [6][B] so that one box per lane reads without bank conflicts
    extern __shared__ __align__(16) uint8_t s_raw[];
    const bool staged = B <= LF_COUNT_SMEM_BOXES;
    float* s_cull = (float*)(s_raw + (size_t)B * sizeof(lf_box));
    if (staged) {
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += blockDim.x) ((double*)s_raw)[i] = ((const double*)boxes)[i];
        for (int b = threadIdx.x; b < B; b += blockDim.x)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                s_cull[k * B + b] = boxes[b].lo[k];
                s_cull[(3 + k) * B + b] = boxes[b].hi[k];
            }
        __syncthreads();
    }
    const lf_box* bxs = staged ? (const lf_box*)s_raw : boxes;
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const float INF = __int_as_float(0x7f800000);
    const int64_t gw = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t base = gw * 128; base < n; base += nw * 128) {
        T x[4], y[4], z[4];
        bool have[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int64_t i = base + 32 * k + lane;
            have[k] = i < n;
            x[k] = y[k] = z[k] = 0;
            if (have[k]) load_xyz<T>(xyz, i, stride_elems, vec4 != 0, x[k], y[k], z[k]);
        }
        // bounding box of the 128 points.  fminf / fmaxf and redux.sync.f32 drop NaN operands: a NaN coordinate does
        // not widen the box, and its own tests below are false everywhere, as in the reference.
        float b0[3] = {INF, INF, INF}, b1[3] = {-INF, -INF, -INF};
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (have[k]) {
                b0[0] = fminf(b0[0], bracket_lo<T>(x[k])); b1[0] = fmaxf(b1[0], bracket_hi<T>(x[k]));
                b0[1] = fminf(b0[1], bracket_lo<T>(y[k])); b1[1] = fmaxf(b1[1], bracket_hi<T>(y[k]));
                b0[2] = fminf(b0[2], bracket_lo<T>(z[k])); b1[2] = fmaxf(b1[2], bracket_hi<T>(z[k]));
            }
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            b0[a] = wmin_f32(b0[a]);
            b1[a] = wmax_f32(b1[a]);
        }
        int sid[4] = {0, 0, 0, 0};
        bool sid_loaded = false;
        for (int g = 0; g < B; g += 32) {
            const int bl = g + lane;
            bool cand = false;
            if (bl < B) {
                float lo[3], hi[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    lo[k] = staged ? s_cull[k * B + bl] : __ldg(&boxes[bl].lo[k]);
                    hi[k] = staged ? s_cull[(3 + k) * B + bl] : __ldg(&boxes[bl].hi[k]);
                }
                cand = !outside_cull(lo, hi, b0[0], b1[0], b0[1], b1[1], b0[2], b1[2]);
            }
            unsigned cmask = __ballot_sync(FULL, cand);
            while (cmask) {
                const int b = g + __ffs((int)cmask) - 1;
                cmask &= cmask - 1;
                const lf_box& bx = bxs[b];
                bool in[4];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    in[k] = have[k] &&
                            !outside_cull(bx.lo, bx.hi, bracket_lo<T>(x[k]), bracket_hi<T>(x[k]), bracket_lo<T>(y[k]),
                                          bracket_hi<T>(y[k]), bracket_lo<T>(z[k]), bracket_hi<T>(z[k])) &&
                            lf_in_box(bx, (double)x[k], (double)y[k], (double)z[k], oriented);
                if (!__any_sync(FULL, in[0] | in[1] | in[2] | in[3])) continue;
                if (!sid_loaded) {                               // warp-uniform
                    sid_loaded = true;
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        if (have[k]) sid[k] = __ldg(set_id + base + 32 * k + lane);
                }
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (!__any_sync(FULL, in[k])) continue;
                    // one atomic per point set present among the hits of this group of 32
                    const unsigned peers = __match_any_sync(FULL, in[k] ? sid[k] : (int)(0x80000000u | (unsigned)lane));
                    if (in[k] && lane == __ffs((int)peers) - 1) atomicAdd(inside + (int64_t)sid[k] * B + b, __popc(peers));
                }
            }
        }
    }
}

}  // namespace

static unsigned grid_for(const lf_ctx* ctx, int64_t n) {
    int64_t blocks = (n + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 32;
    return (unsigned)(blocks > cap ? cap : (blocks < 1 ? 1 : blocks));
}

extern "C" int lf_mask_gather(lf_ctx* ctx, const uint32_t* packed, int m, int H, int W, const int64_t* u,
                              const int64_t* v, int64_t n, uint64_t* bits, void* stream) {
    if (!ctx || m < 0 || m > LF_MAX_CARS || H <= 0 || W <= 0 || n < 0) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!u || !v || !bits || (m > 0 && !packed)) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int wpr = (W + 31) / 32;
    if ((int64_t)(m > 0 ? m : 1) * H * wpr < ((int64_t)1 << 31))
        lf_mask_gather_lean_kernel<<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            packed, m, H, W, wpr, (const long long*)u, (const long long*)v, n, (unsigned long long*)bits);
    else
        lf_mask_gather_kernel<<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            packed, m, H, W, wpr, (const long long*)u, (const long long*)v, n, (unsigned long long*)bits);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_points_in_box(lf_ctx* ctx, const void* xyz, int is_f64, int64_t n, int stride_elems,
                                const lf_box* box, int use_oriented, uint8_t* out, void* stream) {
    if (!ctx || n < 0 || stride_elems < 3) return LF_E_ARG;
    if (n == 0) return LF_OK;
    if (!xyz || !box || !out) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int vec4 = !is_f64 && stride_elems == 4 && ((uintptr_t)xyz & 15) == 0;   // float4 records: one 128-bit load per point
    if (is_f64)
        lf_points_in_box_kernel<double><<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            (const double*)xyz, n, stride_elems, 0, box, use_oriented, out);
    else
        lf_points_in_box_kernel<float><<<grid_for(ctx, n), 256, 0, (cudaStream_t)stream>>>(
            (const float*)xyz, n, stride_elems, vec4, box, use_oriented, out);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_count_in_boxes(lf_ctx* ctx, const void* xyz, int is_f64, int64_t n, int stride_elems,
                                 const int32_t* set_id, const lf_box* boxes, int B, int use_oriented,
                                 int32_t* inside, void* stream) {
    if (!ctx || n < 0 || B < 0 || stride_elems < 3) return LF_E_ARG;
    if (n == 0 || B == 0) return LF_OK;
    if (!xyz || !set_id || !boxes || !inside) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t smem = B <= LF_COUNT_SMEM_BOXES ? (size_t)B * (sizeof(lf_box) + 24) : 0;
    const int vec4 = !is_f64 && stride_elems == 4 && ((uintptr_t)xyz & 15) == 0;
    if (is_f64) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_count_in_boxes_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(LF_COUNT_SMEM_BOXES * (sizeof(lf_box) + 24))));
        lf_count_in_boxes_kernel<double><<<grid_for(ctx, n), 256, smem, (cudaStream_t)stream>>>(
            (const double*)xyz, n, stride_elems, 0, set_id, boxes, B, use_oriented, inside);
    } else {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_count_in_boxes_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)(LF_COUNT_SMEM_BOXES * (sizeof(lf_box) + 24))));
        lf_count_in_boxes_kernel<float><<<grid_for(ctx, n), 256, smem, (cudaStream_t)stream>>>(
            (const float*)xyz, n, stride_elems, vec4, set_id, boxes, B, use_oriented, inside);
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
