// K1 projection + gate (per-point outputs) and K2 depth maps.
#include "common.cuh"

// ---------------------------------------------------------------------------
// K1: one thread per point, one coalesced 128-bit load, fp64 chain, up to 28 B out.
// Algorithmic bytes: 16 in + (8+8+8+4) out per point.  HBM-bound.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
lf_project_kernel(const __grid_constant__ LfCam cam, const float4* __restrict__ pts, int64_t n,
                  double max_depth, long long* __restrict__ u, long long* __restrict__ v,
                  double* __restrict__ depth, int32_t* __restrict__ pix) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const bool pix_only = pix && !u && !v && !depth;            // only the gated pixel: the fp32 pre-gate may answer
    const PreGate pg(cam, max_depth);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 p = ld_stream_f4(pts + i);
        if (pix_only && !pg.maybe(p)) {
            pix[i] = -1;
            continue;
        }
        double uq, vq, d;
        lf_project_point(cam, p.x, p.y, p.z, uq, vq, d);
        if (u) u[i] = lf_np_f64_to_i64(uq);
        if (v) v[i] = lf_np_f64_to_i64(vq);
        if (depth) depth[i] = d;
        if (pix) pix[i] = lf_gate(cam, uq, vq, d, max_depth) ? (int32_t)vq * cam.W + (int32_t)uq : -1;
    }
}

extern "C" int lf_project(lf_ctx* ctx, const float* pts, int64_t n, double max_depth, int64_t* u,
                          int64_t* v, double* depth, int32_t* pix, void* stream) {
    if (!ctx || n < 0 || (n > 0 && !pts)) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (((uintptr_t)pts & 15) != 0) return LF_E_ARG;
    if (n == 0) return LF_OK;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int threads = 256;
    int64_t blocks = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)ctx->sm_count * 8 * 4;
    if (blocks > cap) blocks = cap;
    lf_project_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream>>>(
        ctx->cam, (const float4*)pts, n, max_depth, (long long*)u, (long long*)v, depth, pix);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

// ---------------------------------------------------------------------------
// K2: depth maps.  Pass 1 scatters one 64-bit key per gated point into an
// H*W image with a single atomic; pass 2 turns keys into f64 depths and fans
// them out to the per-car maps through the packed masks.
//   nearest: key = bits(depth f64)          atomicMin  (depth > 0 => IEEE bits are ordered)
//   last:    key = (idx+1) << 0 .. see below atomicMax  (highest point index wins, which is
//            what the index-ordered loop of seg_with_pointcloud.py:162-168 leaves behind)
// For `last` the winning index is kept and the depth recomputed in pass 2 (the
// projection is deterministic), so the stored value is the f64 depth, not a
// float32 copy.  The scratch image is L2-resident (4.2 MB at 1408x376).
// ---------------------------------------------------------------------------
__global__ void lf_fill_u64_kernel(unsigned long long* p, int64_t n, unsigned long long val) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) p[i] = val;
}

template <int POLICY>
__global__ void __launch_bounds__(256)
lf_zscatter_kernel(const __grid_constant__ LfCam cam, const float4* __restrict__ pts, int64_t n,
                   double max_depth, unsigned long long* __restrict__ img) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 p = ld_stream_f4(pts + i);
        double uq, vq, d;
        lf_project_point(cam, p.x, p.y, p.z, uq, vq, d);
        if (!lf_gate(cam, uq, vq, d, max_depth)) continue;
        const int px = (int)vq * cam.W + (int)uq;
        if (POLICY == LF_POLICY_NEAREST)
            atomicMin(img + px, (unsigned long long)__double_as_longlong(d));
        else
            atomicMax(img + px, (unsigned long long)(i + 1));
    }
}

template <int POLICY>
__global__ void __launch_bounds__(256)
lf_zresolve_kernel(const __grid_constant__ LfCam cam, const float4* __restrict__ pts,
                   const unsigned long long* __restrict__ img, const uint32_t* __restrict__ packed, int m,
                   double* __restrict__ full, double* __restrict__ per_car) {
    const int npix = cam.W * cam.H;
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= npix) return;
    const unsigned long long key = img[px];
    double d = 0.0;
    if (POLICY == LF_POLICY_NEAREST) {
        if (key != ~0ULL) d = __longlong_as_double((long long)key);
    } else if (key != 0ULL) {
        const float4 p = pts[key - 1];
        double uq, vq;
        lf_project_point(cam, p.x, p.y, p.z, uq, vq, d);
    }
    if (full) full[px] = d;
    if (per_car) {
        const int y = px / cam.W, x = px - y * cam.W;
        const int64_t widx = (int64_t)y * cam.wpr + (x >> 5);
        const int64_t plane = (int64_t)cam.H * cam.wpr;
        for (int c = 0; c < m; ++c) {
            const uint32_t w = __ldg(packed + c * plane + widx);
            per_car[(int64_t)c * npix + px] = ((w >> (x & 31)) & 1u) ? d : 0.0;
        }
    }
}

extern "C" int lf_depth_maps(lf_ctx* ctx, const float* pts, int64_t n, double max_depth, int policy,
                             const uint32_t* packed, int m, uint64_t* scratch, double* full,
                             double* per_car, void* stream) {
    if (!ctx || n < 0 || (n > 0 && !pts) || !scratch || m < 0) return LF_E_ARG;
    if (policy != LF_POLICY_NEAREST && policy != LF_POLICY_LAST) return LF_E_ARG;
    if (per_car && m > 0 && !packed) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (((uintptr_t)pts & 15) != 0) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int npix = ctx->cam.W * ctx->cam.H;
    const int threads = 256;
    lf_fill_u64_kernel<<<lf_div_up(npix, threads), threads, 0, s>>>(
        (unsigned long long*)scratch, npix, policy == LF_POLICY_NEAREST ? ~0ULL : 0ULL);
    LF_CUDA(ctx, cudaGetLastError());
    if (n > 0) {
        int64_t blocks = (n + threads - 1) / threads;
        const int64_t cap = (int64_t)ctx->sm_count * 8 * 4;
        if (blocks > cap) blocks = cap;
        if (policy == LF_POLICY_NEAREST)
            lf_zscatter_kernel<LF_POLICY_NEAREST><<<(unsigned)blocks, threads, 0, s>>>(
                ctx->cam, (const float4*)pts, n, max_depth, (unsigned long long*)scratch);
        else
            lf_zscatter_kernel<LF_POLICY_LAST><<<(unsigned)blocks, threads, 0, s>>>(
                ctx->cam, (const float4*)pts, n, max_depth, (unsigned long long*)scratch);
        LF_CUDA(ctx, cudaGetLastError());
    }
    if (full || (per_car && m > 0)) {
        if (policy == LF_POLICY_NEAREST)
            lf_zresolve_kernel<LF_POLICY_NEAREST><<<lf_div_up(npix, threads), threads, 0, s>>>(
                ctx->cam, (const float4*)pts, (const unsigned long long*)scratch, packed, m, full, per_car);
        else
            lf_zresolve_kernel<LF_POLICY_LAST><<<lf_div_up(npix, threads), threads, 0, s>>>(
                ctx->cam, (const float4*)pts, (const unsigned long long*)scratch, packed, m, full, per_car);
        LF_CUDA(ctx, cudaGetLastError());
    }
    return LF_OK;
}

// ---------------------------------------------------------------------------
// lf_depth_maps without any caller- or library-side allocation: the key image lives in a
// file-scope device array (LF_NOALLOC_MAX_PIXELS 64-bit keys = 8 MiB per GPU that loads the
// library, held until exit).  No counterpart in the reference; added
// for callers that cannot call cudaMalloc (CUDA graph capture, tight serving
// loops, harnesses that forbid it).  In host code the array's name is only a stand-in: its device
// address is taken once per device with cudaGetSymbolAddress.  Every call on a device shares
// the array (the fill launch at the start of each call clears it), so calls that overlap on
// the GPU -- different streams, or different contexts on one device -- race.
// ---------------------------------------------------------------------------
static __device__ __align__(256) unsigned long long lf_noalloc_keys[LF_NOALLOC_MAX_PIXELS];
static unsigned long long* lf_noalloc_ptr[64];             // device address of the array, per device ordinal

extern "C" int64_t lf_depth_maps_noalloc_max_pixels(void) { return LF_NOALLOC_MAX_PIXELS; }

extern "C" int lf_depth_maps_noalloc(lf_ctx* ctx, const float* pts, int64_t n, double max_depth, int policy,
                                     const uint32_t* packed, int m, double* full, double* per_car, void* stream) {
    if (!ctx) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if ((int64_t)ctx->cam.W * ctx->cam.H > (int64_t)LF_NOALLOC_MAX_PIXELS) return LF_E_LIMIT;
    if (ctx->device < 0 || ctx->device >= 64) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    if (!lf_noalloc_ptr[ctx->device]) {                     // set-up, once per device, that device current
        void* p = nullptr;
        LF_CUDA(ctx, cudaGetSymbolAddress(&p, lf_noalloc_keys));
        if (!p) return LF_E_CUDA;
        lf_noalloc_ptr[ctx->device] = (unsigned long long*)p;
    }
    return lf_depth_maps(ctx, pts, n, max_depth, policy, packed, m, (uint64_t*)lf_noalloc_ptr[ctx->device], full, per_car,
                         stream);
}

// ---------------------------------------------------------------------------
// K2 for a batch of frames: the same two passes over F frames at once.
//   scatter: CTA = (frame, segment of DM_SEG points); eight independent 128-bit loads per
//            thread in flight, the fp32 pre-gate (common.cuh) drops the ~75 % of points that
//            cannot pass the gate before any fp64 work, then one 64-bit atomic per gated
//            point into the frame's key image (keys are per-frame point indices + 1 / depth bits).
//   resolve: thread = two neighbouring pixels of one frame; 16-byte stores to the frame map
//            and to the maps of that frame's cars.  Output-bound: 8 B per pixel and map.
// Algorithmic bytes: 16 per point + 8 (key) + 8 * (full + M_f) per pixel.
// ---------------------------------------------------------------------------
namespace {

constexpr int DM_THREADS = 256;
constexpr int DM_PER_THREAD = 8;
constexpr int DM_SEG = DM_THREADS * DM_PER_THREAD;

template <int POLICY>
__global__ void __launch_bounds__(DM_THREADS)
lf_zscatter_batch_kernel(const __grid_constant__ LfCam cam, const float4* __restrict__ pts,
                         const int64_t* __restrict__ pt_off, double max_depth,
                         unsigned long long* __restrict__ img) {
    const int f = blockIdx.y;
    const int64_t p0 = pt_off[f], n = pt_off[f + 1] - p0;
    const int64_t base = (int64_t)blockIdx.x * DM_SEG;
    if (base >= n) return;
    const PreGate pg(cam, max_depth);
    unsigned long long* fimg = img + (int64_t)f * cam.W * cam.H;
    float4 p[DM_PER_THREAD];
#pragma unroll
    for (int j = 0; j < DM_PER_THREAD; ++j) {
        const int64_t i = base + j * DM_THREADS + threadIdx.x;
        p[j] = i < n ? ld_stream_f4(pts + p0 + i) : make_float4(0.f, 0.f, -1e30f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < DM_PER_THREAD; ++j) {
        const int64_t i = base + j * DM_THREADS + threadIdx.x;
        if (i >= n || !pg.maybe(p[j])) continue;
        double uq, vq, d;
        lf_project_point(cam, p[j].x, p[j].y, p[j].z, uq, vq, d);
        if (!lf_gate(cam, uq, vq, d, max_depth)) continue;
        const int px = (int)vq * cam.W + (int)uq;
        if (POLICY == LF_POLICY_NEAREST)
            atomicMin(fimg + px, (unsigned long long)__double_as_longlong(d));
        else
            atomicMax(fimg + px, (unsigned long long)(i + 1));
    }
}

template <int POLICY>
__device__ __forceinline__ double dm_depth_of(const LfCam& cam, const float4* __restrict__ fpts, unsigned long long key) {
    double d = 0.0;
    if (POLICY == LF_POLICY_NEAREST) {
        if (key != ~0ULL) d = __longlong_as_double((long long)key);
    } else if (key != 0ULL) {
        const float4 p = fpts[key - 1];
        double uq, vq;
        lf_project_point(cam, p.x, p.y, p.z, uq, vq, d);
    }
    return d;
}

template <int POLICY>
__global__ void __launch_bounds__(DM_THREADS)
lf_zresolve_batch_kernel(const __grid_constant__ LfCam cam, const float4* __restrict__ pts,
                         const int64_t* __restrict__ pt_off, const unsigned long long* __restrict__ img,
                         const uint32_t* __restrict__ packed, const int32_t* __restrict__ mask_off,
                         double* __restrict__ full, double* __restrict__ per_car) {
    const int f = blockIdx.y;
    const int npix = cam.W * cam.H;                       // even: the host checks it
    const int px = 2 * (blockIdx.x * DM_THREADS + threadIdx.x);
    if (px >= npix) return;
    const int64_t fo = (int64_t)f * npix;
    const ulonglong2 key = *(const ulonglong2*)(img + fo + px);
    const float4* fpts = pts + pt_off[f];
    double2 d;
    d.x = dm_depth_of<POLICY>(cam, fpts, key.x);
    d.y = dm_depth_of<POLICY>(cam, fpts, key.y);
    if (full) __stcs((double2*)(full + fo + px), d);
    if (per_car) {
        const int m0 = mask_off[f], m1 = mask_off[f + 1];
        const int y = px / cam.W, x = px - y * cam.W;     // W even: both pixels sit in one word
        const int64_t widx = (int64_t)y * cam.wpr + (x >> 5);
        const int64_t plane = (int64_t)cam.H * cam.wpr;
        const int sh = x & 31;
        for (int c = m0; c < m1; ++c) {
            const uint32_t w = __ldg(packed + c * plane + widx) >> sh;
            __stcs((double2*)(per_car + (int64_t)c * npix + px), make_double2((w & 1u) ? d.x : 0.0, (w & 2u) ? d.y : 0.0));
        }
    }
}

}  // namespace

extern "C" int lf_depth_maps_batch(lf_ctx* ctx, const float* pts, const int64_t* pt_off, int F,
                                   int64_t max_points_per_frame, double max_depth, int policy,
                                   const uint32_t* packed, const int32_t* mask_off, uint64_t* scratch,
                                   double* full, double* per_car, void* stream) {
    if (!ctx || F < 0 || max_points_per_frame < 0) return LF_E_ARG;
    if (policy != LF_POLICY_NEAREST && policy != LF_POLICY_LAST) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0) return LF_OK;
    if (!pt_off || !scratch || (max_points_per_frame > 0 && !pts)) return LF_E_ARG;
    if (per_car && (!packed || !mask_off)) return LF_E_ARG;
    if (((uintptr_t)pts & 15) != 0 || ((uintptr_t)scratch & 15) != 0 || ((uintptr_t)full & 15) != 0 ||
        ((uintptr_t)per_car & 15) != 0)
        return LF_E_ARG;
    if (F > 65535) return LF_E_UNSUPPORTED;                       // gridDim.y; callers split larger batches
    if ((ctx->cam.W & 1) != 0) return LF_E_UNSUPPORTED;           // pixel pairs; odd widths take lf_depth_maps
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t npix = (int64_t)ctx->cam.W * ctx->cam.H;
    // ~0 (nearest: atomicMin) or 0 (last: atomicMax) in every key
    LF_CUDA(ctx, cudaMemsetAsync(scratch, policy == LF_POLICY_NEAREST ? 0xFF : 0, (size_t)F * npix * 8, s));
    const int64_t chunks = (max_points_per_frame + DM_SEG - 1) / DM_SEG;
    if (chunks > 0x7FFFFFFF) return LF_E_UNSUPPORTED;
    if (chunks > 0) {
        dim3 grid((unsigned)chunks, (unsigned)F);
        if (policy == LF_POLICY_NEAREST)
            lf_zscatter_batch_kernel<LF_POLICY_NEAREST><<<grid, DM_THREADS, 0, s>>>(
                ctx->cam, (const float4*)pts, pt_off, max_depth, (unsigned long long*)scratch);
        else
            lf_zscatter_batch_kernel<LF_POLICY_LAST><<<grid, DM_THREADS, 0, s>>>(
                ctx->cam, (const float4*)pts, pt_off, max_depth, (unsigned long long*)scratch);
        LF_CUDA(ctx, cudaGetLastError());
    }
    if (full || per_car) {
        dim3 grid((unsigned)lf_div_up(npix / 2, DM_THREADS), (unsigned)F);
        if (policy == LF_POLICY_NEAREST)
            lf_zresolve_batch_kernel<LF_POLICY_NEAREST><<<grid, DM_THREADS, 0, s>>>(
                ctx->cam, (const float4*)pts, pt_off, (const unsigned long long*)scratch, packed, mask_off, full, per_car);
        else
            lf_zresolve_batch_kernel<LF_POLICY_LAST><<<grid, DM_THREADS, 0, s>>>(
                ctx->cam, (const float4*)pts, pt_off, (const unsigned long long*)scratch, packed, mask_off, full, per_car);
        LF_CUDA(ctx, cudaGetLastError());
    }
    return LF_OK;
}
