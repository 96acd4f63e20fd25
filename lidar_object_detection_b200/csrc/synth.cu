// Synthetic HDL-64-shaped workload generator (bench / tests input only).
// Not part of the reference's path: it only manufactures inputs of the right
// shape on the device so that 10 GB batches need no host staging.
#include "common.cuh"

namespace {

__device__ __forceinline__ uint64_t mix64(uint64_t x) {     // splitmix64 finaliser, counter-based
    x += 0x9E3779B97F4A7C15ULL;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
    return x ^ (x >> 31);
}
__device__ __forceinline__ float u01(uint64_t h) { return (float)((h >> 40) + 1) * (1.0f / 16777217.0f); }

// ray / oriented box (slab method in the box frame); returns entry distance or -1
__device__ __forceinline__ float ray_car(const double* car, float ox, float oy, float oz, float dx, float dy,
                                         float dz, float inflate) {
    const float cx = (float)car[0], cy = (float)car[1], cz = (float)car[2];
    const float yaw = (float)car[3];
    const float hl = 0.5f * (float)car[4] * inflate, hw = 0.5f * (float)car[5] * inflate,
                hh = 0.5f * (float)car[6] * inflate;
    float s, c;
    __sincosf(yaw, &s, &c);
    const float rx = ox - cx, ry = oy - cy, rz = oz - cz;
    const float lo[3] = {c * rx + s * ry, -s * rx + c * ry, rz};
    const float ld[3] = {c * dx + s * dy, -s * dx + c * dy, dz};
    const float hx[3] = {hl, hw, hh};
    float t0 = 0.f, t1 = 1e30f;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        if (fabsf(ld[a]) < 1e-9f) {
            if (fabsf(lo[a]) > hx[a]) return -1.f;
        } else {
            const float inv = 1.f / ld[a];
            float ta = (-hx[a] - lo[a]) * inv, tb = (hx[a] - lo[a]) * inv;
            if (ta > tb) { const float t = ta; ta = tb; tb = t; }
            t0 = fmaxf(t0, ta);
            t1 = fminf(t1, tb);
            if (t0 > t1) return -1.f;
        }
    }
    return t0;
}

__global__ void __launch_bounds__(256)
lf_synth_points_kernel(const double* __restrict__ cars, int ncars, int rings, int azimuths, uint64_t seed,
                       float4* __restrict__ pts) {
    const int f = blockIdx.y;
    const int n = rings * azimuths;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int ring = i / azimuths, az = i - ring * azimuths;
    const uint64_t key = mix64(seed + (uint64_t)f * 0x100000001B3ULL) ^ ((uint64_t)i * 0x9E3779B97F4A7C15ULL);
    const float j0 = u01(mix64(key ^ 1)), j1 = u01(mix64(key ^ 2)), j2 = u01(mix64(key ^ 3)),
                j3 = u01(mix64(key ^ 4)), j4 = u01(mix64(key ^ 5));
    const float deg = 0.017453292519943295f;
    const float el = (2.0f - 26.8f * (float)ring / (float)(rings - 1 > 0 ? rings - 1 : 1)) * deg;
    const float a = ((float)az * (360.0f / (float)azimuths) + (j0 - 0.5f) * 0.1f - 180.0f) * deg;
    float se, ce, sa, ca;
    sincosf(el, &se, &ce);
    sincosf(a, &sa, &ca);
    const float dx = ce * ca, dy = ce * sa, dz = se;
    float t = 1e30f;
    if (dz < -1e-4f) t = -1.73f / dz;
    const double* fc = cars + (size_t)f * ncars * 8;
    for (int c = 0; c < ncars; ++c) {
        const float tc = ray_car(fc + c * 8, 0.f, 0.f, 0.f, dx, dy, dz, 1.0f);
        if (tc > 0.f && tc < t) t = tc;
    }
    if (t > 120.f) t = 5.f + 75.f * j1;                       // clutter
    // Box-Muller range noise, sigma 2 cm
    t += 0.02f * sqrtf(-2.f * __logf(j2)) * __cosf(6.2831853f * j3);
    pts[(size_t)f * n + i] = make_float4(t * dx, t * dy, t * dz, j4);
}

struct MaskCam { float Ri[9]; float o[3]; float fx, fy, cx, cy; int W, H; };

__global__ void __launch_bounds__(256)
lf_synth_masks_kernel(const double* __restrict__ cars, int ncars, const __grid_constant__ MaskCam mc, float inflate,
                      uint8_t* __restrict__ masks) {
    const int f = blockIdx.y;
    const int npix = mc.W * mc.H;
    const int px = blockIdx.x * blockDim.x + threadIdx.x;
    if (px >= npix) return;
    const int y = px / mc.W, x = px - y * mc.W;
    const float rx = ((float)x - mc.cx) / mc.fx, ry = ((float)y - mc.cy) / mc.fy, rz = 1.f;
    const float dx = mc.Ri[0] * rx + mc.Ri[1] * ry + mc.Ri[2] * rz;
    const float dy = mc.Ri[3] * rx + mc.Ri[4] * ry + mc.Ri[5] * rz;
    const float dz = mc.Ri[6] * rx + mc.Ri[7] * ry + mc.Ri[8] * rz;
    const double* fc = cars + (size_t)f * ncars * 8;
    uint8_t* out = masks + (size_t)f * ncars * npix + px;
    for (int c = 0; c < ncars; ++c) {
        const float t = ray_car(fc + c * 8, mc.o[0], mc.o[1], mc.o[2], dx, dy, dz, inflate);
        out[(size_t)c * npix] = t >= 0.f ? 1 : 0;
    }
}

}  // namespace

extern "C" int lf_synth_points(lf_ctx* ctx, const double* cars, int ncars, int F, int rings, int azimuths,
                               uint64_t seed, float* pts, void* stream) {
    if (!ctx || !cars || !pts || ncars < 0 || F < 0 || rings <= 0 || azimuths <= 0) return LF_E_ARG;
    if (F == 0) return LF_OK;
    if (F > 65535) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int n = rings * azimuths;
    dim3 grid((n + 255) / 256, F);
    lf_synth_points_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cars, ncars, rings, azimuths, seed, (float4*)pts);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_synth_masks(lf_ctx* ctx, const double* cars, int ncars, int F, const double Tinv[16],
                              double inflate, uint8_t* masks, void* stream) {
    if (!ctx || !cars || !masks || !Tinv || ncars < 0 || F < 0) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0 || ncars == 0) return LF_OK;
    if (F > 65535) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    MaskCam mc;
    for (int r = 0; r < 3; ++r) {
        for (int c = 0; c < 3; ++c) mc.Ri[3 * r + c] = (float)Tinv[4 * r + c];
        mc.o[r] = (float)Tinv[4 * r + 3];
    }
    mc.fx = (float)ctx->cam.K[0]; mc.cx = (float)ctx->cam.K[2];
    mc.fy = (float)ctx->cam.K[4]; mc.cy = (float)ctx->cam.K[5];
    mc.W = ctx->cam.W; mc.H = ctx->cam.H;
    const int npix = mc.W * mc.H;
    dim3 grid((npix + 255) / 256, F);
    lf_synth_masks_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(cars, ncars, mc, (float)inflate, masks);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
