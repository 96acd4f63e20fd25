// Shared declarations for the lidarfuse kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "lidarfuse.h"

// Calibration as the kernels consume it: the three used rows of TrVeloToRect
// and the 3x3 block of K (cvs_erosion.py:314, 344-345).  Passed by value as a
// __grid_constant__ kernel parameter (constant bank, no global state).
struct LfCam {
    double T[12];   // rows 0..2 of the 4x4 velo->rect transform
    double K[9];    // K[:3,:3]
    int W, H, wpr;  // image size, packed words per row
    // fp32 pre-gate of the fused kernel (never decides anything exact, see fuse.cu):
    // rows of K[:3,:3]*T with the image centre folded in.  A point with
    // max(|x|,|y|,|z|) <= gR can only pass the gate if |Pg[0..3].p| <= hwx*d + tX,
    // |Pg[4..7].p| <= hwy*d + tY and d = Pg[8..11].p >= -tD.
    float Pg[12];
    float hwx, hwy;        // (W+1)/2, (H+1)/2: half extents with a one-pixel margin
    float tX, tY, tD;      // bounds on the fp32 evaluation error for points within gR
    float gR;              // points beyond it skip the pre-gate (exact path decides)
    // fused point pass (fuse_body.cuh): one magnitude-relative tolerance for every fp32 row,
    //   tol(p) = kT * (|x|+|y|+|z|+1) >= 4 x the worst fp32 evaluation error of any row at p,
    // evaluated as fmaf((|x|+|y|+|z|) * gS, kTs, kT) with gS = 2^108, kTs = kT * 2^-108 so that it
    // overflows to +inf (= "never certain") for points beyond 2^20 m; cxm, cym = image centre
    // ((W-1)/2, (H-1)/2) the rows 0 / 1 of Pg are shifted by.
    float kT, kTs, gS, cxm, cym;
};

struct lf_ctx {
    int device;
    int has_camera;
    int last_cuda;
    int sm_count;
    LfCam cam;
    // library-owned device workspace, allocated once by lf_create: a ring of claim counters for the
    // persistent point pass (one slot per launch, zeroed on the caller's stream right before it)
    unsigned int* claim_ws;
    int claim_slot;
};
#define LF_CLAIM_SLOTS 1024

#define LF_CUDA(ctx, call)                                                  \
    do {                                                                    \
        cudaError_t e__ = (call);                                           \
        if (e__ != cudaSuccess) {                                           \
            (ctx)->last_cuda = (int)e__;                                    \
            (void)cudaGetLastError();                                       \
            return LF_E_CUDA;                                               \
        }                                                                   \
    } while (0)

// Streaming 128-bit load: the point cloud is read exactly once.
__device__ __forceinline__ float4 ld_stream_f4(const float4* p) {
    float4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}

// Projection of one velodyne point, bit-compatible with the reference's numpy:
//   pointsCam = matmul(TrVeloToRect[f64], [x y z 1]^T[f32->f64])   cvs_erosion.py:342-344
//   proj = matmul(K[:3,:3], pointsCam); depth = proj.z; depth==0 -> -1e-6
//   u = round(proj.x/|depth|), v = round(proj.y/|depth|)           (cam2image)
// OpenBLAS dgemm accumulates k = 0..3 with fused multiply-adds, so
// the chains below are spelled with __fma_rn / __dmul_rn (never re-associated).
// uq, vq are the rounded quotients as doubles (may be NaN / huge).
__device__ __forceinline__ void lf_project_point(const LfCam& c, float x, float y, float z,
                                                 double& uq, double& vq, double& depth) {
    const double X = (double)x, Y = (double)y, Z = (double)z;
    double p[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = __dmul_rn(c.T[4 * i + 0], X);
        a = __fma_rn(c.T[4 * i + 1], Y, a);
        a = __fma_rn(c.T[4 * i + 2], Z, a);
        a = __fma_rn(c.T[4 * i + 3], 1.0, a);
        p[i] = a;
    }
    double q[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = __dmul_rn(c.K[3 * i + 0], p[0]);
        a = __fma_rn(c.K[3 * i + 1], p[1], a);
        a = __fma_rn(c.K[3 * i + 2], p[2], a);
        q[i] = a;
    }
    depth = q[2];
    if (depth == 0.0) depth = -1e-6;
    const double ad = fabs(depth);
    uq = rint(__ddiv_rn(q[0], ad));
    vq = rint(__ddiv_rn(q[1], ad));
}

// The same projection in two steps, for kernels that gate on depth before
// paying for the two divisions: step 1 gives the undivided image coordinates
// and the depth (identical chains, identical bits), step 2 the rounded quotients.
__device__ __forceinline__ void lf_project_xyd(const LfCam& c, float x, float y, float z, double& qx, double& qy,
                                               double& depth) {
    const double X = (double)x, Y = (double)y, Z = (double)z;
    double p[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = __dmul_rn(c.T[4 * i + 0], X);
        a = __fma_rn(c.T[4 * i + 1], Y, a);
        a = __fma_rn(c.T[4 * i + 2], Z, a);
        a = __fma_rn(c.T[4 * i + 3], 1.0, a);
        p[i] = a;
    }
    double q[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = __dmul_rn(c.K[3 * i + 0], p[0]);
        a = __fma_rn(c.K[3 * i + 1], p[1], a);
        a = __fma_rn(c.K[3 * i + 2], p[2], a);
        q[i] = a;
    }
    qx = q[0]; qy = q[1];
    depth = q[2];
    if (depth == 0.0) depth = -1e-6;
}
__device__ __forceinline__ void lf_project_uv(double qx, double qy, double depth, double& uq, double& vq) {
    const double ad = fabs(depth);
    uq = rint(__ddiv_rn(qx, ad));
    vq = rint(__ddiv_rn(qy, ad));
}

// numpy's float64 -> int64 cast (x86 cvttsd2si): NaN and out-of-range give INT64_MIN.
__device__ __forceinline__ long long lf_np_f64_to_i64(double r) {
    if (!(fabs(r) < 9.2233720368547758e18)) return (long long)0x8000000000000000ULL;
    return __double2ll_rz(r);
}

// gate of cvs_erosion.py:359 on the rounded quotients
__device__ __forceinline__ bool lf_gate(const LfCam& c, double uq, double vq, double depth, double max_depth) {
    return (uq >= 0.0) && (uq < (double)c.W) && (vq >= 0.0) && (vq < (double)c.H) &&
           (depth > 0.0) && (depth < max_depth);
}

// fp32 look at a point before any fp64 work (SURVEY A4 gate, cvs_erosion.py:359):
// false only for points that CERTAINLY fail the gate -- behind the camera, beyond
// the range limit, or more than a pixel outside the image.  The margins (LfCam
// tX/tY/tD, lf_set_camera) cover the fp32 evaluation error for points within the
// radius gR; farther points, NaN and inf stay "maybe", so the exact fp64 path decides
// everything that matters.  Branch-free on purpose.
struct PreGate {
    float c[12], hwx, hwy, tX, tY, ntD, dhi, gR;
    __device__ __forceinline__ PreGate(const LfCam& cam, double max_depth) {
#pragma unroll
        for (int i = 0; i < 12; ++i) c[i] = cam.Pg[i];
        hwx = cam.hwx; hwy = cam.hwy; tX = cam.tX; tY = cam.tY; ntD = -cam.tD; gR = cam.gR;
        float dmax32 = (float)max_depth;
        dmax32 += fabsf(dmax32) * 1e-6f;                   // the cast may round down
        dhi = dmax32 + cam.tD;
    }
    __device__ __forceinline__ bool maybe(const float4 p) const {
        const float d = fmaf(c[8], p.x, fmaf(c[9], p.y, fmaf(c[10], p.z, c[11])));
        const float x = fmaf(c[0], p.x, fmaf(c[1], p.y, fmaf(c[2], p.z, c[3])));
        const float y = fmaf(c[4], p.x, fmaf(c[5], p.y, fmaf(c[6], p.z, c[7])));
        const float big = fmaxf(fmaxf(fabsf(p.x), fabsf(p.y)), fabsf(p.z));
        const bool out = (d < ntD) | (d > dhi) | (fabsf(x) > fmaf(hwx, d, tX)) | (fabsf(y) > fmaf(hwy, d, tY));
        return !out | !(big <= gR);
    }
};

static inline int lf_div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
