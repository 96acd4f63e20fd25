// Context, error strings, camera upload, host-side box preparation.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

extern "C" int lf_version(void) { return LF_VERSION; }

extern "C" const char* lf_strerror(int code) {
    switch (code) {
        case LF_OK: return "ok";
        case LF_E_ARG: return "bad argument";
        case LF_E_CUDA: return "CUDA error (see lf_last_cuda_error)";
        case LF_E_UNSUPPORTED: return "size not supported by the kernels";
        case LF_E_NOCAMERA: return "lf_set_camera has not been called";
        case LF_E_LIMIT: return "beyond the fixed capacity of the no-allocation entry point";
        default: return "unknown lidarfuse error";
    }
}

extern "C" int lf_create(lf_ctx** out, int device) {
    if (!out) return LF_E_ARG;
    *out = nullptr;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        (void)cudaGetLastError();
        return LF_E_CUDA;
    }
    lf_ctx* c = (lf_ctx*)calloc(1, sizeof(lf_ctx));
    if (!c) return LF_E_ARG;
    c->device = device;
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) {
        c->last_cuda = (int)cudaGetLastError();
        free(c);
        return LF_E_CUDA;
    }
    c->sm_count = sms > 0 ? sms : 148;
    int prev = -1;
    (void)cudaGetDevice(&prev);
    if (cudaSetDevice(device) != cudaSuccess ||
        cudaMalloc((void**)&c->claim_ws, LF_CLAIM_SLOTS * sizeof(unsigned int)) != cudaSuccess) {
        (void)cudaGetLastError();
        if (prev >= 0) (void)cudaSetDevice(prev);
        free(c);
        return LF_E_CUDA;
    }
    if (prev >= 0) (void)cudaSetDevice(prev);
    *out = c;
    return LF_OK;
}

extern "C" int lf_destroy(lf_ctx* ctx) {
    if (!ctx) return LF_E_ARG;
    if (ctx->claim_ws) {
        int prev = -1;
        (void)cudaGetDevice(&prev);
        if (cudaSetDevice(ctx->device) == cudaSuccess) (void)cudaFree(ctx->claim_ws);
        (void)cudaGetLastError();
        if (prev >= 0) (void)cudaSetDevice(prev);
    }
    free(ctx);
    return LF_OK;
}

extern "C" int lf_last_cuda_error(const lf_ctx* ctx) { return ctx ? ctx->last_cuda : 0; }

extern "C" int lf_set_camera(lf_ctx* ctx, const double TrVeloToRect[16], const double K[12], int W, int H) {
    if (!ctx || !TrVeloToRect || !K || W <= 0 || H <= 0) return LF_E_ARG;
    if ((int64_t)W * H >= (int64_t)1 << 30) return LF_E_UNSUPPORTED;
    for (int i = 0; i < 12; ++i) ctx->cam.T[i] = TrVeloToRect[i];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) ctx->cam.K[3 * r + c] = K[4 * r + c];
    ctx->cam.W = W;
    ctx->cam.H = H;
    ctx->cam.wpr = (W + 31) / 32;
    // fp32 pre-gate rows: P = K[:3,:3] * T (fp64), rows 0/1 shifted by the image centre
    // ((W-1)/2, (H-1)/2) times row 2.  Error of the fp32 evaluation of one row:
    // coefficient rounding + three fma roundings <= 4 * 2^-24 * sum|P_j||x_j|
    // <= 4 * 2^-24 * max|P_j| * mag with mag = |x|+|y|+|z|+1; the bounds below carry a
    // further factor 4, are taken at the gate radius gR (points beyond it bypass the
    // pre-gate), and the x / y bounds include the effect of the depth error on the
    // (W+1)/2 * d term.
    double P[12];
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            double a = 0.0;
            for (int k = 0; k < 3; ++k) a += K[4 * r + k] * TrVeloToRect[4 * k + c];
            P[4 * r + c] = a;
        }
    const double cxm = 0.5 * (W - 1), cym = 0.5 * (H - 1);
    double pm[3] = {0.0, 0.0, 0.0};
    for (int c = 0; c < 4; ++c) {
        P[c] -= cxm * P[8 + c];
        P[4 + c] -= cym * P[8 + c];
    }
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            ctx->cam.Pg[4 * r + c] = (float)P[4 * r + c];
            pm[r] = fmax(pm[r], fabs(P[4 * r + c]));
        }
    const double eps = 16.0 * 5.9604644775390625e-08;      // 16 * 2^-24
    const double gR = 1024.0;                              // metres; HDL-64 range is ~120 m
    const double mag = 3.0 * gR + 1.0;                     // >= |x|+|y|+|z|+1 inside the gate radius
    ctx->cam.gR = (float)gR;
    ctx->cam.hwx = (float)(0.5 * (W + 1));
    ctx->cam.hwy = (float)(0.5 * (H + 1));
    ctx->cam.tD = (float)(eps * fmax(pm[2], 1e-30) * mag * 1.0001);
    ctx->cam.tX = (float)(eps * (pm[0] + (W + 2.0) * pm[2]) * mag * 1.0001 + 1e-30);
    ctx->cam.tY = (float)(eps * (pm[1] + (H + 2.0) * pm[2]) * mag * 1.0001 + 1e-30);
    // Fused point pass: |fp32 row - exact row| <= 4 * 2^-24 * max_j|P_rj| * mag(p), mag = |x|+|y|+|z|+1
    // (one coefficient rounding + three fma roundings, no overflow below 2^20 m).  kT carries
    // a further factor 4 and the largest amplification any test applies to the depth error
    // ((W+2) resp. (H+2) times row 2), so one tolerance serves the depth gate, the one-pixel
    // image margin and the half-pixel residual test of the rounded pixel.
    double kT = eps * (fmax(pm[0], pm[1]) + (fmax(W, H) + 2.0) * fmax(pm[2], 1e-30)) * 1.0001 + 1e-30;
    if (!(kT >= 7.62939453125e-06)) kT = 7.62939453125e-06;   // 2^-17: a larger tolerance is always safe
    ctx->cam.kT = (float)kT;
    if ((double)ctx->cam.kT < kT) ctx->cam.kT = nextafterf(ctx->cam.kT, INFINITY);
    ctx->cam.gS = 3.2451855365842673e+32f;                 // 2^108
    ctx->cam.kTs = ctx->cam.kT * 3.0814879110195774e-33f;  // kT * 2^-108 (exact scaling; kT >= 2^-18 keeps it normal)
    ctx->cam.cxm = (float)cxm;
    ctx->cam.cym = (float)cym;
    ctx->has_camera = 1;
    return LF_OK;
}

extern "C" int lf_zero_counters(lf_ctx* ctx, void* buf, int64_t bytes, void* stream) {
    if (!ctx || bytes < 0 || (bytes > 0 && !buf)) return LF_E_ARG;
    if (bytes == 0) return LF_OK;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    LF_CUDA(ctx, cudaMemsetAsync(buf, 0, (size_t)bytes, (cudaStream_t)stream));
    return LF_OK;
}

// 3x3 inverse by cofactors; returns false when the determinant is (near) zero.
static bool inv3(const double m[9], double out[9]) {
    const double a = m[0], b = m[1], c = m[2], d = m[3], e = m[4], f = m[5], g = m[6], h = m[7], i = m[8];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    const double scale = fabs(a * e * i) + fabs(b * f * g) + fabs(c * d * h) + 1e-300;
    if (!(fabs(det) > 1e-9 * scale)) return false;
    const double r = 1.0 / det;
    out[0] = (e * i - f * h) * r; out[1] = (c * h - b * i) * r; out[2] = (b * f - c * e) * r;
    out[3] = (f * g - d * i) * r; out[4] = (a * i - c * g) * r; out[5] = (c * d - a * f) * r;
    out[6] = (d * h - e * g) * r; out[7] = (b * g - a * h) * r; out[8] = (a * e - b * d) * r;
    return true;
}

// The three closed slabs 0 <= dot(p-c0, e_i)/ee_i <= 1 intersect in the
// parallelepiped c0 + E^-1 t, t in [0,1]^3 (rows of E are e_i/ee_i).  Its
// axis-aligned bounds, padded far beyond fp64 rounding of the exact test and
// rounded outward to float32, let the kernel reject most points in fp32.
extern "C" int lf_prepare_boxes(const double* corners, int64_t n, lf_box* out) {
    if (n < 0 || (n > 0 && (!corners || !out))) return LF_E_ARG;
    for (int64_t b = 0; b < n; ++b) {
        const double* c = corners + b * 24;
        lf_box* o = out + b;
        double E[9];
        bool ok = true;
        const int idx[3] = {1, 3, 4};
        for (int i = 0; i < 3; ++i) {
            double* e = i == 0 ? o->e1 : (i == 1 ? o->e2 : o->e3);
            for (int k = 0; k < 3; ++k) e[k] = c[3 * idx[i] + k] - c[k];
            double ee = e[0] * e[0];
            ee = fma(e[1], e[1], ee);
            ee = fma(e[2], e[2], ee);
            o->ee[i] = ee;
            if (!(ee > 0.0) || !isfinite(ee)) ok = false;
            for (int k = 0; k < 3; ++k) E[3 * i + k] = ok ? e[k] / ee : 0.0;
        }
        for (int k = 0; k < 3; ++k) {
            o->c0[k] = c[k];
            double mn = c[k], mx = c[k];
            for (int j = 1; j < 8; ++j) {
                mn = fmin(mn, c[3 * j + k]);   // np.min / np.max semantics for finite input
                mx = fmax(mx, c[3 * j + k]);
            }
            o->amin[k] = mn;
            o->amax[k] = mx;
        }
        double Ei[9];
        if (ok) ok = inv3(E, Ei);
        for (int k = 0; k < 3; ++k) {
            if (!ok) { o->lo[k] = -INFINITY; o->hi[k] = INFINITY; continue; }
            double lo = o->c0[k], hi = o->c0[k];
            for (int j = 0; j < 3; ++j) {
                const double w = Ei[3 * k + j];
                if (w < 0) lo += w; else hi += w;
            }
            lo = fmin(lo, o->amin[k]);
            hi = fmax(hi, o->amax[k]);
            const double pad = 1e-3 + 1e-4 * (hi - lo) + 1e-6 * (fabs(lo) + fabs(hi));
            lo -= pad; hi += pad;
            if (!isfinite(lo) || !isfinite(hi)) { o->lo[k] = -INFINITY; o->hi[k] = INFINITY; continue; }
            o->lo[k] = nextafterf((float)lo, -INFINITY);
            o->hi[k] = nextafterf((float)hi, INFINITY);
        }
    }
    return LF_OK;
}
