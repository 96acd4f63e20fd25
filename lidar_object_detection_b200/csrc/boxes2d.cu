// Batched 3D-box projection and detection <-> box IoU / score matrices on the device
// (SURVEY 8(f) rank 2): the caller-side consumers of the same cam0 projection --
//   filter_visible_bboxes            cvs_erosion.py:55-74
//   project_3d_bbox_to_2d            V5_ProjectingBBoxes.py:215-252
//   is_bbox_in_camera_view           secondtest.py:277-359
//   calculate_iou_2d                 V4_BBox_IoU_filtering.py:118-137
//   calculate_matching_score         V5_ProjectingBBoxes.py:277-304
// The Hungarian assignment itself stays on the host (scipy, V5:360).
#include "common.cuh"

namespace {

// camera.cam2image on one cam0 corner: K[:3,:3] @ p with the fma order of
// the oracle, depth == 0 -> -1e-6, u = rint(x / |depth|) cast like numpy's astype(int).
__device__ __forceinline__ void cam2image(const LfCam& c, double x, double y, double z, long long& u, long long& v,
                                          double& depth) {
    double q[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        double a = __dmul_rn(c.K[3 * i + 0], x);
        a = __fma_rn(c.K[3 * i + 1], y, a);
        a = __fma_rn(c.K[3 * i + 2], z, a);
        q[i] = a;
    }
    depth = q[2];
    if (depth == 0.0) depth = -1e-6;
    const double ad = fabs(depth);
    u = lf_np_f64_to_i64(rint(__ddiv_rn(q[0], ad)));
    v = lf_np_f64_to_i64(rint(__ddiv_rn(q[1], ad)));
}

// np.mean of k <= 8 doubles: numpy's pairwise sum adds fewer than 8 elements in order
// and exactly 8 as ((a0+a1)+(a2+a3))+((a4+a5)+(a6+a7)).
__device__ __forceinline__ double np_mean8(const double* a, int k) {
    if (k == 0) return 0.0;
    double s;
    if (k == 8) s = __dadd_rn(__dadd_rn(__dadd_rn(a[0], a[1]), __dadd_rn(a[2], a[3])),
                              __dadd_rn(__dadd_rn(a[4], a[5]), __dadd_rn(a[6], a[7])));
    else {
        s = a[0];
        for (int i = 1; i < k; ++i) s = __dadd_rn(s, a[i]);
    }
    return __ddiv_rn(s, (double)k);
}

__global__ void __launch_bounds__(128)
lf_project_boxes_kernel(const __grid_constant__ LfCam cam, const double* __restrict__ corners, int64_t n,
                        double dmin, double dmax, int min_in_view, lf_box2d* __restrict__ out,
                        long long* __restrict__ u_out, long long* __restrict__ v_out, double* __restrict__ d_out) {
    const int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= n) return;
    const double* c = corners + b * 24;
    long long u[8], v[8];
    double d[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) cam2image(cam, c[3 * i], c[3 * i + 1], c[3 * i + 2], u[i], v[i], d[i]);
    if (u_out) {
#pragma unroll
        for (int i = 0; i < 8; ++i) { u_out[b * 8 + i] = u[i]; v_out[b * 8 + i] = v[i]; d_out[b * 8 + i] = d[i]; }
    }
    lf_box2d r;
    const long long BIG = 0x7fffffffffffffffLL;
    // V5 project_3d_bbox_to_2d: rectangle over the corners in front of the camera
    long long x0 = BIG, y0 = BIG, x1 = -BIG - 1, y1 = -BIG - 1;
    double fr[8];
    int nf = 0, nvis = 0;
    // secondtest is_bbox_in_camera_view: depth range, image bounds
    long long rx0 = BIG, ry0 = BIG, rx1 = -BIG - 1, ry1 = -BIG - 1;
    double rd[8];
    int nr = 0, nin = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const bool in_img = u[i] >= 0 && u[i] < cam.W && v[i] >= 0 && v[i] < cam.H;
        if (d[i] > 0.0) {
            x0 = min(x0, u[i]); x1 = max(x1, u[i]); y0 = min(y0, v[i]); y1 = max(y1, v[i]);
            fr[nf++] = d[i];
        }
        if (d[i] > 0.1 && in_img) ++nvis;                          // cvs_erosion.py:66-68
        if (d[i] >= dmin && d[i] <= dmax) {
            rx0 = min(rx0, u[i]); rx1 = max(rx1, u[i]); ry0 = min(ry0, v[i]); ry1 = max(ry1, v[i]);
            rd[nr++] = d[i];
            if (in_img) ++nin;
        }
    }
    r.x0 = nf ? x0 : 0; r.y0 = nf ? y0 : 0; r.x1 = nf ? x1 : 0; r.y1 = nf ? y1 : 0;
    r.avg_depth = np_mean8(fr, nf);
    r.n_front = nf;
    r.n_visible = nvis;
    r.corners_in_view = nin;
    r.n_in_range = nr;
    int code = LF_VIEW_VALID;
    if (nr == 0) code = LF_VIEW_ALL_BEHIND;
    else {
        if (nin < min_in_view) {
            const bool intersects = !(rx1 < 0 || rx0 >= cam.W || ry1 < 0 || ry0 >= cam.H);
            if (!intersects) code = LF_VIEW_NO_INTERSECTION;
        }
        if (code == LF_VIEW_VALID && nr >= 2) {
            const long long area = (rx1 - rx0) * (ry1 - ry0);      // int64 like numpy
            if (area < 100) code = LF_VIEW_TOO_SMALL;
        }
    }
    r.view_code = code;
    r.range_avg_depth = np_mean8(rd, nr);
    out[b] = r;
}

// One thread per (detection, box) pair of a frame.  A CTA lives for two dependent memory latencies (offsets, then
// records) and a few hundred float64 instructions, so what bounds the kernel is how many CTAs an SM holds:
// 32 registers -> 8 CTAs of 256 threads.
__global__ void __launch_bounds__(256, 8)
lf_match_matrices_kernel(const double* __restrict__ dets, const int32_t* __restrict__ det_off,
                         const lf_box2d* __restrict__ rects, const int32_t* __restrict__ box_off,
                         const int64_t* __restrict__ out_off, double w_iou, double w_center, double w_size,
                         double* __restrict__ iou_out, double* __restrict__ score_out) {
    const int f = blockIdx.y;
    const int d0 = det_off[f], D = det_off[f + 1] - d0;
    const int b0 = box_off[f], B = box_off[f + 1] - b0;
    const int64_t o = out_off[f];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < D * B; i += gridDim.x * blockDim.x) {
        const int di = i / B, bi = i - di * B;
        const double* q = dets + (int64_t)(d0 + di) * 4;
        const lf_box2d r = rects[b0 + bi];
        double iou = 0.0, score = -1.0 / 0.0;
        if (r.n_front > 0) {
            const double ax1 = q[0], ay1 = q[1], ax2 = q[2], ay2 = q[3];
            const double bx1 = (double)r.x0, by1 = (double)r.y0, bx2 = (double)r.x1, by2 = (double)r.y1;
            const double left = fmax(ax1, bx1), top = fmax(ay1, by1);
            const double right = fmin(ax2, bx2), bottom = fmin(ay2, by2);
            if (!(right <= left || bottom <= top)) {                   // V4:126-128
                const double inter = __dmul_rn(right - left, bottom - top);
                const double a1 = __dmul_rn(ax2 - ax1, ay2 - ay1), a2 = __dmul_rn(bx2 - bx1, by2 - by1);
                const double uni = __dadd_rn(__dadd_rn(a1, a2), -inter);
                iou = uni > 0.0 ? __ddiv_rn(inter, uni) : 0.0;
            }
            // V5:277-304
            const double dcx = (ax1 + ax2) / 2 - (bx1 + bx2) / 2, dcy = (ay1 + ay2) / 2 - (by1 + by2) / 2;
            const double dist = sqrt(__fma_rn(dcy, dcy, __dmul_rn(dcx, dcx)));
            const double center = fmax(0.0, 1.0 - dist / 1000.0);
            const double da = __dmul_rn(ax2 - ax1, ay2 - ay1);
            const double ba = __dmul_rn(bx2 - bx1, by2 - by1);
            const double size = (da > 0.0 && ba > 0.0) ? fmin(da, ba) / fmax(da, ba) : 0.0;
            score = __dadd_rn(__dadd_rn(__dmul_rn(w_iou, iou), __dmul_rn(w_center, center)), __dmul_rn(w_size, size));
        }
        iou_out[o + i] = iou;
        score_out[o + i] = score;
    }
}

}  // namespace

extern "C" int lf_project_boxes(lf_ctx* ctx, const double* corners_cam0, int64_t n, double depth_min, double depth_max,
                                int min_points_in_view, lf_box2d* out, int64_t* u, int64_t* v, double* depth,
                                void* stream) {
    if (!ctx || n < 0) return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (n == 0) return LF_OK;
    if (!corners_cam0 || !out) return LF_E_ARG;
    if ((u || v || depth) && !(u && v && depth)) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const unsigned grid = (unsigned)((n + 127) / 128);
    lf_project_boxes_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(ctx->cam, corners_cam0, n, depth_min, depth_max,
                                                                    min_points_in_view, out, (long long*)u,
                                                                    (long long*)v, depth);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_match_matrices(lf_ctx* ctx, const double* dets, const int32_t* det_off, const lf_box2d* rects,
                                 const int32_t* box_off, const int64_t* out_off, int F, int max_pairs_per_frame,
                                 double w_iou, double w_center, double w_size, double* iou, double* score,
                                 void* stream) {
    if (!ctx || F < 0 || max_pairs_per_frame < 0) return LF_E_ARG;
    if (F == 0 || max_pairs_per_frame == 0) return LF_OK;
    if (!dets || !det_off || !rects || !box_off || !out_off || !iou || !score) return LF_E_ARG;
    if (F > 65535) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    int gx = (max_pairs_per_frame + 255) / 256;
    if (gx > 64) gx = 64;
    dim3 grid((unsigned)gx, (unsigned)F);
    lf_match_matrices_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dets, det_off, rects, box_off, out_off, w_iou,
                                                                     w_center, w_size, iou, score);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
