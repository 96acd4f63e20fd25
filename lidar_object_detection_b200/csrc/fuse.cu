// K1+K4+K5 fused pass over the points of a batch of frames, and K6 per-car decision.
//
// One CTA = (frame, chunk of FUSE_THREADS*FUSE_PPT consecutive points).  Each
// point is read once (one 128-bit streaming load, 16 B of HBM traffic); the
// packed masks (66 KB each) and the boxes are L2 / shared-memory traffic.
// Counters are privatised in shared memory per CTA and flushed with one global
// atomic per non-zero counter.
#include "common.cuh"

namespace {

constexpr int FUSE_THREADS = 256;
constexpr int FUSE_PPT = 4;
constexpr int FUSE_CHUNK = FUSE_THREADS * FUSE_PPT;

struct FuseArgs {
    const float4* pts;
    const int64_t* pt_off;
    const uint32_t* packed;
    const int32_t* mask_off;
    const lf_box* boxes;
    const int32_t* box_off;
    const int64_t* cnt_off;
    int32_t* total;
    int32_t* inside;
    int32_t* n_in_box;
    int32_t* n_valid;
    unsigned long long* car_bits;
    double max_depth;
    int use_oriented;
    int smem_boxes;      // boxes staged in shared memory (max over frames)
    int smem_cars;       // min(max masks per frame, 64)
};

__device__ __forceinline__ unsigned long long warp_or64(unsigned long long v) {
    const unsigned lo = __reduce_or_sync(0xFFFFFFFFu, (unsigned)v);
    const unsigned hi = __reduce_or_sync(0xFFFFFFFFu, (unsigned)(v >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

// closed-interval slab test 0 <= dot(rel,e)/ee <= 1 (cvs_erosion.py:133-138).
// d/ee <= 1 <=> d <= ee and d/ee >= 0 <=> d >= 0 for ee > 0, except when a
// negative d/ee underflows to -0 (which numpy accepts): that corner keeps the
// division.  ee == 0 gives inf/nan in numpy -> outside.
__device__ __forceinline__ bool slab(double rx, double ry, double rz, const double* e, double ee) {
    double d = __dmul_rn(rx, e[0]);
    d = __fma_rn(ry, e[1], d);
    d = __fma_rn(rz, e[2], d);
    if (!(ee > 0.0)) return false;
    if (d >= 0.0) return d <= ee;
    return __ddiv_rn(d, ee) >= 0.0;
}

__global__ void __launch_bounds__(FUSE_THREADS)
lf_fuse_kernel(const __grid_constant__ LfCam cam, const __grid_constant__ FuseArgs A) {
    extern __shared__ __align__(16) uint8_t smem[];
    const int f = blockIdx.y;
    const int64_t p0 = A.pt_off[f], p1 = A.pt_off[f + 1];
    const int64_t base = p0 + (int64_t)blockIdx.x * FUSE_CHUNK;
    if (base >= p1) return;
    const int m0 = A.mask_off[f];
    const int M = A.mask_off[f + 1] - m0;
    const int b0 = A.box_off[f];
    const int B = A.box_off[f + 1] - b0;

    lf_box* s_box = (lf_box*)smem;
    int32_t* s_inside = (int32_t*)(smem + (size_t)A.smem_boxes * sizeof(lf_box));
    int32_t* s_nin = s_inside + (size_t)A.smem_cars * A.smem_boxes;
    int32_t* s_total = s_nin + A.smem_boxes;
    int32_t* s_nvalid = s_total + A.smem_cars;

    {   // stage the frame's boxes and clear the private counters
        const double* gsrc = (const double*)(A.boxes + b0);
        double* sdst = (double*)s_box;
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += FUSE_THREADS) sdst[i] = gsrc[i];
        const int nc = A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 1;
        for (int i = threadIdx.x; i < nc; i += FUSE_THREADS) s_inside[i] = 0;
    }

    // all loads of the chunk first: FUSE_PPT independent 16-B requests per thread
    float4 p[FUSE_PPT];
    bool live[FUSE_PPT];
#pragma unroll
    for (int j = 0; j < FUSE_PPT; ++j) {
        const int64_t i = base + j * FUSE_THREADS + threadIdx.x;
        live[j] = i < p1;
        p[j] = live[j] ? ld_stream_f4(A.pts + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    __syncthreads();

    const int lane = threadIdx.x & 31;
    const int64_t plane = (int64_t)cam.H * cam.wpr;
    int my_valid = 0;

    for (int g0 = 0; g0 < (M > 0 ? M : 1); g0 += LF_MAX_CARS) {     // >64 masks: extra passes
        const int Mg = min(LF_MAX_CARS, M - g0);
#pragma unroll
        for (int j = 0; j < FUSE_PPT; ++j) {
            double uq, vq, d;
            lf_project_point(cam, p[j].x, p[j].y, p[j].z, uq, vq, d);
            const bool valid = live[j] && lf_gate(cam, uq, vq, d, A.max_depth);
            if (g0 == 0) my_valid += valid ? 1 : 0;
            if (!__any_sync(0xFFFFFFFFu, valid)) {
                if (A.car_bits && live[j] && g0 == 0) A.car_bits[base + j * FUSE_THREADS + threadIdx.x] = 0ULL;
                continue;
            }
            // ---- K4: point-in-mask gather (cvs_erosion.py:152-160) -------------
            unsigned long long bits = 0ULL;
            if (valid && Mg > 0) {
                const int ui = (int)uq, vi = (int)vq;
                const uint32_t* w = A.packed + (int64_t)(m0 + g0) * plane + (int64_t)vi * cam.wpr + (ui >> 5);
                const int sh = ui & 31;
                for (int c = 0; c < Mg; ++c)
                    bits |= (unsigned long long)((__ldg(w + c * plane) >> sh) & 1u) << c;
            }
            if (A.car_bits && live[j] && g0 == 0) A.car_bits[base + j * FUSE_THREADS + threadIdx.x] = bits;
            unsigned long long un = warp_or64(bits);
            while (un) {
                const int c = __ffsll((long long)un) - 1;
                un &= un - 1;
                const unsigned bal = __ballot_sync(0xFFFFFFFFu, (bits >> c) & 1ULL);
                if (lane == 0) atomicAdd(&s_total[c], __popc(bal));
            }
            // ---- K5: containment (cvs_erosion.py:114-145 / V1:118-139) ----------
            for (int b = 0; b < B; ++b) {
                const lf_box& bx = s_box[b];
                const bool near = valid && p[j].x >= bx.lo[0] && p[j].x <= bx.hi[0] && p[j].y >= bx.lo[1] &&
                                  p[j].y <= bx.hi[1] && p[j].z >= bx.lo[2] && p[j].z <= bx.hi[2];
                if (!__any_sync(0xFFFFFFFFu, near)) continue;
                bool in = false;
                if (near) {
                    const double X = (double)p[j].x, Y = (double)p[j].y, Z = (double)p[j].z;
                    if (A.use_oriented) {
                        const double rx = X - bx.c0[0], ry = Y - bx.c0[1], rz = Z - bx.c0[2];
                        in = slab(rx, ry, rz, bx.e1, bx.ee[0]) && slab(rx, ry, rz, bx.e2, bx.ee[1]) &&
                             slab(rx, ry, rz, bx.e3, bx.ee[2]);
                    } else {
                        in = X >= bx.amin[0] && X <= bx.amax[0] && Y >= bx.amin[1] && Y <= bx.amax[1] &&
                             Z >= bx.amin[2] && Z <= bx.amax[2];
                    }
                }
                const unsigned bal = __ballot_sync(0xFFFFFFFFu, in);
                if (!bal) continue;
                if (lane == 0 && g0 == 0) atomicAdd(&s_nin[b], __popc(bal));
                unsigned long long un2 = warp_or64(in ? bits : 0ULL);
                while (un2) {
                    const int c = __ffsll((long long)un2) - 1;
                    un2 &= un2 - 1;
                    const unsigned bb = __ballot_sync(0xFFFFFFFFu, in && ((bits >> c) & 1ULL));
                    if (lane == 0) atomicAdd(&s_inside[c * A.smem_boxes + b], __popc(bb));
                }
            }
        }
        // flush this group of cars
        const unsigned vb = __reduce_add_sync(0xFFFFFFFFu, (unsigned)my_valid);
        if (g0 == 0 && lane == 0 && vb) atomicAdd(s_nvalid, (int)vb);
        __syncthreads();
        for (int i = threadIdx.x; i < Mg * B; i += FUSE_THREADS) {
            const int c = i / B, b = i - c * B;
            const int v = s_inside[c * A.smem_boxes + b];
            if (v) {
                atomicAdd(A.inside + A.cnt_off[f] + (int64_t)(g0 + c) * B + b, v);
                s_inside[c * A.smem_boxes + b] = 0;
            }
        }
        for (int c = threadIdx.x; c < Mg; c += FUSE_THREADS) {
            const int v = s_total[c];
            if (v) { atomicAdd(A.total + m0 + g0 + c, v); s_total[c] = 0; }
        }
        if (g0 == 0) {
            for (int b = threadIdx.x; b < B; b += FUSE_THREADS) {
                const int v = s_nin[b];
                if (v) atomicAdd(A.n_in_box + b0 + b, v);
            }
            if (threadIdx.x == 0 && *s_nvalid) atomicAdd(A.n_valid + f, *s_nvalid);
        }
        __syncthreads();
    }
}

// K6: one warp per frame; lanes stride over the frame's cars, each walks the
// boxes in order keeping the first strict maximum (cvs_erosion.py:181-195).
__global__ void __launch_bounds__(128)
lf_car_stats_kernel(const int32_t* __restrict__ total, const int32_t* __restrict__ inside,
                    const int32_t* __restrict__ mask_off, const int32_t* __restrict__ box_off,
                    const int64_t* __restrict__ cnt_off, int F, int min_points, int32_t* __restrict__ best_idx,
                    int32_t* __restrict__ best_cnt, unsigned long long* __restrict__ agg) {
    const int lane = threadIdx.x & 31;
    const int f = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (f >= F) return;
    const int m0 = mask_off[f], M = mask_off[f + 1] - m0;
    const int B = box_off[f + 1] - box_off[f];
    const int32_t* ins = inside + cnt_off[f];
    unsigned rows = 0, matched = 0;
    unsigned long long st_m = 0, si_m = 0, st_all = 0;
    for (int c = lane; c < M; c += 32) {
        int best = 0, bi = -1;
        for (int b = 0; b < B; ++b) {
            const int v = ins[(int64_t)c * B + b];
            if (v > best) { best = v; bi = b; }
        }
        best_idx[m0 + c] = bi;
        best_cnt[m0 + c] = best;
        const int t = total[m0 + c];
        if (t > 0 && B > 0) {              // a row is written (cvs_erosion.py:169-173)
            rows++;
            st_all += (unsigned long long)t;
            if (bi >= 0 && best >= min_points) { matched++; st_m += (unsigned long long)t; si_m += (unsigned long long)best; }
        }
    }
    if (!agg) return;
    rows = __reduce_add_sync(0xFFFFFFFFu, rows);
    matched = __reduce_add_sync(0xFFFFFFFFu, matched);
    for (int o = 16; o; o >>= 1) {
        st_m += __shfl_xor_sync(0xFFFFFFFFu, st_m, o);
        si_m += __shfl_xor_sync(0xFFFFFFFFu, si_m, o);
        st_all += __shfl_xor_sync(0xFFFFFFFFu, st_all, o);
    }
    if (lane == 0 && rows) {
        atomicAdd(agg + 0, 1ULL);
        atomicAdd(agg + 1, (unsigned long long)rows);
        atomicAdd(agg + 2, (unsigned long long)matched);
        atomicAdd(agg + 3, st_m);
        atomicAdd(agg + 4, si_m);
        atomicAdd(agg + 5, st_all);
    }
}

}  // namespace

extern "C" int lf_fuse_batch(lf_ctx* ctx, const float* pts, const int64_t* pt_off, const uint32_t* packed,
                             const int32_t* mask_off, const lf_box* boxes, const int32_t* box_off,
                             const int64_t* cnt_off, int F, int64_t max_points_per_frame,
                             int max_masks_per_frame, int max_boxes_per_frame, double max_depth,
                             int use_oriented, int32_t* total, int32_t* inside, int32_t* n_in_box,
                             int32_t* n_valid, uint64_t* car_bits, void* stream) {
    if (!ctx || F < 0 || max_points_per_frame < 0 || max_masks_per_frame < 0 || max_boxes_per_frame < 0)
        return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0 || max_points_per_frame == 0) return LF_OK;
    if (!pts || !pt_off || !mask_off || !box_off || !cnt_off || !total || !inside || !n_in_box || !n_valid)
        return LF_E_ARG;
    if (max_masks_per_frame > 0 && !packed) return LF_E_ARG;
    if (max_boxes_per_frame > 0 && !boxes) return LF_E_ARG;
    if (((uintptr_t)pts & 15) != 0) return LF_E_ARG;
    if (F > 65535) return LF_E_UNSUPPORTED;      // gridDim.y; callers split larger batches
    LF_CUDA(ctx, cudaSetDevice(ctx->device));

    FuseArgs A;
    A.pts = (const float4*)pts; A.pt_off = pt_off; A.packed = packed; A.mask_off = mask_off;
    A.boxes = boxes; A.box_off = box_off; A.cnt_off = cnt_off;
    A.total = total; A.inside = inside; A.n_in_box = n_in_box; A.n_valid = n_valid;
    A.car_bits = (unsigned long long*)car_bits;
    A.max_depth = max_depth; A.use_oriented = use_oriented;
    A.smem_boxes = max_boxes_per_frame > 0 ? max_boxes_per_frame : 1;
    A.smem_cars = max_masks_per_frame > LF_MAX_CARS ? LF_MAX_CARS : (max_masks_per_frame > 0 ? max_masks_per_frame : 1);
    const size_t smem = (size_t)A.smem_boxes * sizeof(lf_box) +
                        4 * ((size_t)A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 4);
    if (smem > 200 * 1024) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaFuncSetAttribute(lf_fuse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t chunks = (max_points_per_frame + FUSE_CHUNK - 1) / FUSE_CHUNK;
    if (chunks > 0x7FFFFFFF) return LF_E_UNSUPPORTED;
    dim3 grid((unsigned)chunks, (unsigned)F);
    lf_fuse_kernel<<<grid, FUSE_THREADS, smem, (cudaStream_t)stream>>>(ctx->cam, A);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_car_stats(lf_ctx* ctx, const int32_t* total, const int32_t* inside, const int32_t* mask_off,
                            const int32_t* box_off, const int64_t* cnt_off, int F, int min_points,
                            int32_t* best_idx, int32_t* best_cnt, int64_t* agg, void* stream) {
    if (!ctx || F < 0) return LF_E_ARG;
    if (F == 0) return LF_OK;
    if (!total || !inside || !mask_off || !box_off || !cnt_off || !best_idx || !best_cnt) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int wpb = 4;
    lf_car_stats_kernel<<<(F + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(
        total, inside, mask_off, box_off, cnt_off, F, min_points, best_idx, best_cnt,
        (unsigned long long*)agg);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
