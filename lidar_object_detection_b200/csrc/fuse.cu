// K1+K4+K5 fused pass over the points of a batch of frames, and K6 per-car decision.
//
// One CTA = (frame, segment of FUSE_SEG consecutive points).  The segment is
// streamed through a 2-stage shared-memory ring by cp.async.bulk (TMA engine):
// 30 KB sub-chunks of 1920 points (16 KB / 1024 points in the wide geometry), so every
// point costs exactly 16 B of HBM traffic and the next sub-chunk is in flight while this
// one is consumed (the geometries were swept on B200).
// Per point: fp64 projection chain -> depth gate -> two divisions only for
// points in front of the camera -> packed-mask gather (L2) -> oriented-box test
// against the boxes a warp-level bounding-box cull leaves over.  Counters are
// privatised in shared memory and flushed with one global atomic per non-zero
// counter per CTA.
#include "fuse_body.cuh"

namespace {

// Two geometries of the same kernel (swept on B200):
//   narrow  12 warps, 2 x 1920-point stages (60 KB), 15 360-point segments, 3 CTAs/SM -- when the
//           frame state (boxes + private counters) leaves room for three such CTAs per SM;
//   wide     8 warps, 2 x 1024-point stages (32 KB), 16 384-point segments, 3 CTAs/SM -- large
//           frame state (cfg5: 64 masks x 64 boxes per frame = 30 KB per CTA).
template <int SUB, int WARPS, int CTAS>
__global__ void __launch_bounds__(32 * WARPS, CTAS)
lf_fuse_kernel(const __grid_constant__ LfCam cam, const __grid_constant__ FuseArgs A) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[FUSE_STAGES];
    __shared__ int stage_cnt[FUSE_STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < FUSE_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            stage_cnt[s] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t fbase = 0;
    fuse_item<32 * WARPS, FuseNoHook, SUB, WARPS>(cam, A, blockIdx.y, blockIdx.x, smem, full_bar, stage_cnt, fbase,
                                                   FuseNoHook());
}

#ifndef LF_NARROW_SUB
#define LF_NARROW_SUB 1920
#endif
#ifndef LF_NARROW_WARPS
#define LF_NARROW_WARPS 12
#endif
#ifndef LF_NARROW_SEG
#define LF_NARROW_SEG 15360
#endif
#ifndef LF_NARROW_CTAS
#define LF_NARROW_CTAS 3
#endif
constexpr int NARROW_SUB = LF_NARROW_SUB, NARROW_WARPS = LF_NARROW_WARPS, NARROW_SEG = LF_NARROW_SEG, NARROW_CTAS = LF_NARROW_CTAS;
constexpr int WIDE_SUB = 1024, WIDE_WARPS = 8, WIDE_SEG = 16384, WIDE_CTAS = 3;
static_assert(NARROW_SEG % NARROW_SUB == 0 && WIDE_SEG % WIDE_SUB == 0, "segment = whole stages");

template <int SUB, int WARPS, int CTAS>
cudaError_t launch_fuse(const lf_ctx* ctx, FuseArgs A, size_t state_bytes, int seg_subs, int F,
                        int64_t max_points_per_frame, cudaStream_t stream) {
    const size_t smem = (size_t)FUSE_STAGES * SUB * sizeof(float4) + state_bytes;
    cudaError_t e = cudaFuncSetAttribute(lf_fuse_kernel<SUB, WARPS, CTAS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(lf_fuse_kernel<SUB, WARPS, CTAS>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) return e;
    // small batches: shorter segments, so that one frame still spreads over the machine
    int subs = seg_subs;                                  // stages per segment
    while (subs > 1 && (int64_t)F * ((max_points_per_frame + (int64_t)subs * SUB - 1) / ((int64_t)subs * SUB)) <
                           (int64_t)4 * ctx->sm_count)
        subs = (subs + 1) / 2;
    const int seg = subs * SUB;
    A.seg_points = seg;
    const int64_t chunks = (max_points_per_frame + seg - 1) / seg;
    if (chunks > 0x7FFFFFFF) return cudaErrorInvalidValue;
    dim3 grid((unsigned)chunks, (unsigned)F);
    lf_fuse_kernel<SUB, WARPS, CTAS><<<grid, 32 * WARPS, smem, stream>>>(ctx->cam, A);
    return cudaGetLastError();
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
                             const int32_t* mask_bbox, const int32_t* mask_off, const lf_box* boxes, const int32_t* box_off,
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
    if (((uintptr_t)pts & 15) != 0 || ((uintptr_t)mask_bbox & 15) != 0) return LF_E_ARG;
    if ((int64_t)ctx->cam.H * ctx->cam.wpr * LF_MAX_CARS >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
    if (F > 65535) return LF_E_UNSUPPORTED;      // gridDim.y; callers split larger batches
    LF_CUDA(ctx, cudaSetDevice(ctx->device));

    FuseArgs A;
    A.pts = (const float4*)pts; A.pt_off = pt_off; A.packed = packed; A.mask_bbox = mask_bbox; A.mask_off = mask_off;
    A.boxes = boxes; A.box_off = box_off; A.cnt_off = cnt_off;
    A.total = total; A.inside = inside; A.n_in_box = n_in_box; A.n_valid = n_valid;
    A.car_bits = (unsigned long long*)car_bits;
    A.max_depth = max_depth; A.use_oriented = use_oriented;
    A.smem_boxes = max_boxes_per_frame > 0 ? max_boxes_per_frame : 1;
    if (max_masks_per_frame > LF_MAX_CARS) return LF_E_UNSUPPORTED;   // callers split the masks of such frames
    A.smem_cars = max_masks_per_frame > 0 ? max_masks_per_frame : 1;
    const size_t state = (size_t)A.smem_boxes * sizeof(lf_box) +
                         4 * ((size_t)A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 4) + 16 * (size_t)A.smem_cars;
    if ((size_t)FUSE_STAGES * WIDE_SUB * sizeof(float4) + state > 200 * 1024) return LF_E_UNSUPPORTED;
    // NARROW_CTAS CTAs of the narrow geometry per SM (1 KB per CTA is the system's)
    // ... and only when the batch fills the machine with them: a frame or two spread better over 1024-point CTAs
    const bool narrow = (size_t)FUSE_STAGES * NARROW_SUB * sizeof(float4) + state + 1024 <= (size_t)227 * 1024 / NARROW_CTAS &&
                        (int64_t)F * ((max_points_per_frame + NARROW_SUB - 1) / NARROW_SUB) >= (int64_t)2 * ctx->sm_count;
    if (narrow)
        LF_CUDA(ctx, (launch_fuse<NARROW_SUB, NARROW_WARPS, NARROW_CTAS>(ctx, A, state, NARROW_SEG / NARROW_SUB, F,
                                                                        max_points_per_frame, (cudaStream_t)stream)));
    else
        LF_CUDA(ctx, (launch_fuse<WIDE_SUB, WIDE_WARPS, WIDE_CTAS>(ctx, A, state, WIDE_SEG / WIDE_SUB, F,
                                                                  max_points_per_frame, (cudaStream_t)stream)));
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
