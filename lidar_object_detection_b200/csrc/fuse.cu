// K1+K4+K5 fused pass over the points of a batch of frames, and K6 per-car decision.
//
// A persistent grid of independent warps takes claims of consecutive points from a global
// counter.  Every thread streams its points through registers (128-bit loads, the next
// iteration's in flight while this one is classified; each point read exactly once: 16 B of
// HBM traffic per point), classifies them in fp32, and only survivors -- compacted into
// per-warp shared-memory queues -- reach the packed-mask gather (L2) and the float64
// containment test (fuse_body.cuh).
#include <math.h>
#include "fuse_body.cuh"

namespace {

#ifndef LF_FUSE_WARPS
#define LF_FUSE_WARPS 8
#endif
#ifndef LF_FUSE_PER
#define LF_FUSE_PER 5            /* loads in flight per thread and iteration (2..6 are instantiated) */
#endif
#ifndef LF_FUSE_PER_DENSE
#define LF_FUSE_PER_DENSE 3      /* the same for frames of millions of points */
#endif
#ifndef LF_FUSE_CTAS
#define LF_FUSE_CTAS 4           /* launch bound (registers); shared memory decides how many really fit */
#endif
#ifndef LF_FUSE_SMEM_BOXES
#define LF_FUSE_SMEM_BOXES 512   /* most boxes per frame whose cull bounds a warp caches in shared memory */
#endif
#ifndef LF_FUSE_CLAIM
#define LF_FUSE_CLAIM 2048       /* points per claim for batches that fill the machine */
#endif

template <int WARPS, int PER, bool BITS, int CTAS, bool ROWS>
__global__ void __launch_bounds__(32 * WARPS, CTAS)
lf_fuse_kernel(const __grid_constant__ LfCam cam, const __grid_constant__ FuseArgs A) {
    extern __shared__ __align__(128) uint8_t smem[];
    fuse_warp_loop<WARPS, PER, BITS, ROWS>(cam, A, smem);
}

// Persistent launch: as many CTAs as fit the machine (or as there are claims), every warp an
// independent worker taking claims of `claim_points` consecutive points from ctx's counter.
template <int WARPS, int PER, bool BITS, int CTAS>
cudaError_t launch_fuse(lf_ctx* ctx, FuseArgs A, int F, int64_t max_points_per_frame, cudaStream_t stream) {
    using G = FuseGeom<WARPS, PER, BITS>;
    const size_t smem = G::warp_bytes(A.smem_cars, A.smem_boxes) * WARPS;
    // the kernels without per-point membership output come in two builds, without and with the row-map lookup in
    // the gather (the lookup costs the dense path 2 % just by being there); the others always carry it
    auto kern = (BITS || A.row_map) ? lf_fuse_kernel<WARPS, PER, BITS, CTAS, true> : lf_fuse_kernel<WARPS, PER, BITS, CTAS, BITS>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, G::NT, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    const int64_t workers = (int64_t)ctx->sm_count * per_sm * WARPS;
    // claim size: LF_FUSE_CLAIM points for batches that keep every warp busy for many claims,
    // down to one iteration (32 * PER points) for a frame or two
    // (measured on B200: 2048 for HDL-64 sweeps, where a claim usually starts a new frame and reloads the
    // frame cache; 512 for multi-sweep frames of millions of points, where dense claims need finer balancing)
    int64_t claim = max_points_per_frame >= (1 << 20) ? LF_FUSE_CLAIM / 4 : LF_FUSE_CLAIM;
    while (claim > G::SPAN && (int64_t)F * max_points_per_frame / claim < 8 * workers) claim /= 2;
    claim = ((claim + G::SPAN - 1) / G::SPAN) * G::SPAN;
    const int64_t cpf = (max_points_per_frame + claim - 1) / claim;
    if (cpf * F >= 0x7FFF0000LL) return cudaErrorInvalidValue;
    A.claim_points = (int)claim;
    A.cpf = (int)cpf;
    A.F = F;
    A.claim = ctx->claim_ws + ctx->claim_slot;
    ctx->claim_slot = (ctx->claim_slot + 1) % LF_CLAIM_SLOTS;
    e = cudaMemsetAsync(A.claim, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    int64_t blocks = (cpf * F + WARPS - 1) / WARPS;
    if (blocks > (int64_t)ctx->sm_count * per_sm) blocks = (int64_t)ctx->sm_count * per_sm;
#ifdef LF_FUSE_GRID_CAP
    if (blocks > (int64_t)ctx->sm_count * LF_FUSE_GRID_CAP) blocks = (int64_t)ctx->sm_count * LF_FUSE_GRID_CAP;   // experiment
#endif
    kern<<<(unsigned)blocks, G::NT, smem, stream>>>(ctx->cam, A);
    return cudaGetLastError();
}

template <bool BITS>
cudaError_t launch_fuse_per(int per, lf_ctx* ctx, const FuseArgs& A, int F, int64_t maxn, cudaStream_t st) {
    switch (per) {
        case 2: return launch_fuse<LF_FUSE_WARPS, 2, BITS, LF_FUSE_CTAS>(ctx, A, F, maxn, st);
        case 3: return launch_fuse<LF_FUSE_WARPS, 3, BITS, LF_FUSE_CTAS>(ctx, A, F, maxn, st);
        case 4: return launch_fuse<LF_FUSE_WARPS, 4, BITS, LF_FUSE_CTAS>(ctx, A, F, maxn, st);
        case 5: return launch_fuse<LF_FUSE_WARPS, 5, BITS, LF_FUSE_CTAS>(ctx, A, F, maxn, st);
        default: return launch_fuse<LF_FUSE_WARPS, 6, BITS, LF_FUSE_CTAS>(ctx, A, F, maxn, st);
    }
}

// per-warp shared memory of the geometry with `per` loads per iteration (FuseGeom::warp_bytes, PER at run time)
template <bool BITS>
size_t fuse_warp_bytes(int per, int cars, int boxes) {
    const size_t qcap = (size_t)per * 32 + 32 + 64;
    return ((qcap + 64) * 16 + (BITS ? qcap * 4 : 0) + 16 * (size_t)cars + 24 * (size_t)boxes + 8 + 15) & ~(size_t)15;
}

template <bool BITS>
cudaError_t launch_fuse_any(lf_ctx* ctx, const FuseArgs& A, int F, int64_t maxn, cudaStream_t st) {
    // loads in flight per thread and iteration (an iteration of a warp covers 32 * per consecutive points):
    // LF_FUSE_PER for ordinary sweeps, as many as leave LF_FUSE_CTAS CTAs per SM when the per-warp caches of the
    // frame state are large.  Small batches (a frame or two: latency, not throughput) take shorter iterations
    // still, so that a claim -- one warp's serial work -- is 128 or 64 points and the frame spreads over every
    // warp of the machine.
    int per = maxn >= (1 << 20) ? LF_FUSE_PER_DENSE : LF_FUSE_PER;
    if (per > 6) per = 6;
    while (per > 2 && fuse_warp_bytes<BITS>(per, A.smem_cars, A.smem_boxes) * LF_FUSE_WARPS + 1024 > (size_t)227 * 1024 / LF_FUSE_CTAS)
        --per;
    const int64_t warps = (int64_t)ctx->sm_count * LF_FUSE_CTAS * LF_FUSE_WARPS;
    while (per > 2 && (int64_t)F * maxn / (32 * per) < 2 * warps) per = per / 2 > 2 ? per / 2 : 2;
    return launch_fuse_per<BITS>(per, ctx, A, F, maxn, st);
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
    return lf_fuse_batch_rows(ctx, pts, pt_off, packed, mask_bbox, mask_off, boxes, box_off, cnt_off, F, max_points_per_frame,
                              max_masks_per_frame, max_boxes_per_frame, max_depth, use_oriented, total, inside, n_in_box,
                              n_valid, car_bits, nullptr, stream);
}

extern "C" int lf_fuse_batch_rows(lf_ctx* ctx, const float* pts, const int64_t* pt_off, const uint32_t* packed,
                                  const int32_t* mask_bbox, const int32_t* mask_off, const lf_box* boxes,
                                  const int32_t* box_off, const int64_t* cnt_off, int F, int64_t max_points_per_frame,
                                  int max_masks_per_frame, int max_boxes_per_frame, double max_depth, int use_oriented,
                                  int32_t* total, int32_t* inside, int32_t* n_in_box, int32_t* n_valid,
                                  uint64_t* car_bits, const uint32_t* row_map, void* stream) {
    if (!ctx || F < 0 || max_points_per_frame < 0 || max_masks_per_frame < 0 || max_boxes_per_frame < 0)
        return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0 || max_points_per_frame == 0) return LF_OK;
    if (!pts || !pt_off || !mask_off || !box_off || !cnt_off || !n_valid) return LF_E_ARG;
    if (max_masks_per_frame > 0 && (!packed || !total)) return LF_E_ARG;
    if (max_boxes_per_frame > 0 && (!boxes || !n_in_box)) return LF_E_ARG;
    if (max_masks_per_frame > 0 && max_boxes_per_frame > 0 && !inside) return LF_E_ARG;
    if (((uintptr_t)pts & 15) != 0 || ((uintptr_t)mask_bbox & 15) != 0) return LF_E_ARG;
    if (row_map && !mask_bbox) return LF_E_ARG;             // the rows that are valid are the bounding box's
    if ((int64_t)ctx->cam.H * ctx->cam.wpr * LF_MAX_CARS >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
    if (ctx->cam.W > 65535 || ctx->cam.H > 65535) return LF_E_UNSUPPORTED;   // queued pixels are 16 + 16 bits
    if (max_masks_per_frame > LF_MAX_CARS && car_bits) return LF_E_UNSUPPORTED;   // one 64-bit membership word per point
    LF_CUDA(ctx, cudaSetDevice(ctx->device));

    FuseArgs A;
    A.pts = (const float4*)pts; A.pt_off = pt_off; A.packed = packed; A.mask_bbox = mask_bbox; A.mask_off = mask_off;
    A.boxes = boxes; A.box_off = box_off; A.cnt_off = cnt_off;
    A.row_map = row_map; A.rm_wpr = (ctx->cam.H + 31) / 32;
    A.total = total; A.inside = inside; A.n_in_box = n_in_box; A.n_valid = n_valid;
    A.car_bits = (unsigned long long*)car_bits;
    A.max_depth = max_depth;
    A.use_oriented = use_oriented & LF_FUSE_ORIENTED;
    A.force_exact = (use_oriented & LF_FUSE_EXACT) != 0;
    // float brackets of the range gate: dmax_hi >= max_depth > dmax_lo (NaN: nothing is ever certain)
    float dh = (float)max_depth, dl = (float)max_depth;
    if ((double)dh < max_depth) dh = nextafterf(dh, INFINITY);
    if (!((double)dl < max_depth)) dl = nextafterf(dl, -INFINITY);
    A.dmax_hi = dh; A.dmax_lo = dl;
    // up to 64 boxes per frame: cull bounds cached per warp in shared memory; more (real frames have up to 314):
    // read from the box records through L1 -- no 7.5 KB reload per claim, and the CTAs keep their occupancy
    A.smem_boxes = max_boxes_per_frame > LF_FUSE_SMEM_BOXES ? 0 : (max_boxes_per_frame > 0 ? max_boxes_per_frame : 1);
    A.smem_cars = max_masks_per_frame > 0 ? (max_masks_per_frame < LF_MAX_CARS ? max_masks_per_frame : LF_MAX_CARS) : 1;
    A.claim = nullptr; A.F = 0; A.cpf = 0; A.claim_points = 0; A.group = 0;
    cudaStream_t st = (cudaStream_t)stream;
    // the per-warp caches of the largest frame must leave room for the half-size queues
    if (fuse_warp_bytes<true>(2, A.smem_cars, A.smem_boxes) * LF_FUSE_WARPS + 1024 > (size_t)227 * 1024)
        return LF_E_UNSUPPORTED;
    // the claim space of one launch is 31 bits: very large batches go out in several launches
    // (the CSR arrays are simply entered at frame f0)
    const int64_t fmax = (int64_t)1 << 20;
    // frames with more than LF_MAX_CARS masks: one pass over the points per group of 64 masks; the passes
    // after the first add to the per-car counters only (n_valid and n_in_box are per frame / per box)
    const int groups = max_masks_per_frame > LF_MAX_CARS ? (max_masks_per_frame + LF_MAX_CARS - 1) / LF_MAX_CARS : 1;
    for (int g = 0; g < groups; ++g)
        for (int64_t f0 = 0; f0 < F; f0 += fmax) {
            const int nf = (int)(F - f0 < fmax ? F - f0 : fmax);
            FuseArgs Af = A;
            Af.group = g;
            Af.pt_off += f0; Af.mask_off += f0; Af.box_off += f0; Af.cnt_off += f0;
            Af.n_valid = g == 0 ? A.n_valid + f0 : nullptr;
            if (g > 0) Af.n_in_box = nullptr;
            if (car_bits)
                LF_CUDA(ctx, launch_fuse_any<true>(ctx, Af, nf, max_points_per_frame, st));
            else
                LF_CUDA(ctx, launch_fuse_any<false>(ctx, Af, nf, max_points_per_frame, st));
        }
    return LF_OK;
}

extern "C" int lf_car_stats(lf_ctx* ctx, const int32_t* total, const int32_t* inside, const int32_t* mask_off,
                            const int32_t* box_off, const int64_t* cnt_off, int F, int min_points,
                            int32_t* best_idx, int32_t* best_cnt, int64_t* agg, void* stream) {
    if (!ctx || F < 0) return LF_E_ARG;
    if (F == 0) return LF_OK;
    if (!mask_off || !box_off || !cnt_off) return LF_E_ARG;
    if (!total || !best_idx || !best_cnt) return LF_OK;      // a batch without masks has no cars to decide
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int wpb = 4;
    lf_car_stats_kernel<<<(F + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(
        total, inside, mask_off, box_off, cnt_off, F, min_points, best_idx, best_cnt,
        (unsigned long long*)agg);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

// erode + fuse + decision in one ABI call (three stand-alone launches).  `sched` is reserved: the
// persistent erode+fuse kernel it once selected was measured slower than the launches and is gone.
extern "C" int lf_process_batch(lf_ctx* ctx, const float* pts, const int64_t* pt_off, const void* masks, int mask_mode,
                                int64_t n_masks, const int32_t* mask_off, const uint8_t* se, int k, int iterations,
                                uint32_t* packed, int32_t* mask_bbox, const lf_box* boxes, const int32_t* box_off,
                                const int64_t* cnt_off, int F, int64_t max_points_per_frame, int max_masks_per_frame,
                                int max_boxes_per_frame, double max_depth, int use_oriented, int min_points,
                                int32_t* total, int32_t* inside, int32_t* n_in_box, int32_t* n_valid,
                                int32_t* best_idx, int32_t* best_cnt, int64_t* agg, int32_t* sched, void* stream) {
    (void)sched;
    if (!ctx || F < 0 || n_masks < 0 || max_points_per_frame < 0 || max_masks_per_frame < 0 ||
        max_boxes_per_frame < 0 || iterations < 0)
        return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0) return LF_OK;
    int rc;
    if (masks && n_masks > 0) {
        rc = lf_erode_pack(ctx, masks, mask_mode, n_masks, ctx->cam.H, ctx->cam.W, se, k, iterations, packed, mask_bbox,
                           stream);
        if (rc != LF_OK) return rc;
    }
    if (max_points_per_frame > 0) {
        rc = lf_fuse_batch(ctx, pts, pt_off, packed, mask_bbox, mask_off, boxes, box_off, cnt_off, F,
                           max_points_per_frame, max_masks_per_frame, max_boxes_per_frame, max_depth, use_oriented,
                           total, inside, n_in_box, n_valid, nullptr, stream);
        if (rc != LF_OK) return rc;
        if (best_idx && best_cnt) {
            rc = lf_car_stats(ctx, total, inside, mask_off, box_off, cnt_off, F, min_points, best_idx, best_cnt, agg,
                              stream);
            if (rc != LF_OK) return rc;
        }
    }
    return LF_OK;
}
