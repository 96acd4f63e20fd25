// One persistent kernel for the whole per-batch body of cvs_erosion.process_frames
// (:342-372): mask erosion + bit-packing (K3) and the fused point pass (K1+K4+K5) share
// the machine instead of running back to back.
//
// Why: the erode pass is HBM-bound (it streams 1 B per mask pixel) while the point pass
// is latency / issue bound (fp64 chains, L2 gathers).  Run one after the other, each
// leaves the other resource idle.  Here every SM holds a mix of CTAs of both kinds:
// a CTA repeatedly takes a ticket -- "erode mask m" or "fuse segment s of frame f" -- from
// one global counter whose sequence interleaves the two kinds (see BatchSched).  A fuse
// item starts only once every mask of its frame has been eroded (per-frame completion
// counter, release/acquire); the sequence puts it far enough behind that this is
// normally already true, and close enough that the packed masks are still in L2.
//
// No deadlock: only fuse items wait, only for erode items with smaller tickets, and a
// ticket that has been taken is held by a running CTA (as its current item, or as the
// prefetched next item of a CTA whose current ticket is smaller still).
#include "erode_body.cuh"
#include "fuse_body.cuh"

namespace {

enum { ROLE_EXIT = 0, ROLE_ERODE = 1, ROLE_FUSE = 2 };

// Work is handed out by ONE ticket counter in a fixed, dependency-ordered sequence.
// Tickets come in groups of (slots_m + chunks): group q holds the erode slots of frame q
// (one per mask, slots_m = max masks per frame) followed by the fuse slots of frame
// q - lag (one per segment).  A frame's points are therefore claimed `lag` groups after
// its masks -- by then they are normally packed (and still in L2) -- and every SM sees a
// steady mix of both kinds of work.  Slots a ragged batch does not use are skipped.
struct BatchSched {
    int* ticket;         // next ticket
    int* done;           // [F] masks eroded per frame
    int F;
    int slots_m;         // erode slots per group
    int chunks;          // fuse slots per group (segments per frame, max over frames)
    int lag;             // groups between a frame's masks and its points
    int n_tickets;       // (F + lag) * (slots_m + chunks)
};

struct WorkItem {        // lives in shared memory, written by the scheduler lane
    int role;
    int a;               // erode: global mask index     fuse: frame
    int b;               // erode: frame                 fuse: segment
    int need;            // fuse: masks of the frame; 0 once seen complete
};

__device__ __forceinline__ int ld_acquire_gpu(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(int* p, int v) {
    asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Take the next valid ticket: one atomic plus one round of loads.  Never waits.
__device__ __forceinline__ void claim(const BatchSched& S, const FuseArgs& A, WorkItem* w) {
    const int G = S.slots_m + S.chunks;
    for (;;) {
        const int k = atomicAdd(S.ticket, 1);
        if (k >= S.n_tickets) { w->role = ROLE_EXIT; return; }
        const int q = k / G, r = k - q * G;
        if (r < S.slots_m) {
            if (q >= S.F) continue;
            const int m0 = A.mask_off[q], m1 = A.mask_off[q + 1];
            if (m0 + r >= m1) continue;
            w->a = m0 + r; w->b = q; w->need = 0; w->role = ROLE_ERODE;
            return;
        }
        const int g = q - S.lag, seg = r - S.slots_m;
        if (g < 0) continue;
        const int64_t p0 = A.pt_off[g], p1 = A.pt_off[g + 1];
        if (p0 + (int64_t)seg * A.seg_points >= p1) continue;
        const int need = A.mask_off[g + 1] - A.mask_off[g];
        w->a = g; w->b = seg;
        w->need = ld_acquire_gpu(S.done + g) >= need ? 0 : need;
        w->role = ROLE_FUSE;
        return;
    }
}

struct ClaimHook {
    const BatchSched& S;
    const FuseArgs& A;
    WorkItem* slot;
    __device__ __forceinline__ void operator()() const { claim(S, A, slot); }
};

template <class SE>
__global__ void __launch_bounds__(BATCH_THREADS, 3)
lf_batch_kernel(const __grid_constant__ LfCam cam, const __grid_constant__ FuseArgs A,
                const __grid_constant__ ErodeParams P, const __grid_constant__ BatchSched S,
                const uint8_t* __restrict__ masks, uint32_t* __restrict__ packed, int32_t* __restrict__ bbox) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t er_full[ER_STAGES_BATCH];
    __shared__ __align__(8) uint64_t fu_full[FUSE_STAGES];
    __shared__ int fu_cnt[FUSE_STAGES];
    __shared__ WorkItem s_next;
    const int tid = threadIdx.x;
    const bool sched = tid == 32 * FUSE_CONSUMERS;         // lane 0 of the extra warp: scheduler + copy issuer
    if (tid == 0) {
        for (int s = 0; s < ER_STAGES_BATCH; ++s) mbar_init(&er_full[s], 1);
        for (int s = 0; s < FUSE_STAGES; ++s) {
            mbar_init(&fu_full[s], 1);
            fu_cnt[s] = 0;
        }
        mbar_fence_init();
    }
    if (sched) claim(S, A, &s_next);
    __syncthreads();
    uint32_t cbase = 0, fbase = 0;
    for (;;) {
        if (sched && s_next.role == ROLE_FUSE && s_next.need) {
            // the frame's masks were claimed `lag` groups ago by CTAs that are running: this ends
            while (ld_acquire_gpu(S.done + s_next.a) < s_next.need) __nanosleep(128);
        }
        const WorkItem cur = s_next;
        __syncthreads();                                   // everybody has the item; smem of the last one is free
        if (cur.role == ROLE_EXIT) break;
        if (cur.role == ROLE_ERODE) {
            if (tid < ER_THREADS) {
                for (int strip = 0; strip < P.strips; ++strip) {
                    if (strip) er_sync();                  // the previous strip's buffers are being stored
                    erode_fast_item<SE, ER_STAGES_BATCH>(P, masks, packed, bbox, (int64_t)cur.a * P.strips + strip, smem, er_full,
                                        cbase);
                }
                __threadfence();                           // packed words + bbox visible device-wide ...
            } else if (sched) {
                claim(S, A, &s_next);                      // overlapped with the erosion
            }
            __syncthreads();
            if (tid == 0) red_release_gpu_add(S.done + cur.b, 1);     // ... before the frame counts this mask
        } else {
            fuse_item<BATCH_THREADS>(cam, A, cur.a, cur.b, smem, fu_full, fu_cnt, fbase, ClaimHook{S, A, &s_next});
            __syncthreads();
        }
    }
}

template <class SE>
cudaError_t launch_batch(const LfCam& cam, const FuseArgs& A, const ErodeParams& P, const BatchSched& S,
                         const uint8_t* masks, uint32_t* packed, int32_t* bbox, size_t smem, int sm_count,
                         cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(lf_batch_kernel<SE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lf_batch_kernel<SE>, BATCH_THREADS, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    const long long grid = (long long)per_sm * sm_count;
    BatchSched S2 = S;
    // lag: twice as many mask slots as there are resident CTAs between a frame's masks and its points
    S2.lag = (int)((2 * grid + S.slots_m - 1) / S.slots_m);
    if (S2.lag > S.F) S2.lag = S.F;
    if (S2.lag < 1) S2.lag = 1;
    const long long nt = ((long long)S.F + S2.lag) * (S.slots_m + S.chunks);
    if (nt >= 0x7FFF0000LL) return cudaErrorInvalidValue;
    S2.n_tickets = (int)nt;
    lf_batch_kernel<SE><<<(unsigned)grid, BATCH_THREADS, smem, st>>>(cam, A, P, S2, masks, packed, bbox);
    return cudaGetLastError();
}

}  // namespace

extern "C" int lf_process_batch(lf_ctx* ctx, const float* pts, const int64_t* pt_off, const void* masks, int mask_mode,
                                int64_t n_masks, const int32_t* mask_off, const uint8_t* se, int k, int iterations,
                                uint32_t* packed, int32_t* mask_bbox, const lf_box* boxes, const int32_t* box_off,
                                const int64_t* cnt_off, int F, int64_t max_points_per_frame, int max_masks_per_frame,
                                int max_boxes_per_frame, double max_depth, int use_oriented, int min_points,
                                int32_t* total, int32_t* inside, int32_t* n_in_box, int32_t* n_valid,
                                int32_t* best_idx, int32_t* best_cnt, int64_t* agg, int32_t* sched, void* stream) {
    if (!ctx || F < 0 || n_masks < 0 || max_points_per_frame < 0 || max_masks_per_frame < 0 ||
        max_boxes_per_frame < 0 || iterations < 0)
        return LF_E_ARG;
    if (!ctx->has_camera) return LF_E_NOCAMERA;
    if (F == 0) return LF_OK;
    const int H = ctx->cam.H, W = ctx->cam.W;
    const bool have_points = max_points_per_frame > 0;
    int rc;
    // The persistent kernel needs: raw masks to erode, the bulk-copy fast path with a
    // compile-time structuring element, points to fuse, scheduler words, and item counts
    // that fit the 30-bit item field.  Anything else takes the two stand-alone launches.
    bool fused = sched && masks && n_masks > 0 && have_points && mask_bbox && n_masks < 0x7FFFFFFF &&
                 max_masks_per_frame <= LF_MAX_CARS;
    ErodeParams P;
    int spec = -1, fast = 0;
    size_t er_smem = 0;
    if (fused) {
        if (iterations > 0 && (!se || k < 1 || (k & 1) == 0)) return LF_E_ARG;
        if (iterations > 0 && k > LF_MAX_SE) return LF_E_UNSUPPORTED;
        rc = lf_erode_plan(ctx, masks, mask_mode, n_masks, H, W, se, k, iterations, ER_MAX_ROWS_BATCH, ER_STAGES_BATCH, &P,
                           &spec, &fast, &er_smem);
        if (rc != LF_OK) return rc;
        fused = fast && spec >= 0 && mask_mode == 0;     // the persistent kernel takes uint8 masks
    }
    const int64_t chunks = (max_points_per_frame + FUSE_SEG - 1) / FUSE_SEG;
    if (fused && ((int64_t)F + 2048) * (chunks + LF_MAX_CARS) >= 0x7FFF0000LL) fused = false;
    if (!fused) {
        if (masks && n_masks > 0) {
            rc = lf_erode_pack(ctx, masks, mask_mode, n_masks, H, W, se, k, iterations, packed, mask_bbox, stream);
            if (rc != LF_OK) return rc;
        }
        if (have_points) {
            rc = lf_fuse_batch(ctx, pts, pt_off, packed, mask_bbox, mask_off, boxes, box_off, cnt_off, F,
                               max_points_per_frame, max_masks_per_frame, max_boxes_per_frame, max_depth,
                               use_oriented, total, inside, n_in_box, n_valid, nullptr, stream);
            if (rc != LF_OK) return rc;
        }
    } else {
        if (!pts || !pt_off || !mask_off || !box_off || !cnt_off || !total || !inside || !n_in_box || !n_valid ||
            !packed)
            return LF_E_ARG;
        if (max_boxes_per_frame > 0 && !boxes) return LF_E_ARG;
        if (((uintptr_t)pts & 15) != 0 || ((uintptr_t)mask_bbox & 15) != 0) return LF_E_ARG;
        if ((int64_t)H * ctx->cam.wpr * LF_MAX_CARS >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
        LF_CUDA(ctx, cudaSetDevice(ctx->device));
        FuseArgs A;
        A.pts = (const float4*)pts; A.pt_off = pt_off; A.packed = packed; A.mask_bbox = mask_bbox; A.mask_off = mask_off;
        A.boxes = boxes; A.box_off = box_off; A.cnt_off = cnt_off;
        A.total = total; A.inside = inside; A.n_in_box = n_in_box; A.n_valid = n_valid;
        A.car_bits = nullptr;
        A.max_depth = max_depth; A.use_oriented = use_oriented;
        A.smem_boxes = max_boxes_per_frame > 0 ? max_boxes_per_frame : 1;
        A.smem_cars = max_masks_per_frame > 0 ? max_masks_per_frame : 1;
        A.seg_points = FUSE_SEG;
        const size_t fu_smem = (size_t)FUSE_STAGES * FUSE_SUB * sizeof(float4) + (size_t)A.smem_boxes * sizeof(lf_box) +
                               4 * ((size_t)A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 4) + 16 * (size_t)A.smem_cars;
        const size_t smem = fu_smem > er_smem ? fu_smem : er_smem;
        if (smem > 200 * 1024) return LF_E_UNSUPPORTED;
        BatchSched S;
        S.ticket = sched; S.done = sched + 2;
        S.F = F; S.slots_m = max_masks_per_frame > 0 ? max_masks_per_frame : 1; S.chunks = (int)chunks;
        S.lag = 1; S.n_tickets = 0;                         // set by launch_batch (they depend on the grid)
        const uint8_t* mk = (const uint8_t*)masks;
        cudaStream_t st = (cudaStream_t)stream;
        cudaError_t e;
        switch (spec) {
            case 0: e = launch_batch<SeNone>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 1: e = launch_batch<SeCross3>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 2: e = launch_batch<SeRect3>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 3: e = launch_batch<SeEllipse5>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 4: e = launch_batch<SeRect5>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 5: e = launch_batch<SeCross5>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 6: e = launch_batch<SeEllipse7>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            case 7: e = launch_batch<SeRect7>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
            default: e = launch_batch<SeCross7>(ctx->cam, A, P, S, mk, packed, mask_bbox, smem, ctx->sm_count, st); break;
        }
        LF_CUDA(ctx, e);
    }
    if (have_points && best_idx && best_cnt) {
        rc = lf_car_stats(ctx, total, inside, mask_off, box_off, cnt_off, F, min_points, best_idx, best_cnt, agg, stream);
        if (rc != LF_OK) return rc;
    }
    return LF_OK;
}
