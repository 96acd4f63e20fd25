// K1+K4+K5 fused pass over the points of a batch of frames, and K6 per-car decision.
//
// One CTA = (frame, segment of FUSE_SEG consecutive points).  The segment is
// streamed through a 3-stage shared-memory ring by cp.async.bulk (TMA engine):
// 16 KB sub-chunks of 1024 points, so every point costs exactly 16 B of HBM
// traffic and two sub-chunks per CTA are always in flight with no register cost.
// Per point: fp64 projection chain -> depth gate -> two divisions only for
// points in front of the camera -> packed-mask gather (L2) -> oriented-box test
// against the boxes a warp-level bounding-box cull leaves over.  Counters are
// privatised in shared memory and flushed with one global atomic per non-zero
// counter per CTA.
#include "common.cuh"
#include "bulk.cuh"

namespace {

constexpr int FUSE_CONSUMERS = 8;                         // consumer warps
constexpr int FUSE_THREADS = 32 * (FUSE_CONSUMERS + 1);   // + one producer warp (bulk-copy issue)
constexpr int FUSE_SUB = 1024;                            // points per stage = 16 KB
constexpr int FUSE_GROUPS = FUSE_SUB / 32;                // 32-point groups per stage, handed out dynamically
constexpr int FUSE_STAGES = 4;                            // power of two: stage / parity by shifts
constexpr int FUSE_SUBS_PER_SEG = 16;
constexpr int FUSE_SEG = FUSE_SUB * FUSE_SUBS_PER_SEG;    // 16384 points per CTA

struct FuseArgs {
    const float4* pts;
    const int64_t* pt_off;
    const uint32_t* packed;
    const int32_t* mask_bbox;   // optional [sum M][4] from lf_erode_pack, NULL = test every mask
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
    int smem_cars;       // max masks per frame (<= LF_MAX_CARS)
};

__device__ __forceinline__ unsigned long long warp_or64(unsigned long long v) {
    const unsigned lo = __reduce_or_sync(0xFFFFFFFFu, (unsigned)v);
    const unsigned hi = __reduce_or_sync(0xFFFFFFFFu, (unsigned)(v >> 32));
    return ((unsigned long long)hi << 32) | lo;
}

// warp-wide float min / max in one instruction (redux.sync.f32 is sm_100a-only; SASS CREDUX)
__device__ __forceinline__ float warp_min_f32(float v) {
    float r;
    asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
}
__device__ __forceinline__ float warp_max_f32(float v) {
    float r;
    asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
    return r;
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

__device__ __forceinline__ bool point_in_box(const lf_box& bx, float x, float y, float z, int oriented) {
    const double X = (double)x, Y = (double)y, Z = (double)z;
    if (oriented) {
        const double rx = X - bx.c0[0], ry = Y - bx.c0[1], rz = Z - bx.c0[2];
        return slab(rx, ry, rz, bx.e1, bx.ee[0]) && slab(rx, ry, rz, bx.e2, bx.ee[1]) &&
               slab(rx, ry, rz, bx.e3, bx.ee[2]);
    }
    return X >= bx.amin[0] && X <= bx.amax[0] && Y >= bx.amin[1] && Y <= bx.amax[1] && Z >= bx.amin[2] &&
           Z <= bx.amax[2];
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

// Everything one warp does for 32 consecutive points (one per lane) of which at
// least one passed the pre-gate (`maybe`).
__device__ __forceinline__ void fuse_warp_points(const LfCam& cam, const FuseArgs& A, float4 p, bool live, bool maybe,
                                                 int64_t gidx, int lane, int M, int B, int m0, const uint32_t* mask_base,
                                                 int plane32, const lf_box* s_box, int32_t* s_inside,
                                                 int32_t* s_nin, int32_t* s_total, int& my_valid) {
    const unsigned FULL = 0xFFFFFFFFu;
    bool valid = false;
    double uq = 0.0, vq = 0.0;
    if (maybe) {
        double qx, qy, d;
        lf_project_xyd(cam, p.x, p.y, p.z, qx, qy, d);
        const bool front = d > 0.0 && d < A.max_depth;        // depth half of cvs_erosion.py:359
        if (front) {
            // u = rint(qx / d): one reciprocal instead of two divisions.  q~ = qx * (1/d) is
            // within 2 ulp of the correctly rounded quotient, so rint() agrees unless q~ sits
            // within |q~| * 2^-50 of a half-integer; only then the exact division decides.
            const double r = __drcp_rn(d);
            const double qu = __dmul_rn(qx, r), qv = __dmul_rn(qy, r);
            uq = rint(qu);
            vq = rint(qv);
            const double eu = fabs(fabs(qu - uq) - 0.5), ev = fabs(fabs(qv - vq) - 0.5);
            if (eu <= fabs(qu) * 8.9e-16 + 1e-300) uq = rint(__ddiv_rn(qx, d));
            if (ev <= fabs(qv) * 8.9e-16 + 1e-300) vq = rint(__ddiv_rn(qy, d));
            valid = (uq >= 0.0) && (uq < (double)cam.W) && (vq >= 0.0) && (vq < (double)cam.H);
        }
    }
    const unsigned vmask = __ballot_sync(FULL, valid);
    if (!vmask) {
        if (A.car_bits && live) A.car_bits[gidx] = 0ULL;
        return;
    }
    my_valid += valid ? 1 : 0;
    // ---- K4: point-in-mask gather (cvs_erosion.py:152-160) ---------------------
    // Lanes first test the masks' bounding boxes against the pixel bounds of the warp's
    // valid points (one mask per lane); only the surviving masks are read.
    unsigned long long bits = 0ULL;
    if (M > 0) {
        const int ui = valid ? (int)uq : 0, vi = valid ? (int)vq : 0;
        const int off = vi * cam.wpr + (ui >> 5);            // word offset inside one packed mask
        const int sh = ui & 31;
        int u0 = 0, u1 = 0, v0 = 0, v1 = 0;
        if (A.mask_bbox) {
            u0 = __reduce_min_sync(FULL, valid ? ui : 0x7fffffff);
            u1 = __reduce_max_sync(FULL, valid ? ui : -1);
            v0 = __reduce_min_sync(FULL, valid ? vi : 0x7fffffff);
            v1 = __reduce_max_sync(FULL, valid ? vi : -1);
        }
        for (int g = 0; g < M; g += 32) {
            const int cl = g + lane;
            bool cand = cl < M;
            if (cand && A.mask_bbox) {
                const int4 bb = __ldg((const int4*)A.mask_bbox + m0 + cl);
                const int x0 = (1 << 20) - bb.x, y0 = (1 << 20) - bb.y, x1 = bb.z - 1, y1 = bb.w - 1;
                cand = u1 >= x0 && u0 <= x1 && v1 >= y0 && v0 <= y1;
            }
            unsigned cmask = __ballot_sync(FULL, cand);
            while (cmask) {
                const int c = g + __ffs((int)cmask) - 1;
                cmask &= cmask - 1;
                if (valid) {
                    const uint32_t w = __ldg(mask_base + (unsigned)(off + c * plane32));
                    bits |= (unsigned long long)((w >> sh) & 1u) << c;
                }
            }
        }
    }
    if (A.car_bits && live) A.car_bits[gidx] = bits;
    unsigned long long un = warp_or64(bits);
    while (un) {
        const int c = __ffsll((long long)un) - 1;
        un &= un - 1;
        const unsigned bal = __ballot_sync(FULL, (bits >> c) & 1ULL);
        if (lane == 0) atomicAdd(&s_total[c], __popc(bal));
    }
    if (B == 0) return;
    // ---- warp bounding box of the valid points, then lanes test boxes in parallel
    const float INF = __int_as_float(0x7f800000);
    const float wx0 = warp_min_f32(valid ? p.x : INF), wx1 = warp_max_f32(valid ? p.x : -INF);
    const float wy0 = warp_min_f32(valid ? p.y : INF), wy1 = warp_max_f32(valid ? p.y : -INF);
    const float wz0 = warp_min_f32(valid ? p.z : INF), wz1 = warp_max_f32(valid ? p.z : -INF);
    for (int g = 0; g < B; g += 32) {
        const int bl = g + lane;
        bool cand = false;
        if (bl < B) {
            const lf_box& bx = s_box[bl];
            cand = wx1 >= bx.lo[0] && wx0 <= bx.hi[0] && wy1 >= bx.lo[1] && wy0 <= bx.hi[1] && wz1 >= bx.lo[2] &&
                   wz0 <= bx.hi[2];
        }
        unsigned cmask = __ballot_sync(FULL, cand);
        // ---- K5: containment (cvs_erosion.py:114-145 / V1:118-139) on the candidates
        while (cmask) {
            const int b = g + __ffs((int)cmask) - 1;
            cmask &= cmask - 1;
            const lf_box& bx = s_box[b];
            const bool near = valid && p.x >= bx.lo[0] && p.x <= bx.hi[0] && p.y >= bx.lo[1] && p.y <= bx.hi[1] &&
                              p.z >= bx.lo[2] && p.z <= bx.hi[2];
            const bool in = near && point_in_box(bx, p.x, p.y, p.z, A.use_oriented);
            const unsigned bal = __ballot_sync(FULL, in);
            if (!bal) continue;
            if (lane == 0) atomicAdd(&s_nin[b], __popc(bal));
            unsigned long long un2 = warp_or64(in ? bits : 0ULL);
            while (un2) {
                const int c = __ffsll((long long)un2) - 1;
                un2 &= un2 - 1;
                const unsigned bb = __ballot_sync(FULL, in && ((bits >> c) & 1ULL));
                if (lane == 0) atomicAdd(&s_inside[c * A.smem_boxes + b], __popc(bb));
            }
        }
    }
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// Warp-specialised: warp 8 only issues bulk copies (one elected lane), warps 0-7
// consume.  A stage holds 32 groups of 32 points; consumer warps take groups in
// order from a shared counter (frustum points cluster in runs of a few hundred,
// so a static split leaves most warps idle at a block barrier), copy their point
// to a register, release the group on the stage's "empty" mbarrier and compute.
// No block-wide barrier inside the streaming loop.
__global__ void __launch_bounds__(FUSE_THREADS, 3)
lf_fuse_kernel(const __grid_constant__ LfCam cam, const __grid_constant__ FuseArgs A) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[FUSE_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[FUSE_STAGES];
    const int f = blockIdx.y;
    const int64_t p0 = A.pt_off[f], p1 = A.pt_off[f + 1];
    const int64_t seg0 = p0 + (int64_t)blockIdx.x * FUSE_SEG;
    if (seg0 >= p1) return;
    const int npts = (int)min((int64_t)FUSE_SEG, p1 - seg0);
    const int nsub = (npts + FUSE_SUB - 1) / FUSE_SUB;
    const int ngroups = (npts + 31) / 32;
    const int m0 = A.mask_off[f];
    const int M = A.mask_off[f + 1] - m0;
    const int b0 = A.box_off[f];
    const int B = A.box_off[f + 1] - b0;

    float4* stage = (float4*)smem;                                               // [FUSE_STAGES][FUSE_SUB]
    lf_box* s_box = (lf_box*)(smem + FUSE_STAGES * FUSE_SUB * sizeof(float4));
    int32_t* s_inside = (int32_t*)((uint8_t*)s_box + (size_t)A.smem_boxes * sizeof(lf_box));
    int32_t* s_nin = s_inside + (size_t)A.smem_cars * A.smem_boxes;
    int32_t* s_total = s_nin + A.smem_boxes;
    int32_t* s_nvalid = s_total + A.smem_cars;
    const float4* src = A.pts + seg0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < FUSE_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], FUSE_CONSUMERS);
        }
        mbar_fence_init();
    }
    __syncthreads();
    if (warp == FUSE_CONSUMERS && lane == 0) {
        for (int i = 0; i < FUSE_STAGES && i < nsub; ++i) {
            const uint32_t bytes = (uint32_t)min(FUSE_SUB, npts - i * FUSE_SUB) * 16u;
            mbar_expect_tx(&full_bar[i], bytes);
            bulk_g2s(stage + i * FUSE_SUB, src + (int64_t)i * FUSE_SUB, bytes, &full_bar[i]);
        }
    }
    {   // while the first sub-chunks fly: stage the frame's boxes, clear the private counters
        const double* gsrc = (const double*)(A.boxes + b0);
        double* sdst = (double*)s_box;
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += FUSE_THREADS) sdst[i] = gsrc[i];
        const int nc = A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 1;
        for (int i = threadIdx.x; i < nc; i += FUSE_THREADS) s_inside[i] = 0;
    }
    __syncthreads();

    if (warp == FUSE_CONSUMERS) {
        if (lane == 0) {                                   // producer: refill a stage once all its groups were read
            for (int i = FUSE_STAGES; i < nsub; ++i) {
                const int s = i % FUSE_STAGES;
                mbar_wait(&empty_bar[s], (uint32_t)((i / FUSE_STAGES - 1) & 1));
                const uint32_t bytes = (uint32_t)min(FUSE_SUB, npts - i * FUSE_SUB) * 16u;
                fence_proxy_async_smem();
                mbar_expect_tx(&full_bar[s], bytes);
                bulk_g2s(stage + s * FUSE_SUB, src + (int64_t)i * FUSE_SUB, bytes, &full_bar[s]);
            }
        }
    } else {
        const int plane32 = cam.H * cam.wpr;               // words per packed mask (< 2^25, see lf_set_camera)
        const uint32_t* mask_base = A.packed + (int64_t)m0 * plane32;
        int my_valid = 0;
        const PreGate pg(cam, A.max_depth);
        // A stage holds 32 groups of 32 points; warp w owns groups w, w+8, w+16, w+24 of every
        // stage (frustum points come in runs of a few hundred, so round-robin dealing gives every
        // warp its share of the expensive ones).  One barrier wait and one release per warp and
        // stage; the fp32 pre-gate of the four groups runs first, and only groups with a
        // surviving point are read again for the exact path.
        const bool all_groups = A.car_bits != nullptr;      // per-point output: every group writes
        for (int sub = 0; sub < nsub; ++sub) {
            const int s = sub & (FUSE_STAGES - 1);
            mbar_wait(&full_bar[s], (uint32_t)((sub / FUSE_STAGES) & 1));
            const float4* sp = stage + s * FUSE_SUB + warp * 32 + lane;
            const int k0 = sub * FUSE_SUB + warp * 32 + lane;          // point index inside the segment
            unsigned mb = 0;
#pragma unroll
            for (int j = 0; j < FUSE_GROUPS / FUSE_CONSUMERS; ++j) {
                const float4 p = sp[j * 32 * FUSE_CONSUMERS];
                const bool live = k0 + j * 32 * FUSE_CONSUMERS < npts;
                mb |= (unsigned)(live & pg.maybe(p)) << j;
            }
            unsigned todo = __reduce_or_sync(0xFFFFFFFFu, mb);
            if (all_groups) todo = (1u << (FUSE_GROUPS / FUSE_CONSUMERS)) - 1u;
            while (todo) {
                const int j = __ffs((int)todo) - 1;
                todo &= todo - 1;
                const int k = k0 + j * 32 * FUSE_CONSUMERS;
                const bool live = k < npts;
                const float4 p = sp[j * 32 * FUSE_CONSUMERS];
                fuse_warp_points(cam, A, p, live, (mb >> j) & 1u, seg0 + k, lane, M, B, m0, mask_base, plane32, s_box,
                                 s_inside, s_nin, s_total, my_valid);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty_bar[s]);
        }
        const unsigned vb = __reduce_add_sync(0xFFFFFFFFu, (unsigned)my_valid);
        if (lane == 0 && vb) atomicAdd(s_nvalid, (int)vb);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < M * B; i += FUSE_THREADS) {
        const int c = i / B, b = i - c * B;
        const int v = s_inside[c * A.smem_boxes + b];
        if (v) atomicAdd(A.inside + A.cnt_off[f] + i, v);
    }
    for (int c = threadIdx.x; c < M; c += FUSE_THREADS) {
        const int v = s_total[c];
        if (v) atomicAdd(A.total + m0 + c, v);
    }
    for (int b = threadIdx.x; b < B; b += FUSE_THREADS) {
        const int v = s_nin[b];
        if (v) atomicAdd(A.n_in_box + b0 + b, v);
    }
    if (threadIdx.x == 0 && *s_nvalid) atomicAdd(A.n_valid + f, *s_nvalid);
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
    const size_t smem = (size_t)FUSE_STAGES * FUSE_SUB * sizeof(float4) + (size_t)A.smem_boxes * sizeof(lf_box) +
                        4 * ((size_t)A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 4);
    if (smem > 200 * 1024) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaFuncSetAttribute(lf_fuse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t chunks = (max_points_per_frame + FUSE_SEG - 1) / FUSE_SEG;
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
