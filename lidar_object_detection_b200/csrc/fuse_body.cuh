// K1+K4+K5 fused point pass as a per-item device function, shared by the stand-alone
// kernel (fuse.cu) and the persistent batch kernel (batch.cu).
#pragma once
#include "common.cuh"
#include "bulk.cuh"

namespace {


#ifndef LF_FUSE_WARPS
#define LF_FUSE_WARPS 8
#endif
constexpr int FUSE_CONSUMERS = LF_FUSE_WARPS;             // consumer warps
constexpr int FUSE_THREADS = 32 * FUSE_CONSUMERS;         // stand-alone kernel: every warp consumes
constexpr int BATCH_THREADS = FUSE_THREADS + 32;          // persistent kernel: + one scheduler warp
#ifndef LF_FUSE_SUB
#define LF_FUSE_SUB 1536         /* persistent kernel (batch.cu); the stand-alone launch has its own geometries */
#endif
#ifndef LF_FUSE_STAGES
#define LF_FUSE_STAGES 2
#endif
#ifndef LF_FUSE_SEG
#define LF_FUSE_SEG 12288
#endif
#ifndef LF_FUSE_CTAS
#define LF_FUSE_CTAS 4
#endif
constexpr int FUSE_SUB = LF_FUSE_SUB;                     // points per stage = 24 KB (a multiple of 256: whole groups per warp)
constexpr int FUSE_GROUPS = FUSE_SUB / 32;                // 32-point groups per stage, handed out dynamically
constexpr int FUSE_STAGES = LF_FUSE_STAGES;
constexpr int FUSE_SUBS_PER_SEG = LF_FUSE_SEG / FUSE_SUB;
constexpr int FUSE_SEG = FUSE_SUB * FUSE_SUBS_PER_SEG;    // 12288 points per CTA
static_assert(FUSE_SUB % (32 * FUSE_CONSUMERS) == 0 && FUSE_SUB / (32 * FUSE_CONSUMERS) <= 32, "stage = whole groups per warp");
static_assert(LF_FUSE_SEG % FUSE_SUB == 0, "segment = whole stages");

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
    int seg_points;      // points per work item: a multiple of FUSE_SUB, FUSE_SEG for large batches
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

// Everything one warp does for 32 consecutive points (one per lane) of which at
// least one passed the pre-gate (`maybe`).
__device__ __forceinline__ void fuse_warp_points(const LfCam& cam, const FuseArgs& A, float4 p, bool live, bool maybe,
                                                 int64_t gidx, int lane, int M, int B, const int4* s_mbb,
                                                 const uint32_t* mask_base, int plane32, const lf_box* s_box,
                                                 int32_t* s_inside,
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
        // pixel bounds of the warp's valid points against the masks' bounding boxes (staged in
        // shared memory per frame; a mask without a box record covers the whole image)
        const int u0 = __reduce_min_sync(FULL, valid ? ui : 0x7fffffff);
        const int u1 = __reduce_max_sync(FULL, valid ? ui : -1);
        const int v0 = __reduce_min_sync(FULL, valid ? vi : 0x7fffffff);
        const int v1 = __reduce_max_sync(FULL, valid ? vi : -1);
        for (int g = 0; g < M; g += 32) {
            const int cl = g + lane;
            bool cand = false;
            if (cl < M) {
                const int4 bb = s_mbb[cl];
                cand = u1 >= bb.x && u0 <= bb.z && v1 >= bb.y && v0 <= bb.w;
            }
            unsigned cmask = __ballot_sync(FULL, cand);
            // up to four candidate masks per round: their words are requested together, so a
            // group pays one L2 round trip, not one per mask
            while (cmask) {
                int c[4];
                uint32_t w[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    c[j] = cmask ? g + __ffs((int)cmask) - 1 : -1;
                    cmask &= cmask - 1;
                }
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    w[j] = (valid && c[j] >= 0) ? __ldg(mask_base + (unsigned)(off + c[j] * plane32)) : 0u;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c[j] >= 0) bits |= (unsigned long long)((w[j] >> sh) & 1u) << c[j];
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

// One work item = (frame f, segment `seg` of FUSE_SEG points), executed by all NT threads
// of the CTA; warps 0..FUSE_CONSUMERS-1 consume, further warps only take part in the block
// barriers.  The segment streams through a ring of FUSE_STAGES bulk-copy stages.  There is
// no producer warp: each consumer warp counts itself out of a stage on a shared counter
// and the LAST one out refills that stage, so nobody ever waits for a stage to drain.
// `full_bar` (count 1) is initialised once per CTA and `stage_cnt` zeroed; the barrier
// phases run on from item to item: `fbase` counts the sub-chunks this CTA has streamed so
// far, sub-chunk c lives in stage c % FUSE_STAGES.  `hook()` is called once by thread
// 32 * FUSE_CONSUMERS, if the CTA has one (the persistent kernel claims its next work item
// there, overlapped with the item).
struct FuseNoHook { __device__ __forceinline__ void operator()() const {} };

template <int NT, class Hook, int SUB = FUSE_SUB, int WARPS = FUSE_CONSUMERS>
__device__ __forceinline__ void fuse_item(const LfCam& cam, const FuseArgs& A, int f, int seg, uint8_t* smem,
                                          uint64_t* full_bar, int* stage_cnt, uint32_t& fbase, Hook hook) {
    static_assert(SUB % (32 * WARPS) == 0 && SUB / (32 * WARPS) <= 32, "stage = whole groups per warp");
    const int64_t p0 = A.pt_off[f], p1 = A.pt_off[f + 1];
    const int64_t seg0 = p0 + (int64_t)seg * A.seg_points;
    if (seg0 >= p1) {
        if (threadIdx.x == 32 * WARPS) hook();
        return;
    }
    const int npts = (int)min((int64_t)A.seg_points, p1 - seg0);
    const int nsub = (npts + SUB - 1) / SUB;
    const int ngroups = (npts + 31) / 32;
    const int m0 = A.mask_off[f];
    const int M = A.mask_off[f + 1] - m0;
    const int b0 = A.box_off[f];
    const int B = A.box_off[f + 1] - b0;

    float4* stage = (float4*)smem;                                               // [FUSE_STAGES][SUB]
    int4* s_mbb = (int4*)(smem + FUSE_STAGES * SUB * sizeof(float4));      // [smem_cars] {x0, y0, x1, y1}
    lf_box* s_box = (lf_box*)(s_mbb + A.smem_cars);
    int32_t* s_inside = (int32_t*)((uint8_t*)s_box + (size_t)A.smem_boxes * sizeof(lf_box));
    int32_t* s_nin = s_inside + (size_t)A.smem_cars * A.smem_boxes;
    int32_t* s_total = s_nin + A.smem_boxes;
    int32_t* s_nvalid = s_total + A.smem_cars;
    const float4* src = A.pts + seg0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        fence_proxy_async_smem();                          // the previous item read these stages
        for (int i = 0; i < FUSE_STAGES && i < nsub; ++i) {
            const int s = (int)((fbase + i) % FUSE_STAGES);
            const uint32_t bytes = (uint32_t)min(SUB, npts - i * SUB) * 16u;
            mbar_expect_tx(&full_bar[s], bytes);
            bulk_g2s(stage + s * SUB, src + (int64_t)i * SUB, bytes, &full_bar[s]);
        }
    }
    {   // while the first sub-chunks fly: stage the frame's boxes, clear the private counters
        const double* gsrc = (const double*)(A.boxes + b0);
        double* sdst = (double*)s_box;
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += NT) sdst[i] = gsrc[i];
        const int nc = A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 1;
        for (int i = threadIdx.x; i < nc; i += NT) s_inside[i] = 0;
        for (int c = threadIdx.x; c < M; c += NT) {
            int4 bb = make_int4(0, 0, cam.W - 1, cam.H - 1);
            if (A.mask_bbox) {                             // record of lf_erode_pack: zeros = empty mask
                const int4 r = __ldg((const int4*)A.mask_bbox + m0 + c);
                bb = make_int4((1 << 20) - r.x, (1 << 20) - r.y, r.z - 1, r.w - 1);
            }
            s_mbb[c] = bb;
        }
    }
    __syncthreads();

    if (warp >= WARPS) {
        if (threadIdx.x == 32 * WARPS) hook();
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
            const uint32_t c = fbase + (uint32_t)sub;
            const int s = (int)(c % FUSE_STAGES);
            mbar_wait(&full_bar[s], (c / FUSE_STAGES) & 1u);
            const float4* sp = stage + s * SUB + warp * 32 + lane;
            const int k0 = sub * SUB + warp * 32 + lane;          // point index inside the segment
            unsigned mb = 0;
#pragma unroll
            for (int j = 0; j < (SUB / 32 / WARPS); ++j) {
                const float4 p = sp[j * 32 * WARPS];
                const bool live = k0 + j * 32 * WARPS < npts;
                mb |= (unsigned)(live & pg.maybe(p)) << j;
            }
            unsigned todo = __reduce_or_sync(0xFFFFFFFFu, mb);
            if (all_groups) todo = (1u << ((SUB / 32 / WARPS))) - 1u;
            while (todo) {
                const int j = __ffs((int)todo) - 1;
                todo &= todo - 1;
                const int k = k0 + j * 32 * WARPS;
                const bool live = k < npts;
                const float4 p = sp[j * 32 * WARPS];
                fuse_warp_points(cam, A, p, live, (mb >> j) & 1u, seg0 + k, lane, M, B, s_mbb, mask_base, plane32, s_box,
                                 s_inside, s_nin, s_total, my_valid);
            }
            __syncwarp();
            if (lane == 0 && atomicAdd(&stage_cnt[s], 1) == WARPS - 1) {
                // last warp out of this stage: every warp's reads are done -> refill it
                atomicExch(&stage_cnt[s], 0);
                const int i = sub + FUSE_STAGES;
                if (i < nsub) {
                    const uint32_t bytes = (uint32_t)min(SUB, npts - i * SUB) * 16u;
                    fence_proxy_async_smem();
                    mbar_expect_tx(&full_bar[s], bytes);
                    bulk_g2s(stage + s * SUB, src + (int64_t)i * SUB, bytes, &full_bar[s]);
                }
            }
        }
        const unsigned vb = __reduce_add_sync(0xFFFFFFFFu, (unsigned)my_valid);
        if (lane == 0 && vb) atomicAdd(s_nvalid, (int)vb);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < M * B; i += NT) {
        const int c = i / B, b = i - c * B;
        const int v = s_inside[c * A.smem_boxes + b];
        if (v) atomicAdd(A.inside + A.cnt_off[f] + i, v);
    }
    for (int c = threadIdx.x; c < M; c += NT) {
        const int v = s_total[c];
        if (v) atomicAdd(A.total + m0 + c, v);
    }
    for (int b = threadIdx.x; b < B; b += NT) {
        const int v = s_nin[b];
        if (v) atomicAdd(A.n_in_box + b0 + b, v);
    }
    if (threadIdx.x == 0 && *s_nvalid) atomicAdd(A.n_valid + f, *s_nvalid);
    fbase += (uint32_t)nsub;
}

}  // namespace
