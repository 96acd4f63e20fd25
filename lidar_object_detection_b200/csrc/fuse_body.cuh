// K1+K4+K5 fused point pass (cvs_erosion.py:342-372 without the erosion): device code of
// lf_fuse_kernel (fuse.cu).
//
// Persistent, independent warps; register streaming; no CTA-level hand-off of any kind.
//   claims   the batch is cut into claims of consecutive points of one frame; every warp of the
//            (machine-filling) grid takes the next claim from one global counter until none is
//            left, so warps never wait for each other and the load balances itself.  Claim ids
//            walk the frames round-robin (chunk 0 of every frame, then chunk 1, ...).
//   phase A  per iteration a warp owns 32 * PER consecutive points: every thread keeps PER
//            independent 128-bit streaming loads in flight, then classifies each 32-point group
//            in fp32 in three stages of rising cost, each ended by a warp vote so that groups
//            without a survivor stop early:
//              1. depth row only                    -> behind the camera / beyond the range gate
//              2. centred image rows                -> more than a pixel outside the image
//              3. pixel = rint(x/d) in fp32 + a residual test that PROVES it equal to the
//                 reference's float64 result        -> "certain" (pixel known) or "uncertain"
//            Survivors are compacted (ballot + popc prefix) into two per-warp shared-memory
//            queues: certain points with their pixel, uncertain ones for the exact chain.
//            Consecutive queue entries are consecutive survivors of the sweep, so a group of 64
//            covers a short arc of one ring: tight pixel and xyz bounds for the culls below.
//   exact    whenever 32 uncertain points are queued the warp runs the float64 dgemm-order
//            chain and the two divisions on all 32 lanes (common.cuh) and moves the points
//            that pass the gate of cvs_erosion.py:359 to the certain queue.
//   phase B  groups of 64 certain points, two per lane: packed-mask gather (L2) behind a mask
//            bounding-box cull, box cull, float64 containment, ballot/popc counts added to the
//            global counters with one reduction-atomic per (car, box) and group.  (Two points per
//            lane: bounds, culls, candidate loops and atomics are paid once per 64 points and the
//            two points' loads overlap; against 32-point groups: cfg3 0.650 -> 0.635 ms, cfg5
//            0.863 -> 0.830, the 20 real frames with up to 314 boxes 0.306 -> 0.262.)
// Nothing fp32 ever decides a result: a point is "certain" only if the fp32 residual proves
// the rounded pixel and the gate (error bounds: lf_set_camera); everything else takes the exact path.
// Measured and dropped: stage 3 as a separate pass over a candidate queue, and a
// per-frame bit map of where masks and boxes are in front of phase B -- the group-level culls of
// phase B already skip what such a map skips, and the extra pass costs more than it saves.
#pragma once
#include "common.cuh"

#ifndef LF_FUSE_GATHER
#define LF_FUSE_GATHER 2           /* mask words requested together per round of the gather */
#endif
#ifndef LF_FUSE_PFDIST
#define LF_FUSE_PFDIST 1           /* iterations the L2 prefetch runs ahead of the loads */
#endif

namespace {

struct FuseArgs {
    const float4* pts;
    const int64_t* pt_off;
    const uint32_t* packed;
    const int32_t* mask_bbox;   // optional [sum M][4] from lf_erode_pack, NULL = test every mask
    const uint32_t* row_map;    // non-NULL: the packed masks come from lf_erode_pack_rows -- only the rows inside a mask's bounding box are valid
    int rm_wpr;
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
    float dmax_hi;       // smallest float >= max_depth (stage 1: certainly beyond the gate)
    float dmax_lo;       // largest float < max_depth    (stage 3: certainly inside the gate)
    int use_oriented;
    int force_exact;     // verification mode: every candidate takes the float64 chain
    int smem_boxes;      // per-warp cache of box cull bounds (max boxes over frames); 0: frames with many boxes read them
                         // from the lf_box records through L1 instead (no per-claim reload, no shared memory)
    int smem_cars;       // per-warp cache of mask bounding boxes (max masks per frame, <= LF_MAX_CARS)
    unsigned int* claim; // global claim counter, zeroed before the launch
    int F;               // frames of this launch
    int cpf;             // claims per frame = ceil(max points per frame / claim_points)
    int claim_points;    // points per claim: a multiple of 32 * PER
    int group;           // masks 64*group .. 64*group+63 of every frame (frames with more than LF_MAX_CARS masks take several passes)
};

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
__device__ __forceinline__ float rcp_approx(float v) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
    return r;
}

// closed-interval slab test 0 <= dot(rel,e)/ee <= 1 (cvs_erosion.py:133-138).
// d/ee <= 1 <=> d <= ee and d/ee >= 0 <=> d >= 0 for ee > 0, except when a
// negative d/ee underflows to -0 (which numpy accepts): that corner keeps the
// division.  ee == 0 gives inf/nan in numpy -> outside.
__device__ __forceinline__ bool slab(double rx, double ry, double rz, double e0, double e1, double e2, double ee) {
    double d = __dmul_rn(rx, e0);
    d = __fma_rn(ry, e1, d);
    d = __fma_rn(rz, e2, d);
    if (!(ee > 0.0)) return false;
    if (d >= 0.0) return d <= ee;
    return __ddiv_rn(d, ee) >= 0.0;
}

// One box read through the read-only path (all lanes the same address: one L1 transaction per load).
__device__ __forceinline__ bool point_in_box(const lf_box* bx, float x, float y, float z, int oriented) {
    const double X = (double)x, Y = (double)y, Z = (double)z;
    const double2* p = (const double2*)bx;                 // lf_box is 192 B, 16-B aligned in the array
    if (oriented) {
        const double2 a0 = __ldg(p), a1 = __ldg(p + 1), a2 = __ldg(p + 2), a3 = __ldg(p + 3), a4 = __ldg(p + 4),
                      a5 = __ldg(p + 5), a6 = __ldg(p + 6), a7 = __ldg(p + 7);
        const double rx = X - a0.x, ry = Y - a0.y, rz = Z - a1.x;                     // c0
        return slab(rx, ry, rz, a1.y, a2.x, a2.y, a6.x) &&                            // e1, ee[0]
               slab(rx, ry, rz, a3.x, a3.y, a4.x, a6.y) &&                            // e2, ee[1]
               slab(rx, ry, rz, a4.y, a5.x, a5.y, a7.x);                              // e3, ee[2]
    }
    const double2 a7 = __ldg(p + 7), a8 = __ldg(p + 8), a9 = __ldg(p + 9), a10 = __ldg(p + 10);
    return X >= a7.y && X <= a9.x && Y >= a8.x && Y <= a9.y && Z >= a8.y && Z <= a10.x;   // amin, amax
}

// What a warp knows about the frame it is working on (per-warp shared-memory caches + CSR entries).
struct FuseFrame {
    const int4* mbb;      // [M] mask bounding boxes {x0, y0, x1, y1} (shared memory)
    const float* cull;    // [6][SB] lo.x lo.y lo.z hi.x hi.y hi.z of the frame's boxes (shared memory, one box per lane)
    int SB;
    int M, B;
    const uint32_t* mask_base;
    const lf_box* boxes;
    int32_t* total;       // global counters of this frame
    int32_t* inside;
    int32_t* nin;
};

// lo.x lo.y lo.z hi.x hi.y hi.z of box b: from the warp's shared-memory cache, or (frames with many boxes)
// from the lf_box record itself -- bytes 168..191 of it, one 32-byte sector, L1-cached
__device__ __forceinline__ void cull_bounds(const FuseFrame& Fr, int b, float (&c)[6]) {
    if (Fr.SB > 0) {
        const float* cu = Fr.cull + b;
#pragma unroll
        for (int k = 0; k < 6; ++k) c[k] = cu[k * Fr.SB];
    } else {
        const float2* lh = (const float2*)(Fr.boxes[b].lo);
        const float2 t0 = __ldg(lh), t1 = __ldg(lh + 1), t2 = __ldg(lh + 2);
        c[0] = t0.x; c[1] = t0.y; c[2] = t1.x; c[3] = t1.y; c[4] = t2.x; c[5] = t2.y;
    }
}

// counts of one 32-mask block for two points per lane: one reduction-atomic per car present in the 64 points
__device__ __forceinline__ void add_per_car2(unsigned blk0, unsigned blk1, bool p0, bool p1, int32_t* dst, int stride,
                                             int lane) {
    unsigned un = __reduce_or_sync(0xFFFFFFFFu, (p0 ? blk0 : 0u) | (p1 ? blk1 : 0u));
    while (un) {
        const int c = __ffs((int)un) - 1;
        un &= un - 1;
        const unsigned b0 = __ballot_sync(0xFFFFFFFFu, p0 && ((blk0 >> c) & 1u));
        const unsigned b1 = __ballot_sync(0xFFFFFFFFu, p1 && ((blk1 >> c) & 1u));
        if (lane == 0) atomicAdd(dst + (int64_t)c * stride, __popc(b0) + __popc(b1));
    }
}

// ---- phase B on 64 work-queue points, two per lane (lane l: entries l and l + 32): the pixel / xyz bounds, both
// culls, the candidate loops and the counter atomics are paid once per 64 points instead of once per 32.
template <bool ROWS>
__device__ __forceinline__ void fuse_group2(const LfCam& cam, const FuseArgs& A, const FuseFrame& Fr, float4 e0, float4 e1,
                                            bool have0, bool have1, int plane32, unsigned long long& bits0,
                                            unsigned long long& bits1) {
    const unsigned FULL = 0xFFFFFFFFu;
    const int lane = threadIdx.x & 31;
    const unsigned uv0 = __float_as_uint(e0.w), uv1 = __float_as_uint(e1.w);
    const int ui0 = (int)(uv0 & 0xFFFFu), vi0 = (int)(uv0 >> 16), ui1 = (int)(uv1 & 0xFFFFu), vi1 = (int)(uv1 >> 16);
    const int M = Fr.M, B = Fr.B;
    unsigned blk0[2] = {0u, 0u}, blk1[2] = {0u, 0u};         // membership bits of the two points, masks 0..31 and 32..63
    if (M > 0) {
        const int off0 = vi0 * cam.wpr + (ui0 >> 5), off1 = vi1 * cam.wpr + (ui1 >> 5);
        const int sh0 = ui0 & 31, sh1 = ui1 & 31;
        const int u0 = __reduce_min_sync(FULL, min(have0 ? ui0 : 0x7fffffff, have1 ? ui1 : 0x7fffffff));
        const int u1 = __reduce_max_sync(FULL, max(have0 ? ui0 : -1, have1 ? ui1 : -1));
        const int v0 = __reduce_min_sync(FULL, min(have0 ? vi0 : 0x7fffffff, have1 ? vi1 : 0x7fffffff));
        const int v1 = __reduce_max_sync(FULL, max(have0 ? vi0 : -1, have1 ? vi1 : -1));
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int g = 32 * h;
            if (g >= M) break;
            const int cl = g + lane;
            bool cand = false;
            if (cl < M) {
                const int4 bb = Fr.mbb[cl];
                cand = u1 >= bb.x && u0 <= bb.z && v1 >= bb.y && v0 <= bb.w;
            }
            unsigned cmask = __ballot_sync(FULL, cand);
            unsigned acc0 = 0u, acc1 = 0u;
            while (cmask) {
                int c[LF_FUSE_GATHER];
                uint32_t w0[LF_FUSE_GATHER], w1[LF_FUSE_GATHER];
#pragma unroll
                for (int j = 0; j < LF_FUSE_GATHER; ++j) {
                    c[j] = cmask ? __ffs((int)cmask) - 1 : -1;
                    cmask &= cmask - 1;
                }
#pragma unroll
                for (int j = 0; j < LF_FUSE_GATHER; ++j) {
                    bool g0 = have0 && c[j] >= 0, g1 = have1 && c[j] >= 0;
                    if (ROWS && A.row_map && c[j] >= 0) {
                        // packed masks of lf_erode_pack_rows: only the rows from the mask's first to its last row with
                        // a set pixel hold valid words (the bounding box says which), the others were never written
                        const int4 bb = Fr.mbb[g + c[j]];
                        g0 &= vi0 >= bb.y && vi0 <= bb.w;
                        g1 &= vi1 >= bb.y && vi1 <= bb.w;
                    }
                    w0[j] = g0 ? __ldg(Fr.mask_base + (unsigned)(off0 + (g + c[j]) * plane32)) : 0u;
                    w1[j] = g1 ? __ldg(Fr.mask_base + (unsigned)(off1 + (g + c[j]) * plane32)) : 0u;
                }
#pragma unroll
                for (int j = 0; j < LF_FUSE_GATHER; ++j)
                    if (c[j] >= 0) {
                        acc0 |= ((w0[j] >> sh0) & 1u) << c[j];
                        acc1 |= ((w1[j] >> sh1) & 1u) << c[j];
                    }
            }
            blk0[h] = acc0;
            blk1[h] = acc1;
            add_per_car2(acc0, acc1, true, true, Fr.total + g, 1, lane);
        }
    }
    bits0 = ((unsigned long long)blk0[1] << 32) | blk0[0];
    bits1 = ((unsigned long long)blk1[1] << 32) | blk1[0];
    if (B == 0) return;
    const float INF = __int_as_float(0x7f800000);
    const float wx0 = warp_min_f32(fminf(have0 ? e0.x : INF, have1 ? e1.x : INF));
    const float wx1 = warp_max_f32(fmaxf(have0 ? e0.x : -INF, have1 ? e1.x : -INF));
    const float wy0 = warp_min_f32(fminf(have0 ? e0.y : INF, have1 ? e1.y : INF));
    const float wy1 = warp_max_f32(fmaxf(have0 ? e0.y : -INF, have1 ? e1.y : -INF));
    const float wz0 = warp_min_f32(fminf(have0 ? e0.z : INF, have1 ? e1.z : INF));
    const float wz1 = warp_max_f32(fmaxf(have0 ? e0.z : -INF, have1 ? e1.z : -INF));
    for (int g = 0; g < B; g += 32) {
        const int bl = g + lane;
        bool cand = false;
        if (bl < B) {
            float c[6];
            cull_bounds(Fr, bl, c);
            cand = wx1 >= c[0] && wx0 <= c[3] && wy1 >= c[1] && wy0 <= c[4] && wz1 >= c[2] && wz0 <= c[5];
        }
        unsigned cmask = __ballot_sync(FULL, cand);
        while (cmask) {
            const int b = g + __ffs((int)cmask) - 1;
            cmask &= cmask - 1;
            float c[6];
            cull_bounds(Fr, b, c);
            const bool near0 = have0 && e0.x >= c[0] && e0.x <= c[3] && e0.y >= c[1] && e0.y <= c[4] && e0.z >= c[2] && e0.z <= c[5];
            const bool near1 = have1 && e1.x >= c[0] && e1.x <= c[3] && e1.y >= c[1] && e1.y <= c[4] && e1.z >= c[2] && e1.z <= c[5];
            if (!__any_sync(FULL, near0 | near1)) continue;
            const bool in0 = near0 && point_in_box(Fr.boxes + b, e0.x, e0.y, e0.z, A.use_oriented);
            const bool in1 = near1 && point_in_box(Fr.boxes + b, e1.x, e1.y, e1.z, A.use_oriented);
            const unsigned bal0 = __ballot_sync(FULL, in0), bal1 = __ballot_sync(FULL, in1);
            if (!(bal0 | bal1)) continue;
            if (lane == 0 && Fr.nin) atomicAdd(Fr.nin + b, __popc(bal0) + __popc(bal1));
            add_per_car2(blk0[0], blk1[0], in0, in1, Fr.inside + b, B, lane);
            if (M > 32) add_per_car2(blk0[1], blk1[1], in0, in1, Fr.inside + (int64_t)32 * B + b, B, lane);
        }
    }
}

// ---- exact path: up to 32 queued uncertain points (one per lane) through the float64 dgemm-order
// chain and the two divisions (common.cuh); points that pass the gate of cvs_erosion.py:359 are
// appended to the work queue.  Rare (a few per cent of the candidates).  Returns the new fill of
// the work queue.
template <bool BITS>
__device__ __noinline__ int fuse_resolve(const LfCam& cam, const FuseArgs& A, const float4* uq, int cnt, float4* q,
                                         int* qi, int qn, int64_t seg0) {
    __syncwarp();
    const int lane = threadIdx.x & 31;
    const bool have = lane < cnt;
    const float4 e = uq[have ? lane : 0];
    double uqd, vqd, d;
    lf_project_point(cam, e.x, e.y, e.z, uqd, vqd, d);
    const bool ok = have && lf_gate(cam, uqd, vqd, d, A.max_depth);
    const unsigned mk = __ballot_sync(0xFFFFFFFFu, ok);
    if (ok) {
        const int pos = qn + __popc(mk & ((1u << lane) - 1u));
        q[pos] = make_float4(e.x, e.y, e.z, __uint_as_float((unsigned)(int)uqd | ((unsigned)(int)vqd << 16)));
        if (BITS) qi[pos] = __float_as_int(e.w);
    } else if (BITS && have) {
        A.car_bits[seg0 + __float_as_int(e.w)] = 0ULL;
    }
    __syncwarp();
    return qn + __popc(mk);
}


// The points of iteration `it` of a claim into registers (dead lanes of the last iteration are masked by `k < npts`
// in stage 1), and PER * 4 lines (128 B each) of a later iteration, one per lane, towards L2.
template <int PER>
__device__ __forceinline__ void fuse_load(float4 (&P)[PER], const float4* src, int it, int npts, int lane) {
    constexpr int SPAN = PER * 32;
    const int kbase = it * SPAN + lane;
    const int kn = (it + LF_FUSE_PFDIST) * SPAN + lane * 8;
    if (lane < PER * 4 && kn < npts) asm volatile("prefetch.global.L2 [%0];" ::"l"(src + kn));
    if (it * SPAN + SPAN <= npts) {
#pragma unroll
        for (int j = 0; j < PER; ++j) P[j] = ld_stream_f4(src + kbase + 32 * j);
    } else {
#pragma unroll
        for (int j = 0; j < PER; ++j)
            P[j] = kbase + 32 * j < npts ? ld_stream_f4(src + kbase + 32 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
}

// Geometry of one warp's working set.
//   WARPS  warps per CTA (independent workers; the CTA only shares the shared-memory allocation)
//   PER    128-bit loads in flight per thread and iteration (an iteration covers PER*32 consecutive points)
//   BITS   per-point membership output (car_bits) wanted: queues carry the point index too
template <int WARPS, int PER, bool BITS>
struct FuseGeom {
    static constexpr int NT = 32 * WARPS;
    static constexpr int SPAN = PER * 32;                  // points per iteration of one warp
    static constexpr int GRP = 64;                         // queue entries phase B takes at a time (two per lane)
    static constexpr int QCAP = PER * 32 + 32 + GRP;       // certain queue: < GRP left over + PER*32 new + < 32 resolved
    static constexpr int UCAP = 64;                        // uncertain queue: < 32 left over + 32 new
    static constexpr size_t queue_bytes = (size_t)(QCAP + UCAP) * 16 + (BITS ? (size_t)QCAP * 4 : 0);
    // + per-warp frame cache: mask bounding boxes (16 B each) and box cull bounds (24 B each)
    static __host__ __device__ size_t warp_bytes(int cars, int boxes) {
        return (queue_bytes + 16 * (size_t)cars + 24 * (size_t)boxes + 8 + 15) & ~(size_t)15;
    }
};

//   ROWS   the packed masks may come with a row map (lf_erode_pack_rows): the gather consults it
template <int WARPS, int PER, bool BITS, bool ROWS>
__device__ __forceinline__ void fuse_warp_loop(const LfCam& cam, const FuseArgs& A, uint8_t* smem) {
    using G = FuseGeom<WARPS, PER, BITS>;
    constexpr int SPAN = G::SPAN, QCAP = G::QCAP, UCAP = G::UCAP;
    const unsigned FULL = 0xFFFFFFFFu;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint8_t* ws = smem + (size_t)warp * G::warp_bytes(A.smem_cars, A.smem_boxes);
    float4* q = (float4*)ws;                               // [QCAP] certain: x, y, z, bits(u | v << 16)
    float4* uq = q + QCAP;                                 // [UCAP] uncertain: x, y, z, bits(index in claim)
    int4* s_mbb = (int4*)(uq + UCAP);                      // [smem_cars]
    float* s_cull = (float*)(s_mbb + A.smem_cars);         // [6][smem_boxes]
    int* qi = (int*)(s_cull + 6 * (size_t)A.smem_boxes);   // [QCAP] index in claim of the certain entries (BITS)

    const int plane32 = cam.H * cam.wpr;                   // words per packed mask (< 2^25, see lf_set_camera)
    const unsigned lt = (1u << lane) - 1u;
    // fp32 classification constants (lf_set_camera)
    const float kT = cam.kT, kTs = cam.kTs, gS = cam.gS, hwx = cam.hwx, hwy = cam.hwy;
    const float cxm = cam.cxm, cym = cam.cym, umax = (float)(cam.W - 1), vmax = (float)(cam.H - 1);
    const float dhi = A.dmax_hi, dlo = A.dmax_lo;
    const bool fast = !A.force_exact;
    const unsigned n_claims = (unsigned)A.F * (unsigned)A.cpf;

    FuseFrame Fr;
    Fr.mbb = s_mbb; Fr.cull = s_cull; Fr.SB = A.smem_boxes;
    Fr.M = Fr.B = 0; Fr.mask_base = nullptr; Fr.boxes = nullptr; Fr.total = Fr.inside = Fr.nin = nullptr;
    int cur_f = -1;

    unsigned id = 0;
    if (lane == 0) id = atomicAdd(A.claim, 1u);
    id = __shfl_sync(FULL, id, 0);
    while (id < n_claims) {
        unsigned nxt = 0;
        if (lane == 0) nxt = atomicAdd(A.claim, 1u);        // next claim requested now, consumed after this one
        // claim id -> (frame id % F, chunk id / F): round-robin over the frames, so that the warps running at one
        // time are spread over every frame of the batch instead of the ~70 neighbouring ones that consecutive
        // chunks of consecutive frames would give (cfg3: 0.636 -> 0.595 ms; cfg5 and the real frames: unchanged
        // within 3 %; measured with windows of 16 / 64 / 256 frames in between: 0.644 / 0.636 / 0.615)
        const unsigned chunk = id / (unsigned)A.F;
        const int f = (int)(id - chunk * (unsigned)A.F);
        const int64_t c0 = (int64_t)chunk * A.claim_points;
        const int64_t p0 = A.pt_off[f], nf = A.pt_off[f + 1] - p0;
        if (c0 < nf) {
            const int npts = (int)min((int64_t)A.claim_points, nf - c0);
            const int64_t seg0 = p0 + c0;
            const float4* src = A.pts + seg0;
            if (f != cur_f) {                               // frame state into the warp's cache
                cur_f = f;
                const int b0 = A.box_off[f];
                const int m0 = min(A.mask_off[f] + LF_MAX_CARS * A.group, A.mask_off[f + 1]);
                Fr.M = min(A.mask_off[f + 1] - m0, LF_MAX_CARS);
                Fr.B = A.box_off[f + 1] - b0;
                Fr.mask_base = A.packed + (int64_t)m0 * plane32;
                Fr.boxes = A.boxes + b0;
                Fr.total = A.total + m0;
                Fr.nin = A.n_in_box ? A.n_in_box + b0 : nullptr;      // passes after the first count per-car only
                Fr.inside = A.inside + A.cnt_off[f] + (int64_t)LF_MAX_CARS * A.group * Fr.B;
                __syncwarp();
                for (int c = lane; c < Fr.M; c += 32) {
                    int4 bb = make_int4(0, 0, cam.W - 1, cam.H - 1);
                    if (A.mask_bbox) {                      // record of lf_erode_pack: zeros = empty mask
                        const int4 r = __ldg((const int4*)A.mask_bbox + m0 + c);
                        bb = make_int4((1 << 20) - r.x, (1 << 20) - r.y, r.z - 1, r.w - 1);
                    }
                    s_mbb[c] = bb;
                }
                for (int b = lane; A.smem_boxes > 0 && b < Fr.B; b += 32) {
                    const float2* lh = (const float2*)(Fr.boxes[b].lo);    // lo[3], hi[3]: 8-B aligned
                    const float2 t0 = __ldg(lh), t1 = __ldg(lh + 1), t2 = __ldg(lh + 2);
                    s_cull[b] = t0.x; s_cull[A.smem_boxes + b] = t0.y; s_cull[2 * A.smem_boxes + b] = t1.x;
                    s_cull[3 * A.smem_boxes + b] = t1.y; s_cull[4 * A.smem_boxes + b] = t2.x;
                    s_cull[5 * A.smem_boxes + b] = t2.y;
                }
                __syncwarp();
            }
            int qn = 0, un = 0;                             // queue fill, warp-uniform
            int nval = 0;                                   // certain points consumed in this claim
            // phase B on the last GRP (or fewer, at the very end) entries of the certain queue
            auto consume = [&](int cnt) {
                const bool have0 = lane < cnt, have1 = lane + 32 < cnt;
                const int at0 = qn - cnt + (have0 ? lane : 0), at1 = qn - cnt + (have1 ? lane + 32 : 0);
                const float4 e0 = q[at0], e1 = q[at1];
                unsigned long long bits0, bits1;
                fuse_group2<ROWS>(cam, A, Fr, e0, e1, have0, have1, plane32, bits0, bits1);
                if (BITS && have0) A.car_bits[seg0 + qi[at0]] = bits0;
                if (BITS && have1) A.car_bits[seg0 + qi[at1]] = bits1;
                nval += cnt;
                qn -= cnt;
            };
            const int niter = (npts + SPAN - 1) / SPAN;
            for (int it = 0;; ++it) {
                const bool last = it == niter;              // pseudo-iteration: flush both queues
                if (!last) {
                    const int kbase = it * SPAN + lane;     // group j: kbase + 32 j
                    float4 P[PER];
                    fuse_load<PER>(P, src, it, npts, lane);
                    // ---- stage 1: depth row.  tol bounds every fp32 evaluation error of this point
                    // (+inf beyond 2^20 m and for NaN / inf input: such points are never decided here).
                    auto stage1 = [&](const float4& p, int k, float& d, float& tol) -> bool {
                        d = fmaf(cam.Pg[8], p.x, fmaf(cam.Pg[9], p.y, fmaf(cam.Pg[10], p.z, cam.Pg[11])));
                        tol = fmaf((fabsf(p.x) + fabsf(p.y) + fabsf(p.z)) * gS, kTs, kT);
                        return !(d < -tol) & !(d > dhi + tol) & (k < npts);                 // NaN stays in
                    };
                    // stages 2 and 3 of a group that has a stage-1 survivor, survivors into the queues
                    auto stages23 = [&](const float4& p, int k, float d, float tol, bool okd) {
                        // ---- stage 2: centred image rows, one pixel of margin
                        const float x = fmaf(cam.Pg[0], p.x, fmaf(cam.Pg[1], p.y, fmaf(cam.Pg[2], p.z, cam.Pg[3])));
                        const float y = fmaf(cam.Pg[4], p.x, fmaf(cam.Pg[5], p.y, fmaf(cam.Pg[6], p.z, cam.Pg[7])));
                        const bool maybe = okd & !(fabsf(x) > fmaf(hwx, d, tol)) & !(fabsf(y) > fmaf(hwy, d, tol));
                        if (!__any_sync(FULL, maybe)) {
                            if (BITS && k < npts) A.car_bits[seg0 + k] = 0ULL;
                            return;
                        }
                        // ---- stage 3: pixel in fp32, proven by the residual |x - (u - cx) d| < d/2 - tol
                        const float r = rcp_approx(d);
                        const float ru = rintf(fmaf(x, r, cxm)), rv = rintf(fmaf(y, r, cym));
                        const float half = fmaf(0.5f, d, -tol);
                        const bool sure = (fabsf(fmaf(cxm - ru, d, x)) < half) & (fabsf(fmaf(cym - rv, d, y)) < half) &
                                          (ru >= 0.f) & (ru <= umax) & (rv >= 0.f) & (rv <= vmax) &
                                          (d - tol > 0.f) & (d + tol < dlo) & fast;
                        const bool certain = maybe & sure, unsure = maybe & !sure;
                        const unsigned mc = __ballot_sync(FULL, certain), mu = __ballot_sync(FULL, unsure);
                        if (certain) {
                            const int pos = qn + __popc(mc & lt);
                            q[pos] = make_float4(p.x, p.y, p.z,
                                                 __uint_as_float((unsigned)(int)ru | ((unsigned)(int)rv << 16)));
                            if (BITS) qi[pos] = k;
                        }
                        qn += __popc(mc);
                        if (BITS && !maybe && k < npts) A.car_bits[seg0 + k] = 0ULL;
                        if (mu) {
                            if (unsure) uq[un + __popc(mu & lt)] = make_float4(p.x, p.y, p.z, __int_as_float(k));
                            un += __popc(mu);
                            if (un >= 32) {
                                qn = fuse_resolve<BITS>(cam, A, uq + (un - 32), 32, q, qi, qn, seg0);
                                un -= 32;
                            }
                        }
                    };
#pragma unroll
                    for (int j = 0; j < PER; ++j) {
                        const int k = kbase + 32 * j;
                        float d, tol;
                        const bool okd = stage1(P[j], k, d, tol);
                        if (__any_sync(FULL, okd)) stages23(P[j], k, d, tol, okd);
                        else if (BITS && k < npts) A.car_bits[seg0 + k] = 0ULL;
                    }
                } else if (un > 0) {
                    qn = fuse_resolve<BITS>(cam, A, uq, un, q, qi, qn, seg0);
                    un = 0;
                }
                __syncwarp();
                while (qn >= G::GRP || (last && qn > 0)) consume(min(qn, G::GRP));
                if (last) break;
                __syncwarp();
            }
            if (lane == 0 && nval && A.n_valid) atomicAdd(A.n_valid + f, nval);
        }
        id = __shfl_sync(FULL, nxt, 0);
    }
}

}  // namespace
