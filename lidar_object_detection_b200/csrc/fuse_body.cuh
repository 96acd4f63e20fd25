// K1+K4+K5 fused point pass (cvs_erosion.py:342-372 without the erosion): device code of
// lf_fuse_kernel (fuse.cu).
//
// Register streaming, no shared-memory staging of the points and no CTA-level hand-off:
//   phase A  every thread keeps PER independent 128-bit streaming loads in flight (a warp
//            owns 32-point groups dealt round-robin over the CTA's warps), then classifies
//            each group in fp32 in three stages of rising cost, each ended by a warp vote
//            so that groups without a survivor stop early:
//              1. depth row only                    -> behind the camera / beyond the range gate
//              2. centred image rows                -> more than a pixel outside the image
//              3. pixel = rint(x/d) in fp32 + a residual test that PROVES it equal to the
//                 reference's float64 result        -> "certain" (pixel known) or "uncertain"
//            Survivors are compacted (ballot + popc prefix) into two per-warp shared-memory
//            queues: certain points with their pixel, uncertain ones for the exact chain.
//   exact    whenever 32 uncertain points are queued the warp runs the float64 dgemm-order
//            chain and the two divisions on all 32 lanes (common.cuh) and moves the points
//            that pass the gate of cvs_erosion.py:359 to the certain queue.
//   phase B  full 32-lane groups of certain points: packed-mask gather (L2), box cull,
//            float64 containment, ballot/popc counters in shared memory.
// Nothing fp32 ever decides a result: a point is "certain" only if the fp32 residual proves
// the rounded pixel and the gate (error bounds below); everything else takes the exact path.
#pragma once
#include "common.cuh"

namespace {

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
    float dmax_hi;       // smallest float >= max_depth (stage 1: certainly beyond the gate)
    float dmax_lo;       // largest float < max_depth    (stage 3: certainly inside the gate)
    int use_oriented;
    int force_exact;     // verification mode: every surviving point takes the float64 chain
    int smem_boxes;      // boxes staged in shared memory (max over frames)
    int smem_cars;       // max masks per frame (<= LF_MAX_CARS)
    int seg_points;      // points per CTA: a multiple of 32 * WARPS
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
__device__ __forceinline__ float rcp_approx(float v) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
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

// Per-CTA shared state of one (frame, segment) work item.
struct FuseShared {
    int4* mbb;            // [smem_cars] mask bounding boxes {x0, y0, x1, y1}
    lf_box* box;          // [smem_boxes]
    float* cull;          // [6][smem_boxes] lo.x lo.y lo.z hi.x hi.y hi.z, one box per lane without bank conflicts
    int32_t* inside;      // [smem_cars][smem_boxes]
    int32_t* nin;         // [smem_boxes]
    int32_t* total;       // [smem_cars]
    int32_t* nvalid;      // [1]
};

// ---- phase B: 32 certain points, one per lane (`have` false only in a warp's last, partial group).
// e = {x, y, z, bits(u | v << 16)}.  Returns the lane's membership bits.
__device__ __forceinline__ unsigned long long fuse_group(const LfCam& cam, const FuseArgs& A, const FuseShared& S,
                                                         float4 e, bool have, int lane, int M, int B,
                                                         const uint32_t* mask_base, int plane32) {
    const unsigned FULL = 0xFFFFFFFFu;
    const unsigned uv = __float_as_uint(e.w);
    const int ui = (int)(uv & 0xFFFFu), vi = (int)(uv >> 16);
    // ---- K4: point-in-mask gather (cvs_erosion.py:152-160).  Lanes first test the masks' bounding
    // boxes against the pixel bounds of the group (one mask per lane); only surviving masks are read.
    unsigned long long bits = 0ULL;
    if (M > 0) {
        const int off = vi * cam.wpr + (ui >> 5);            // word offset inside one packed mask
        const int sh = ui & 31;
        const int u0 = __reduce_min_sync(FULL, have ? ui : 0x7fffffff);
        const int u1 = __reduce_max_sync(FULL, have ? ui : -1);
        const int v0 = __reduce_min_sync(FULL, have ? vi : 0x7fffffff);
        const int v1 = __reduce_max_sync(FULL, have ? vi : -1);
        for (int g = 0; g < M; g += 32) {
            const int cl = g + lane;
            bool cand = false;
            if (cl < M) {
                const int4 bb = S.mbb[cl];
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
                    w[j] = (have && c[j] >= 0) ? __ldg(mask_base + (unsigned)(off + c[j] * plane32)) : 0u;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c[j] >= 0) bits |= (unsigned long long)((w[j] >> sh) & 1u) << c[j];
            }
        }
        unsigned long long un = warp_or64(bits);
        while (un) {
            const int c = __ffsll((long long)un) - 1;
            un &= un - 1;
            const unsigned bal = __ballot_sync(FULL, (bits >> c) & 1ULL);
            if (lane == 0) atomicAdd(&S.total[c], __popc(bal));
        }
    }
    if (B == 0) return bits;
    // ---- bounding box of the group's points, then lanes test the boxes' cull bounds in parallel
    const float INF = __int_as_float(0x7f800000);
    const float wx0 = warp_min_f32(have ? e.x : INF), wx1 = warp_max_f32(have ? e.x : -INF);
    const float wy0 = warp_min_f32(have ? e.y : INF), wy1 = warp_max_f32(have ? e.y : -INF);
    const float wz0 = warp_min_f32(have ? e.z : INF), wz1 = warp_max_f32(have ? e.z : -INF);
    const int SB = A.smem_boxes;
    for (int g = 0; g < B; g += 32) {
        const int bl = g + lane;
        bool cand = false;
        if (bl < B) {
            const float* cu = S.cull + bl;
            cand = wx1 >= cu[0] && wx0 <= cu[3 * SB] && wy1 >= cu[SB] && wy0 <= cu[4 * SB] && wz1 >= cu[2 * SB] &&
                   wz0 <= cu[5 * SB];
        }
        unsigned cmask = __ballot_sync(FULL, cand);
        // ---- K5: containment (cvs_erosion.py:114-145 / V1:118-139) on the candidates
        while (cmask) {
            const int b = g + __ffs((int)cmask) - 1;
            cmask &= cmask - 1;
            const lf_box& bx = S.box[b];
            const bool near = have && e.x >= bx.lo[0] && e.x <= bx.hi[0] && e.y >= bx.lo[1] && e.y <= bx.hi[1] &&
                              e.z >= bx.lo[2] && e.z <= bx.hi[2];
            const bool in = near && point_in_box(bx, e.x, e.y, e.z, A.use_oriented);
            const unsigned bal = __ballot_sync(FULL, in);
            if (!bal) continue;
            if (lane == 0) atomicAdd(&S.nin[b], __popc(bal));
            unsigned long long un2 = warp_or64(in ? bits : 0ULL);
            while (un2) {
                const int c = __ffsll((long long)un2) - 1;
                un2 &= un2 - 1;
                const unsigned bb = __ballot_sync(FULL, in && ((bits >> c) & 1ULL));
                if (lane == 0) atomicAdd(&S.inside[c * SB + b], __popc(bb));
            }
        }
    }
    return bits;
}

// ---- exact path: up to 32 queued uncertain points (one per lane) through the float64 dgemm-order
// chain and the two divisions (common.cuh); points that pass the gate of cvs_erosion.py:359 are
// appended to the certain queue.  Rare (a few per cent of the survivors), hence not inlined into
// the eight unrolled classification stages.  Returns the new fill of the certain queue.
template <bool BITS>
__device__ __noinline__ int fuse_resolve(const LfCam& cam, const FuseArgs& A, const float4* uq, int cnt, float4* q,
                                         int* qi, int qn, int64_t seg0, int npts) {
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
    } else if (BITS && have && __float_as_int(e.w) < npts) {
        A.car_bits[seg0 + __float_as_int(e.w)] = 0ULL;
    }
    __syncwarp();
    return qn + __popc(mk);
}

// One CTA = (frame f, segment `seg` of A.seg_points points).
//   WARPS  warps per CTA, all of them stream and consume
//   PER    128-bit loads in flight per thread and iteration (an iteration covers WARPS*PER*32 points)
//   BITS   per-point membership output (car_bits) wanted: queues carry the point index too
template <int WARPS, int PER, bool BITS>
struct FuseGeom {
    static constexpr int NT = 32 * WARPS;
    static constexpr int SPAN = WARPS * PER * 32;          // points per iteration
    static constexpr int QCAP = PER * 32 + 64;             // certain queue: < 32 left over + PER*32 new + < 32 resolved
    static constexpr int UCAP = 64;                        // uncertain queue: < 32 left over + 32 new
    static constexpr size_t warp_bytes = (size_t)(QCAP + UCAP) * 16 + (BITS ? (size_t)QCAP * 4 : 0);
    static constexpr size_t queue_bytes = warp_bytes * WARPS;
};

__host__ __device__ inline size_t fuse_state_bytes(int cars, int boxes) {
    // mbb + lf_box + cull + counters (inside, nin, total, nvalid + pad)
    return 16 * (size_t)cars + (size_t)boxes * sizeof(lf_box) + 24 * (size_t)boxes +
           4 * ((size_t)cars * boxes + boxes + cars + 4);
}

template <int WARPS, int PER, bool BITS>
__device__ __forceinline__ void fuse_segment(const LfCam& cam, const FuseArgs& A, int f, int seg, uint8_t* smem) {
    using G = FuseGeom<WARPS, PER, BITS>;
    constexpr int NT = G::NT, SPAN = G::SPAN, QCAP = G::QCAP, UCAP = G::UCAP;
    const unsigned FULL = 0xFFFFFFFFu;
    const int64_t p0 = A.pt_off[f], p1 = A.pt_off[f + 1];
    const int64_t seg0 = p0 + (int64_t)seg * A.seg_points;
    if (seg0 >= p1) return;
    const int npts = (int)min((int64_t)A.seg_points, p1 - seg0);
    const int m0 = A.mask_off[f];
    const int M = A.mask_off[f + 1] - m0;
    const int b0 = A.box_off[f];
    const int B = A.box_off[f + 1] - b0;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // ---- shared memory: per-warp queues, then the frame state
    uint8_t* wq = smem + (size_t)warp * G::warp_bytes;
    float4* q = (float4*)wq;                               // [QCAP] certain: x, y, z, bits(u | v << 16)
    float4* uq = q + QCAP;                                 // [UCAP] uncertain: x, y, z, bits(index in segment)
    int* qi = (int*)(uq + UCAP);                           // [QCAP] index in segment of the certain entries (BITS)
    FuseShared S;
    S.mbb = (int4*)(smem + G::queue_bytes);
    S.box = (lf_box*)(S.mbb + A.smem_cars);
    S.cull = (float*)((uint8_t*)S.box + (size_t)A.smem_boxes * sizeof(lf_box));
    S.inside = (int32_t*)(S.cull + 6 * (size_t)A.smem_boxes);
    S.nin = S.inside + (size_t)A.smem_cars * A.smem_boxes;
    S.total = S.nin + A.smem_boxes;
    S.nvalid = S.total + A.smem_cars;
    {   // stage the frame's boxes and mask bounding boxes, clear the private counters
        const double* gsrc = (const double*)(A.boxes + b0);
        double* sdst = (double*)S.box;
        const int nd = B * (int)(sizeof(lf_box) / 8);
        for (int i = threadIdx.x; i < nd; i += NT) sdst[i] = gsrc[i];
        for (int i = threadIdx.x; i < 6 * B; i += NT) {
            const int k = i / B, b = i - k * B;
            const lf_box* gb = A.boxes + b0 + b;
            S.cull[k * A.smem_boxes + b] = k < 3 ? gb->lo[k] : gb->hi[k - 3];
        }
        const int nc = A.smem_cars * A.smem_boxes + A.smem_boxes + A.smem_cars + 1;
        for (int i = threadIdx.x; i < nc; i += NT) S.inside[i] = 0;
        for (int c = threadIdx.x; c < M; c += NT) {
            int4 bb = make_int4(0, 0, cam.W - 1, cam.H - 1);
            if (A.mask_bbox) {                             // record of lf_erode_pack: zeros = empty mask
                const int4 r = __ldg((const int4*)A.mask_bbox + m0 + c);
                bb = make_int4((1 << 20) - r.x, (1 << 20) - r.y, r.z - 1, r.w - 1);
            }
            S.mbb[c] = bb;
        }
    }
    __syncthreads();

    const int plane32 = cam.H * cam.wpr;                   // words per packed mask (< 2^25, see lf_set_camera)
    const uint32_t* mask_base = A.packed + (int64_t)m0 * plane32;
    const float4* src = A.pts + seg0;
    const unsigned lt = (1u << lane) - 1u;
    // fp32 classification constants (lf_set_camera; error analysis there and in DESIGN.md section 4)
    const float kT = cam.kT, kTs = cam.kTs, gS = cam.gS, hwx = cam.hwx, hwy = cam.hwy;
    const float cxm = cam.cxm, cym = cam.cym, umax = (float)(cam.W - 1), vmax = (float)(cam.H - 1);
    const float dhi = A.dmax_hi, dlo = A.dmax_lo;
    const bool fast = !A.force_exact;
    int qn = 0, un = 0;                                    // queue fill, warp-uniform
    int nval = 0;                                          // certain points consumed by this warp

    // the exact chain on up to 32 queued uncertain points; survivors join the certain queue
    auto resolve = [&](int cnt) {
        qn = fuse_resolve<BITS>(cam, A, uq + (un - cnt), cnt, q, qi, qn, seg0, npts);
        un -= cnt;
    };
    // phase B on the last 32 (or fewer, at the very end) entries of the certain queue
    auto consume = [&](int cnt) {
        const bool have = lane < cnt;
        const int at = qn - cnt + (have ? lane : 0);
        const float4 e = q[at];
        const unsigned long long bits = fuse_group(cam, A, S, e, have, lane, M, B, mask_base, plane32);
        if (BITS && have) A.car_bits[seg0 + qi[at]] = bits;
        nval += cnt;
        qn -= cnt;
    };

    const int niter = (npts + SPAN - 1) / SPAN;
    for (int it = 0;; ++it) {
        const bool last = it == niter;                      // pseudo-iteration: flush both queues
        if (!last) {
            const int kbase = it * SPAN + warp * 32 + lane; // group j of this warp: kbase + j * NT
            float4 P[PER];
            if (it * SPAN + SPAN <= npts) {
#pragma unroll
                for (int j = 0; j < PER; ++j) P[j] = ld_stream_f4(src + kbase + j * NT);
            } else {
                const float qnan = __int_as_float(0x7fc00000);   // dead lanes: NaN takes the exact path and fails there
#pragma unroll
                for (int j = 0; j < PER; ++j)
                    P[j] = kbase + j * NT < npts ? ld_stream_f4(src + kbase + j * NT) : make_float4(qnan, qnan, qnan, 0.f);
            }
#pragma unroll
            for (int j = 0; j < PER; ++j) {
                const float4 p = P[j];
                const int k = kbase + j * NT;
                // ---- stage 1: depth row.  tol bounds every fp32 evaluation error of this point
                // (+inf beyond 2^20 m and for NaN / inf input: such points are never decided here).
                const float d = fmaf(cam.Pg[8], p.x, fmaf(cam.Pg[9], p.y, fmaf(cam.Pg[10], p.z, cam.Pg[11])));
                const float tol = fmaf((fabsf(p.x) + fabsf(p.y) + fabsf(p.z)) * gS, kTs, kT);
                const bool okd = !(d < -tol) & !(d > dhi + tol);     // NaN stays in
                if (!__any_sync(FULL, okd)) {
                    if (BITS && k < npts) A.car_bits[seg0 + k] = 0ULL;
                    continue;
                }
                // ---- stage 2: centred image rows, one pixel of margin
                const float x = fmaf(cam.Pg[0], p.x, fmaf(cam.Pg[1], p.y, fmaf(cam.Pg[2], p.z, cam.Pg[3])));
                const float y = fmaf(cam.Pg[4], p.x, fmaf(cam.Pg[5], p.y, fmaf(cam.Pg[6], p.z, cam.Pg[7])));
                const bool maybe = okd & !(fabsf(x) > fmaf(hwx, d, tol)) & !(fabsf(y) > fmaf(hwy, d, tol));
                if (!__any_sync(FULL, maybe)) {
                    if (BITS && k < npts) A.car_bits[seg0 + k] = 0ULL;
                    continue;
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
                    q[pos] = make_float4(p.x, p.y, p.z, __uint_as_float((unsigned)(int)ru | ((unsigned)(int)rv << 16)));
                    if (BITS) qi[pos] = k;
                }
                qn += __popc(mc);
                if (BITS && !maybe && k < npts) A.car_bits[seg0 + k] = 0ULL;
                if (mu) {
                    if (unsure) uq[un + __popc(mu & lt)] = make_float4(p.x, p.y, p.z, __int_as_float(k));
                    un += __popc(mu);
                    if (un >= 32) resolve(32);
                }
            }
        } else if (un > 0) {
            resolve(un);
        }
        __syncwarp();
        while (qn >= 32 || (last && qn > 0)) consume(min(qn, 32));
        if (last) break;
        __syncwarp();
    }
    if (lane == 0 && nval) atomicAdd(S.nvalid, nval);
    __syncthreads();
    for (int i = threadIdx.x; i < M * B; i += NT) {
        const int c = i / B, b = i - c * B;
        const int v = S.inside[c * A.smem_boxes + b];
        if (v) atomicAdd(A.inside + A.cnt_off[f] + i, v);
    }
    for (int c = threadIdx.x; c < M; c += NT) {
        const int v = S.total[c];
        if (v) atomicAdd(A.total + m0 + c, v);
    }
    for (int b = threadIdx.x; b < B; b += NT) {
        const int v = S.nin[b];
        if (v) atomicAdd(A.n_in_box + b0 + b, v);
    }
    if (threadIdx.x == 0 && *S.nvalid) atomicAdd(A.n_valid + f, *S.nvalid);
}

}  // namespace
