// Mask resize to camera resolution: cv2.resize(mask.astype(uint8),
// (W, H)) with the default INTER_LINEAR, the step right before the gather in
// extract_car_points_by_mask (cvs_erosion.py:153; V4:293; V5:531).  Bit-exact with
// OpenCV's uint8 path (checked against cv2 4.13 for binary and grey images):
//   * same size            -> copy
//   * exact 2x downscale   -> OpenCV switches INTER_LINEAR to the fast INTER_AREA: (sum of 2x2 + 2) >> 2
//   * otherwise            -> fixed-point bilinear, 11-bit coefficients:
//       fx = (float)((dx + 0.5) * scale_x - 0.5); sx = floor(fx); fx -= sx   (float32)
//       left / right border: fx = 0 and sx clamped; rows are clamped, fy is kept
//       a = { rint((1 - fx) * 2048), rint(fx * 2048) }, b likewise
//       row value r = S[sx] * a0 + S[sx + 1] * a1                           (int32)
//       out = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2
// Every float / double step is spelled with round-to-nearest intrinsics so that nothing is
// contracted into an fma the CPU code does not have.
#include "common.cuh"
#include "erode_body.cuh"

namespace {

struct Tap {
    int s0, s1;     // source indices
    int c0, c1;     // 11-bit coefficients
};

__device__ __forceinline__ Tap tap(int d, double scale, int n, bool zero_at_border) {
    float f = (float)__dadd_rn(__dmul_rn((double)d + 0.5, scale), -0.5);
    int s = (int)floorf(f);
    f = __fsub_rn(f, (float)s);
    Tap t;
    if (zero_at_border) {
        if (s < 0) { f = 0.f; s = 0; }
        if (s >= n - 1) { f = 0.f; s = n - 1; }
        t.s0 = s;
        t.s1 = min(s + 1, n - 1);
    } else {
        t.s0 = min(max(s, 0), n - 1);
        t.s1 = min(max(s + 1, 0), n - 1);
    }
    t.c0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
    t.c1 = __float2int_rn(__fmul_rn(f, 2048.f));
    return t;
}

// Persistent CTAs over (image, output row) pairs.  The column taps depend only on dx, so a
// CTA computes them once into shared memory (8 bytes per output column); after that an
// output pixel costs two shared loads, four cached byte loads and a dozen integer
// instructions.  One thread = 4 consecutive output pixels (one 32-bit store when W % 4 == 0).
__global__ void __launch_bounds__(256)
lf_resize_linear_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst, int H,
                        int W, double scale_x, double scale_y) {
    extern __shared__ int2 s_tap[];                         // per output column: {s0, (s1 - s0) | c0 << 8 | c1 << 20}; 4 * qpr entries
    int2* taps = s_tap;
    const int qpr = (W + 3) >> 2;
    // stored as [dx & 3][dx >> 2]: lane q reads taps[j * qpr + q], consecutive lanes hit consecutive banks
    for (int dx = threadIdx.x; dx < W; dx += blockDim.x) {
        const Tap t = tap(dx, scale_x, w, true);
        taps[(dx & 3) * qpr + (dx >> 2)] = make_int2(t.s0, (int)((unsigned)(t.s1 - t.s0) | ((unsigned)t.c0 << 8) | ((unsigned)t.c1 << 20)));   // s1 - s0 in {0,1}; c <= 2048 (12 bits)
    }
    __syncthreads();
    const int64_t rows = m * (int64_t)H;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int dy = (int)(row % H);
        const int64_t img = row / H;
        const Tap ty = tap(dy, scale_y, h, false);
        const uint8_t* r0 = src + (img * h + ty.s0) * (int64_t)w;
        const uint8_t* r1 = src + (img * h + ty.s1) * (int64_t)w;
        uint8_t* orow = dst + row * (int64_t)W;
        for (int q = threadIdx.x; q < qpr; q += blockDim.x) {
            uint8_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int dx = 4 * q + j;
                if (dx >= W) { o[j] = 0; continue; }
                const int2 t = taps[j * qpr + q];
                const int s0 = t.x, s1 = t.x + (t.y & 1), c0 = ((unsigned)t.y >> 8) & 0xFFF, c1 = (unsigned)t.y >> 20;
                const int a = (int)__ldg(r0 + s0) * c0 + (int)__ldg(r0 + s1) * c1;
                const int b = (int)__ldg(r1 + s0) * c0 + (int)__ldg(r1 + s1) * c1;
                o[j] = (uint8_t)((((ty.c0 * (a >> 4)) >> 16) + ((ty.c1 * (b >> 4)) >> 16) + 2) >> 2);
            }
            if ((W & 3) == 0) *(uchar4*)(orow + 4 * q) = make_uchar4(o[0], o[1], o[2], o[3]);
            else
                for (int j = 0; j < 4 && 4 * q + j < W; ++j) orow[4 * q + j] = o[j];
        }
    }
}

// The same arithmetic with the source rows staged in shared memory: a CTA takes RS_ROWS
// output rows of one image at a time, copies the two source rows each of them needs
// (coalesced 16-byte loads when the row pitch allows), and the taps then read bytes from
// shared memory instead of issuing four byte-wide global loads per pixel (the load/store
// unit, not HBM, bounds the plain kernel).  Used when 2 * RS_ROWS source rows fit.
constexpr int RS_ROWS = 8;

__global__ void __launch_bounds__(256)
lf_resize_linear_smem_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst,
                             int H, int W, double scale_x, double scale_y, int pitch) {
    extern __shared__ __align__(16) uint8_t rs_smem[];
    int2* taps = (int2*)rs_smem;                              // [W]
    uint8_t* rows = rs_smem + (((size_t)4 * ((W + 3) >> 2) * sizeof(int2) + 15) & ~(size_t)15);   // [RS_ROWS][2][pitch]
    __shared__ int s_c0[RS_ROWS], s_c1[RS_ROWS];
    const int qpr = (W + 3) >> 2;
    for (int dx = threadIdx.x; dx < W; dx += blockDim.x) {
        const Tap t = tap(dx, scale_x, w, true);
        taps[(dx & 3) * qpr + (dx >> 2)] =
            make_int2(t.s0, (int)((unsigned)(t.s1 - t.s0) | ((unsigned)t.c0 << 8) | ((unsigned)t.c1 << 20)));
    }
    const int blocks_per_img = (H + RS_ROWS - 1) / RS_ROWS;
    const int64_t items = m * (int64_t)blocks_per_img;
    const bool vec = (w & 15) == 0 && (((uintptr_t)src) & 15) == 0;
    for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
        const int64_t img = it / blocks_per_img;
        const int y0 = (int)(it - img * blocks_per_img) * RS_ROWS;
        const int ny = min(RS_ROWS, H - y0);
        __syncthreads();                                      // taps ready / previous tile consumed
        for (int r = 0; r < 2 * ny; ++r) {
            const Tap ty = tap(y0 + (r >> 1), scale_y, h, false);
            const uint8_t* g = src + (img * h + ((r & 1) ? ty.s1 : ty.s0)) * (int64_t)w;
            uint8_t* srow = rows + (size_t)r * pitch;
            if (vec)
                for (int i = threadIdx.x; i < (w >> 4); i += blockDim.x) ((uint4*)srow)[i] = __ldg((const uint4*)g + i);
            else
                for (int i = threadIdx.x; i < w; i += blockDim.x) srow[i] = __ldg(g + i);
            if (threadIdx.x == 0 && (r & 1) == 0) { s_c0[r >> 1] = ty.c0; s_c1[r >> 1] = ty.c1; }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < ny * qpr; i += blockDim.x) {
            const int ry = i / qpr, q = i - ry * qpr;
            const uint8_t* r0 = rows + (size_t)(2 * ry) * pitch;
            const uint8_t* r1 = r0 + pitch;
            const int b0 = s_c0[ry], b1 = s_c1[ry];
            uint8_t o[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int dx = 4 * q + j;
                if (dx >= W) { o[j] = 0; continue; }
                const int2 t = taps[j * qpr + q];
                const int s0 = t.x, s1 = t.x + (t.y & 1), c0 = ((unsigned)t.y >> 8) & 0xFFF, c1 = (unsigned)t.y >> 20;
                const int a = (int)r0[s0] * c0 + (int)r0[s1] * c1;
                const int b = (int)r1[s0] * c0 + (int)r1[s1] * c1;
                o[j] = (uint8_t)((((b0 * (a >> 4)) >> 16) + ((b1 * (b >> 4)) >> 16) + 2) >> 2);
            }
            uint8_t* orow = dst + (img * H + y0 + ry) * (int64_t)W;
            if ((W & 3) == 0) *(uchar4*)(orow + 4 * q) = make_uchar4(o[0], o[1], o[2], o[3]);
            else
                for (int j = 0; j < 4 && 4 * q + j < W; ++j) orow[4 * q + j] = o[j];
        }
    }
}

// Separable form of the same arithmetic for enlarging (or equal) widths, the reference's case (a 192x704 or
// 384x1408 network mask to the 376x1408 camera image).  The plain kernels above recompute the horizontal
// interpolation for both source rows of EVERY output row (each source row serves ~4 output rows when the
// image is enlarged 2x) and pay four byte-wide loads per pixel.  Here a CTA takes RL_ROWS output rows of one
// image: pass 1 interpolates every source row the tile touches once -- a thread does four neighbouring
// columns from one 8-byte window of the row (two aligned 32-bit loads), each pixel = one PRMT that puts its
// two source bytes into the low half of a register + one DP2A with the packed coefficients (c0 | c1 << 16),
// which IS S0*c0 + S1*c1 -- and leaves the 15-bit values (a >> 4) in shared memory; pass 2 combines two of
// those rows per output pixel, (b*x) >> 16 as the high half of x << 16 times b (IMAD.HI with the running sum
// as addend).  Same integers in the same order as above: bit-exact with cv2.
// Needs w % 4 == 0, W % 4 == 0, 4-byte aligned images, scale_x <= 1 (the window of four columns then spans
// at most 3 + 3 + 2 = 8 bytes) and a vertical scale that keeps the tile's source rows in shared memory.
constexpr int RL_ROWS = 16;

__global__ void __launch_bounds__(256)
lf_resize_linear_sep_kernel(const uint8_t* __restrict__ src, int64_t m, int h, int w, uint8_t* __restrict__ dst, int H,
                            int W, double scale_x, double scale_y, int ns_max) {
    extern __shared__ __align__(16) uint8_t rl_smem[];
    const int G = W >> 2;                                     // groups of four output columns
    uint2* hrow = (uint2*)rl_smem;                            // [ns_max][G] four 16-bit values a >> 4
    uint4* t_coef = (uint4*)(hrow + (size_t)ns_max * G);      // [G] c0 | c1 << 16 of the four columns
    int2* t_base = (int2*)(t_coef + G);                       // [G] {word index of the window, word index of its second half}
    uint32_t* t_sel = (uint32_t*)(t_base + G);                // [G] four PRMT selectors, 8 bits each
    __shared__ int4 s_row[RL_ROWS];                           // per output row {s0 - r_lo, s1 - r_lo, b0, b1}
    __shared__ int s_lo, s_ns;
    for (int g = threadIdx.x; g < G; g += blockDim.x) {
        Tap t[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) t[j] = tap(4 * g + j, scale_x, w, true);
        const int base = t[0].s0 & ~3;
        uint32_t sel = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) sel |= (uint32_t)((t[j].s0 - base) | ((t[j].s1 - base) << 4)) << (8 * j);
        t_sel[g] = sel;
        t_base[g] = make_int2(base >> 2, (base + 4 < w ? base + 4 : base) >> 2);   // never past the row's end
        t_coef[g] = make_uint4((uint32_t)t[0].c0 | ((uint32_t)t[0].c1 << 16), (uint32_t)t[1].c0 | ((uint32_t)t[1].c1 << 16),
                               (uint32_t)t[2].c0 | ((uint32_t)t[2].c1 << 16), (uint32_t)t[3].c0 | ((uint32_t)t[3].c1 << 16));
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    const int tiles_per_img = (H + RL_ROWS - 1) / RL_ROWS;
    const int64_t items = m * (int64_t)tiles_per_img;
    for (int64_t it = blockIdx.x; it < items; it += gridDim.x) {
        const int64_t img = it / tiles_per_img;
        const int y0 = (int)(it - img * tiles_per_img) * RL_ROWS;
        const int ny = min(RL_ROWS, H - y0);
        __syncthreads();                                      // tables ready / previous tile consumed
        if (threadIdx.x < ny) {
            const Tap ty = tap(y0 + threadIdx.x, scale_y, h, false);
            const int lo = tap(y0, scale_y, h, false).s0;     // rows are monotonic: the tile starts at its first row's s0
            s_row[threadIdx.x] = make_int4(ty.s0 - lo, ty.s1 - lo, ty.c0, ty.c1);
            if (threadIdx.x == 0) s_lo = lo;
            if (threadIdx.x == ny - 1) s_ns = ty.s1 - lo + 1;
        }
        __syncthreads();
        const int r_lo = s_lo, ns = s_ns;
        // ---- pass 1: horizontal interpolation of the tile's source rows.  The (row, group) pairs are one flat range
        // cut into rounds of 128 (four groups per lane: all eight window loads go out before the first one is
        // used), dealt to the warps in turn -- a whole row per warp would leave five warps idle for the last rows
        const int total = ns * G;                            // < 2^16 (the host checks): idx / G by one multiplication
        const unsigned gmagic = (unsigned)((0x100000000ULL + (unsigned)G - 1) / (unsigned)G);
        const uint8_t* tile_src = src + (img * h + r_lo) * (int64_t)w;
        for (int c0 = warp * 128 + lane; c0 < total; c0 += nwarps * 128) {
            uint32_t lo[4], hi[4];
            int gk[4], rk[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int idx = c0 + 32 * k;
                lo[k] = hi[k] = 0u;
                rk[k] = mdiv(idx, gmagic);
                gk[k] = idx - rk[k] * G;
                if (idx < total) {
                    const uint32_t* rp = (const uint32_t*)(tile_src + (int64_t)rk[k] * w);
                    const int2 tb = t_base[gk[k]];
                    lo[k] = __ldg(rp + tb.x);
                    hi[k] = __ldg(rp + tb.y);
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (c0 + 32 * k >= total) break;
                const uint32_t sel = t_sel[gk[k]];
                const uint4 cf = t_coef[gk[k]];
                const uint32_t a0 = __dp2a_lo(cf.x, __byte_perm(lo[k], hi[k], sel), 0u);
                const uint32_t a1 = __dp2a_lo(cf.y, __byte_perm(lo[k], hi[k], sel >> 8), 0u);
                const uint32_t a2 = __dp2a_lo(cf.z, __byte_perm(lo[k], hi[k], sel >> 16), 0u);
                const uint32_t a3 = __dp2a_lo(cf.w, __byte_perm(lo[k], hi[k], sel >> 24), 0u);
                hrow[c0 + 32 * k] = make_uint2(__byte_perm(a0 >> 4, a1 << 12, 0x7610), __byte_perm(a2 >> 4, a3 << 12, 0x7610));
            }
        }
        __syncthreads();
        // ---- pass 2: vertical combination, four pixels per thread and store
        for (int y = warp; y < ny; y += nwarps) {
            const int4 rw = s_row[y];
            const uint2* h0 = hrow + (size_t)rw.x * G;
            const uint2* h1 = hrow + (size_t)rw.y * G;
            const uint32_t b0 = (uint32_t)rw.z, b1 = (uint32_t)rw.w;
            uint32_t* orow = (uint32_t*)(dst + (img * H + y0 + y) * (int64_t)W);
#pragma unroll 2
            for (int g = lane; g < G; g += 32) {
                const uint2 x0 = h0[g], x1 = h1[g];
                const uint32_t o0 = __umulhi(x1.x << 16, b1) + (__umulhi(x0.x << 16, b0) + 2u);
                const uint32_t o1 = __umulhi(x1.x & 0xFFFF0000u, b1) + (__umulhi(x0.x & 0xFFFF0000u, b0) + 2u);
                const uint32_t o2 = __umulhi(x1.y << 16, b1) + (__umulhi(x0.y << 16, b0) + 2u);
                const uint32_t o3 = __umulhi(x1.y & 0xFFFF0000u, b1) + (__umulhi(x0.y & 0xFFFF0000u, b0) + 2u);
                orow[g] = (o0 >> 2) | ((o1 >> 2) << 8) | ((o2 >> 2) << 16) | ((o3 >> 2) << 24);
            }
        }
    }
}

// exact 2x reduction, eight output pixels per thread: 16 bytes of each of the two source rows, the byte pairs added
// in 16-bit lanes ((x & 0x00FF00FF) + ((x >> 8) & 0x00FF00FF)), rows added, + 2 >> 2, repacked
__global__ void __launch_bounds__(256)
lf_resize_half_vec_kernel(const uint8_t* __restrict__ src, int64_t m, int H, int W, uint8_t* __restrict__ dst) {
    const int w = 2 * W, qpr = W >> 3;                        // W % 8 == 0
    const int64_t total = m * (int64_t)H * qpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int q = (int)(i % qpr);
        const int64_t row = i / qpr;                          // img * H + dy
        const uint4* p = (const uint4*)(src + row * 2 * (int64_t)w) + q;
        const uint4 a = __ldg(p), b = __ldg(p + (w >> 4));
        const uint32_t av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
        uint32_t o[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t s = (av[k] & 0x00FF00FFu) + ((av[k] >> 8) & 0x00FF00FFu) + (bv[k] & 0x00FF00FFu) +
                               ((bv[k] >> 8) & 0x00FF00FFu) + 0x00020002u;
            o[k] = (s >> 2) & 0x00FF00FFu;                    // two results, in bytes 0 and 2
        }
        ((uint2*)(dst + row * (int64_t)W))[q] = make_uint2(__byte_perm(o[0], o[1], 0x6420), __byte_perm(o[2], o[3], 0x6420));
    }
}

__global__ void __launch_bounds__(256)
lf_resize_half_kernel(const uint8_t* __restrict__ src, int64_t m, int H, int W, uint8_t* __restrict__ dst) {
    const int w = 2 * W;
    const int64_t total = m * (int64_t)H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int dx = (int)(i % W);
        const int64_t row = i / W;                           // img * H + dy
        const uint8_t* p = src + row * 2 * (int64_t)w + 2 * dx;
        dst[i] = (uint8_t)(((int)__ldg(p) + (int)__ldg(p + 1) + (int)__ldg(p + w) + (int)__ldg(p + w + 1) + 2) >> 2);
    }
}


// ---------------------------------------------------------------------------
// Bit-packed masks in, bit-packed "resized != 0" out (what the gather reads after
// cv2.resize(mask.astype(np.uint8), (W, H)) ... > 0.5, cvs_erosion.py:153-154).  For a 0/1 image
// the row value a = S0*c0 + S1*c1 takes one of four values per column, so the column table holds
// (a >> 4) for the three non-zero patterns; an output pixel is set iff
//     ((b0 * At) >> 16) + ((b1 * Ab) >> 16) >= 2        (the "+ 2 >> 2" of the uint8 formula).
// One thread = one output word.  Instance masks are mostly background: when the source span of the
// word is empty in both rows the word is zero without looking at a single tap.
constexpr int RP_ROWS = 64;                               // output rows per work item (one image)

__global__ void __launch_bounds__(256)
lf_resize_pack_kernel(const uint32_t* __restrict__ src, int64_t m, int h, int w, int wpr_s, uint32_t* __restrict__ dst,
                      int32_t* __restrict__ bbox, int H, int W, int wpr_d, double scale_x, double scale_y) {
    extern __shared__ int2 rp_taps[];                       // per output column {s0, d | A(s0 only) << 1 | A(s1 only) << 9 | A(both) << 17}
    __shared__ int4 s_row[RP_ROWS];                         // per output row of the item {s0, s1, b0, b1}
    for (int dx = threadIdx.x; dx < W; dx += blockDim.x) {
        const Tap t = tap(dx, scale_x, w, true);
        rp_taps[dx] = make_int2(t.s0, (int)((unsigned)(t.s1 - t.s0) | ((unsigned)(t.c0 >> 4) << 1) |
                                            ((unsigned)(t.c1 >> 4) << 9) | ((unsigned)((t.c0 + t.c1) >> 4) << 17)));
    }
    const unsigned blocks_per_img = (unsigned)((H + RP_ROWS - 1) / RP_ROWS);
    const unsigned items = (unsigned)m * blocks_per_img;    // < 2^31 (lf_resize_pack)
    for (unsigned it = blockIdx.x; it < items; it += gridDim.x) {
        const unsigned img32 = it / blocks_per_img;
        const int64_t img = img32;
        const int y0 = (int)(it - img32 * blocks_per_img) * RP_ROWS;
        const int ny = min(RP_ROWS, H - y0);
        __syncthreads();                                    // taps ready / previous item's rows consumed
        if (threadIdx.x < ny) {
            const Tap ty = tap(y0 + threadIdx.x, scale_y, h, false);
            s_row[threadIdx.x] = make_int4(ty.s0, ty.s1, ty.c0, ty.c1);
        }
        __syncthreads();
        BBoxAcc acc;
        const uint32_t* ibase = src + img * h * (int64_t)wpr_s;
        uint32_t* obase = dst + (img * H + y0) * (int64_t)wpr_d;
        const int lane = threadIdx.x & 31;
        // thread = (word column, row group): everything that depends on the column only -- the span of
        // source columns, its first / last word and their masks -- is computed once per item, the row
        // loop then costs the loads, two ANDs / ORs and the store.  Loops are warp-uniform (ballots).
        const int cols = min(wpr_d, (int)blockDim.x), rgroups = (int)blockDim.x / cols;
        const int my_col = (int)threadIdx.x % cols, my_rg = (int)threadIdx.x / cols;
        for (int wx0 = 0; wx0 < wpr_d; wx0 += cols) {
            const int wx = wx0 + my_col;
            const bool col_ok = wx < wpr_d && my_rg < rgroups;
            int dx0 = 0, npx = 0, sa = 0, sb = 0, wa = 0, wb = 0;
            uint32_t keep_a = 0, keep_b = 0;
            if (col_ok) {
                dx0 = 32 * wx;
                npx = min(32, W - dx0);
                const int2 tf = rp_taps[dx0], tl = rp_taps[dx0 + npx - 1];
                sa = tf.x; sb = tl.x + (tl.y & 1);            // source columns the word can touch (taps are monotonic)
                wa = sa >> 5; wb = sb >> 5;
                keep_a = 0xFFFFFFFFu << (sa & 31);
                keep_b = 0xFFFFFFFFu >> (31 - (sb & 31));
                if (wa == wb) { keep_a &= keep_b; keep_b = 0u; }
            }
            const uint32_t ones = npx == 32 ? 0xFFFFFFFFu : (1u << npx) - 1u;
            for (int rb = 0; rb < ny; rb += rgroups) {
                const int ry = rb + my_rg;
                const bool active = col_ok && ry < ny;
                int off0 = 0, off1 = 0, b0 = 0, b1 = 0;
                uint32_t out = 0;
                bool mixed = false;
                if (active) {
                    const int4 rw = s_row[ry];
                    off0 = rw.x * wpr_s; off1 = rw.y * wpr_s; b0 = rw.z; b1 = rw.w;
                    uint32_t any, all;
                    {
                        const uint32_t v0 = __ldg(ibase + off0 + wa), v1 = __ldg(ibase + off1 + wa);
                        any = (v0 | v1) & keep_a;
                        all = (v0 & v1) | ~keep_a;
                    }
                    if (wb > wa) {
                        const uint32_t v0 = __ldg(ibase + off0 + wb), v1 = __ldg(ibase + off1 + wb);
                        any |= (v0 | v1) & keep_b;
                        all &= (v0 & v1) | ~keep_b;
                        for (int q = wa + 1; q < wb; ++q) {   // shrinking by more than 2: whole words in between
                            const uint32_t u0 = __ldg(ibase + off0 + q), u1 = __ldg(ibase + off1 + q);
                            any |= u0 | u1;
                            all &= u0 & u1;
                        }
                    }
                    // nothing set in the span: 0.  Everything set: every pixel has At = Ab = (c0 + c1) >> 4 >= 127
                    // and b0 + b1 >= 2047, so floor(b0 * At / 65536) + floor(b1 * Ab / 65536) >= 3 - 1 = 2: all ones.
                    if (any == 0u) out = 0u;
                    else if (all == 0xFFFFFFFFu) out = ones;
                    else mixed = true;
                }
                // words with an edge inside: the warp does them together, one pixel per lane
                unsigned mm = __ballot_sync(0xFFFFFFFFu, mixed);
                while (mm) {
                    const int k = __ffs((int)mm) - 1;
                    mm &= mm - 1;
                    const int k0 = __shfl_sync(0xFFFFFFFFu, off0, k), k1 = __shfl_sync(0xFFFFFFFFu, off1, k);
                    const int kb0 = __shfl_sync(0xFFFFFFFFu, b0, k), kb1 = __shfl_sync(0xFFFFFFFFu, b1, k);
                    const int kdx = __shfl_sync(0xFFFFFFFFu, dx0, k), kn = __shfl_sync(0xFFFFFFFFu, npx, k);
                    bool bit = false;
                    if (lane < kn) {
                        const int2 t = rp_taps[kdx + lane];
                        const int s0 = t.x, s1 = t.x + (t.y & 1);
                        const unsigned pt = ((__ldg(ibase + k0 + (s0 >> 5)) >> (s0 & 31)) & 1u) |
                                            (((__ldg(ibase + k0 + (s1 >> 5)) >> (s1 & 31)) & 1u) << 1);
                        const unsigned pb = ((__ldg(ibase + k1 + (s0 >> 5)) >> (s0 & 31)) & 1u) |
                                            (((__ldg(ibase + k1 + (s1 >> 5)) >> (s1 & 31)) & 1u) << 1);
                        const int At = pt ? ((unsigned)t.y >> (8 * pt - 7)) & 0xFF : 0;
                        const int Ab = pb ? ((unsigned)t.y >> (8 * pb - 7)) & 0xFF : 0;
                        bit = (((kb0 * At) >> 16) + ((kb1 * Ab) >> 16)) >= 2;
                    }
                    const unsigned w = __ballot_sync(0xFFFFFFFFu, bit);
                    if (lane == k) out = w;
                }
                if (active) {
                    obase[ry * wpr_d + wx] = out;
                    acc.add(out, y0 + ry, wx);
                }
            }
        }
        if (bbox) acc.flush(bbox + 4 * img);
    }
}

// keep the even bits of x (bit 2j -> bit j)
__device__ __forceinline__ uint32_t even_bits(unsigned long long x) {
    x &= 0x5555555555555555ULL;
    x = (x | (x >> 1)) & 0x3333333333333333ULL;
    x = (x | (x >> 2)) & 0x0F0F0F0F0F0F0F0FULL;
    x = (x | (x >> 4)) & 0x00FF00FF00FF00FFULL;
    x = (x | (x >> 8)) & 0x0000FFFF0000FFFFULL;
    x = (x | (x >> 16)) & 0x00000000FFFFFFFFULL;
    return (uint32_t)x;
}

// exact 2x reduction: OpenCV's INTER_AREA substitute, (sum of 2x2 + 2) >> 2 != 0 <=> at least two of the four set
__global__ void __launch_bounds__(256)
lf_resize_pack_half_kernel(const uint32_t* __restrict__ src, int64_t m, int wpr_s, uint32_t* __restrict__ dst,
                           int32_t* __restrict__ bbox, int H, int W, int wpr_d) {
    const int64_t rows = m * (int64_t)H;
    for (int64_t row = blockIdx.x; row < rows; row += gridDim.x) {
        const int64_t img = row / H;
        const int dy = (int)(row - img * H);
        const uint32_t* r0 = src + (img * 2 * H + 2 * dy) * (int64_t)wpr_s;
        const uint32_t* r1 = r0 + wpr_s;
        BBoxAcc acc;
        for (int wx = threadIdx.x; wx < wpr_d; wx += blockDim.x) {
            const int q = 2 * wx;
            const unsigned long long a = (unsigned long long)__ldg(r0 + q) | ((q + 1 < wpr_s ? (unsigned long long)__ldg(r0 + q + 1) : 0ULL) << 32);
            const unsigned long long b = (unsigned long long)__ldg(r1 + q) | ((q + 1 < wpr_s ? (unsigned long long)__ldg(r1 + q + 1) : 0ULL) << 32);
            const unsigned long long a1 = a >> 1, b1 = b >> 1;
            uint32_t out = even_bits((a & a1) | (b & b1) | ((a | a1) & (b | b1)));
            if (32 * wx + 32 > W) out &= (1u << (W & 31)) - 1u;
            dst[row * wpr_d + wx] = out;
            acc.add(out, dy, wx);
        }
        if (bbox) acc.flush(bbox + 4 * img);
    }
}

}  // namespace

extern "C" int lf_resize_masks(lf_ctx* ctx, const uint8_t* src, int64_t m, int h, int w, uint8_t* dst, int H, int W,
                               void* stream) {
    if (!ctx || m < 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return LF_E_ARG;
    if (m == 0) return LF_OK;
    if (!src || !dst) return LF_E_ARG;
    if ((int64_t)h * w >= ((int64_t)1 << 31) || (int64_t)H * W >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    if (h == H && w == W) {
        LF_CUDA(ctx, cudaMemcpyAsync(dst, src, (size_t)m * h * w, cudaMemcpyDeviceToDevice, st));
        return LF_OK;
    }
    const int64_t cap = (int64_t)ctx->sm_count * 16;
    if (w == 2 * W && h == 2 * H) {
        if ((W & 7) == 0 && (((uintptr_t)src | (uintptr_t)dst) & 15) == 0) {
            int64_t blocks = (m * H * (W >> 3) + 255) / 256;
            if (blocks > cap) blocks = cap;
            lf_resize_half_vec_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, m, H, W, dst);
        } else {
            int64_t blocks = (m * H * W + 255) / 256;
            if (blocks > cap) blocks = cap;
            lf_resize_half_kernel<<<(unsigned)blocks, 256, 0, st>>>(src, m, H, W, dst);
        }
    } else {
        // OpenCV: inv_scale = (double)dsize / ssize; scale = 1. / inv_scale
        const double scale_x = 1.0 / ((double)W / (double)w), scale_y = 1.0 / ((double)H / (double)h);
        // enlarging widths (the reference's case): the separable kernel, when its rows fit shared memory
        const int ns_max = (int)((RL_ROWS - 1) * scale_y) + 4;
        const size_t smem_sep = (size_t)ns_max * (W >> 2) * sizeof(uint2) + (size_t)(W >> 2) * (16 + 8 + 4);
        if ((w & 3) == 0 && (W & 3) == 0 && w <= W && (((uintptr_t)src | (uintptr_t)dst) & 3) == 0 &&
            smem_sep <= 100 * 1024 && (int64_t)ns_max * (W >> 2) + 128 < 65536) {
            int64_t b3 = m * ((H + RL_ROWS - 1) / RL_ROWS);
            LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_linear_sep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem_sep));
            int per_sm = 0;
            LF_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lf_resize_linear_sep_kernel, 256, smem_sep));
            const int64_t c3 = (int64_t)ctx->sm_count * (per_sm > 0 ? per_sm : 1);
            if (b3 > c3) b3 = c3;
            lf_resize_linear_sep_kernel<<<(unsigned)b3, 256, smem_sep, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y, ns_max);
            LF_CUDA(ctx, cudaGetLastError());
            return LF_OK;
        }
        if ((size_t)W * sizeof(int2) > 96 * 1024) return LF_E_UNSUPPORTED;      // column taps live in shared memory
        int64_t blocks = m * H;
        const int64_t pcap = (int64_t)ctx->sm_count * 8;
        if (blocks > pcap) blocks = pcap;
        const size_t smem = (size_t)4 * ((W + 3) / 4) * sizeof(int2);
        const int pitch = (w + 15) & ~15;
        const size_t smem2 = ((smem + 15) & ~(size_t)15) + (size_t)2 * RS_ROWS * pitch;
        if (smem2 <= 64 * 1024) {
            int64_t b2 = m * ((H + RS_ROWS - 1) / RS_ROWS);
            const int64_t c2 = (int64_t)ctx->sm_count * 6;
            if (b2 > c2) b2 = c2;
            LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_linear_smem_kernel,
                                              cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2));
            lf_resize_linear_smem_kernel<<<(unsigned)b2, 256, smem2, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y,
                                                                           pitch);
        } else {
            LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_linear_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              (int)smem));
            lf_resize_linear_kernel<<<(unsigned)blocks, 256, smem, st>>>(src, m, h, w, dst, H, W, scale_x, scale_y);
        }
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}

extern "C" int lf_resize_pack(lf_ctx* ctx, const uint32_t* src, int64_t m, int h, int w, uint32_t* dst, int32_t* bbox,
                              int H, int W, void* stream) {
    if (!ctx || m < 0 || h <= 0 || w <= 0 || H <= 0 || W <= 0) return LF_E_ARG;
    if (m == 0) return LF_OK;
    if (!src || !dst || (const void*)src == (const void*)dst) return LF_E_ARG;
    if ((int64_t)h * w >= ((int64_t)1 << 31) || (int64_t)H * W >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    const int wpr_s = (w + 31) / 32, wpr_d = (W + 31) / 32;
    if (w == 2 * W && h == 2 * H) {
        int64_t blocks = m * H;
        const int64_t cap = (int64_t)ctx->sm_count * 16;
        if (blocks > cap) blocks = cap;
        lf_resize_pack_half_kernel<<<(unsigned)blocks, wpr_d >= 128 ? 256 : 64, 0, st>>>(src, m, wpr_s, dst, bbox, H, W, wpr_d);
    } else {
        const double scale_x = 1.0 / ((double)W / (double)w), scale_y = 1.0 / ((double)H / (double)h);
        const size_t smem = (size_t)W * sizeof(int2);
        if (smem > 96 * 1024 || (int64_t)RP_ROWS * wpr_d >= 65536) return LF_E_UNSUPPORTED;
        if (m * ((H + RP_ROWS - 1) / RP_ROWS) >= ((int64_t)1 << 31)) return LF_E_UNSUPPORTED;
        int64_t blocks = m * ((H + RP_ROWS - 1) / RP_ROWS);
        const int64_t cap = (int64_t)ctx->sm_count * 8;
        if (blocks > cap) blocks = cap;
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_resize_pack_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        lf_resize_pack_kernel<<<(unsigned)blocks, 256, smem, st>>>(src, m, h, w, wpr_s, dst, bbox, H, W, wpr_d, scale_x,
                                                                  scale_y);
    }
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
