// K3: mask binarisation + binary erosion + bit-packing  (cvs_erosion.py:96-106).
//
// One CTA = (mask, strip of rows).  The strip plus its halo (iterations * k/2
// rows either side) is streamed from HBM exactly once, binarised and packed to
// one bit per pixel in shared memory (176 B per 1408-px row), eroded there with
// word-wide shifts and ANDs, and only the packed result goes back out:
// 1 byte in, 1/8 byte out per pixel.  Algorithmic bytes per mask: H*W in.
//
// Fast path (uint8, W % 32 == 0, 16-B aligned): the strip is one contiguous byte
// run, pulled in by cp.async.bulk (TMA engine, SASS UBLKCP) through a 4-stage
// mbarrier ring, so ~24-32 KB per CTA are in flight with no register cost.
// Generic path (any W, uint8 or float32): coalesced loads + __ballot_sync.
#include "common.cuh"
#include "bulk.cuh"

#include "erode_body.cuh"

namespace {

// One erosion pass, smem -> smem, over the loaded rows [ga, gb) of the image.
// Pixels outside the image are ones (cv2 border value +inf for erosion); rows
// inside the image but outside the loaded window are treated as ones too --
// the garbage this creates stays inside the halo.
__device__ __forceinline__ void erode_pass(const ErodeParams& P, const uint32_t* __restrict__ src,
                                           uint32_t* __restrict__ dst, int ga, int gb) {
    const int nrows = gb - ga;
    const int items = nrows * P.wpr;
    const uint32_t padmask = (P.W & 31) ? ~((1u << (P.W & 31)) - 1u) : 0u;
    for (int it = threadIdx.x; it < items; it += ER_THREADS) {
        const int lr = it / P.wpr, wx = it - lr * P.wpr;
        uint32_t acc = 0xFFFFFFFFu;
        for (int i = 0; i < P.k; ++i) {
            const unsigned bits = P.se_rows[i];
            if (!bits) continue;
            const int lr2 = lr + i - P.r;
            if (lr2 < 0 || lr2 >= nrows) continue;
            const uint32_t* row = src + lr2 * P.wpr;
            const uint32_t c = row[wx];
            const uint32_t l = wx > 0 ? row[wx - 1] : 0xFFFFFFFFu;
            const uint32_t rr = wx + 1 < P.wpr ? row[wx + 1] : 0xFFFFFFFFu;
            for (int j = 0; j < P.k; ++j) {
                if (!((bits >> j) & 1u)) continue;
                const int dx = j - P.r;
                uint32_t s;
                if (dx == 0) s = c;
                else if (dx > 0) s = __funnelshift_r(c, rr, dx);
                else s = __funnelshift_l(l, c, -dx);
                acc &= s;
            }
        }
        if (wx == P.wpr - 1) acc |= padmask;
        dst[it] = acc;
    }
}

// AND over dx = 0..b of bit (x + dx), for the 32 pixels of word c (r = next word): window doubling
// on the 64-bit pair.  Zeros shifted in at the top never reach the low word for b <= 15.
__device__ __forceinline__ uint32_t and_right(uint32_t c, uint32_t r, int b) {
    unsigned long long v = ((unsigned long long)r << 32) | c;
    for (int w = 1; w < b + 1;) {
        const int s = min(w, b + 1 - w);
        v &= v >> s;
        w += s;
    }
    return (uint32_t)v;
}
// AND over dx = 0..a of bit (x - dx) (l = previous word)
__device__ __forceinline__ uint32_t and_left(uint32_t l, uint32_t c, int a) {
    unsigned long long v = ((unsigned long long)c << 32) | l;
    for (int w = 1; w < a + 1;) {
        const int s = min(w, a + 1 - w);
        v &= v << s;
        w += s;
    }
    return (uint32_t)(v >> 32);
}

// Row-span erosion: every SE row i is one run [-a_i, +b_i], so
// out[y] = AND_i hspan_i(in[y + i - r]).  K > 0 unrolls the row loop.
template <int K>
__device__ __forceinline__ void erode_pass_spans(const ErodeParams& P, const uint32_t* __restrict__ src,
                                                 uint32_t* __restrict__ dst, int ga, int gb) {
    const int nrows = gb - ga;
    const int items = nrows * P.wpr;
    const int k = K > 0 ? K : P.k;
    const int rad = k / 2;
    const uint32_t padmask = (P.W & 31) ? ~((1u << (P.W & 31)) - 1u) : 0u;
    for (int it = threadIdx.x; it < items; it += ER_THREADS) {
        const int lr = mdiv(it, P.wpr_magic);
        const int wx = it - lr * P.wpr;
        uint32_t acc = 0xFFFFFFFFu;
#pragma unroll
        for (int i = 0; i < k; ++i) {
            const int a = P.span_a[i], b = P.span_b[i];
            if (a < 0) continue;
            const int lr2 = lr + i - rad;
            if (lr2 < 0 || lr2 >= nrows) continue;
            const uint32_t* row = src + lr2 * P.wpr;
            uint32_t m = row[wx];
            if (b > 0) m = and_right(m, wx + 1 < P.wpr ? row[wx + 1] : 0xFFFFFFFFu, b);
            if (a > 0) m &= and_left(wx > 0 ? row[wx - 1] : 0xFFFFFFFFu, row[wx], a);
            acc &= m;
        }
        if (wx == P.wpr - 1) acc |= padmask;
        dst[it] = acc;
    }
}

__device__ __forceinline__ void erode_dispatch(const ErodeParams& P, const uint32_t* src, uint32_t* dst, int ga,
                                               int gb) {
    if (!P.spans_ok) erode_pass(P, src, dst, ga, gb);
    else if (P.k == 3) erode_pass_spans<3>(P, src, dst, ga, gb);
    else if (P.k == 5) erode_pass_spans<5>(P, src, dst, ga, gb);
    else if (P.k == 7) erode_pass_spans<7>(P, src, dst, ga, gb);
    else erode_pass_spans<0>(P, src, dst, ga, gb);
}

__device__ __forceinline__ void erode_and_store(const ErodeParams& P, uint32_t* bufA, uint32_t* bufB, int ga,
                                                int gb, int r0, int r1, uint32_t* __restrict__ out_mask,
                                                int32_t* bbox) {
    uint32_t* cur = bufA;
    uint32_t* nxt = bufB;
    for (int itn = 0; itn < P.iterations; ++itn) {
        __syncthreads();
        erode_dispatch(P, cur, nxt, ga, gb);
        uint32_t* t = cur; cur = nxt; nxt = t;
    }
    __syncthreads();
    const uint32_t keep = (P.W & 31) ? ((1u << (P.W & 31)) - 1u) : 0xFFFFFFFFu;
    const int items = (r1 - r0) * P.wpr;
    const uint32_t* srcw = cur + (r0 - ga) * P.wpr;
    uint32_t* dstw = out_mask + (int64_t)r0 * P.wpr;
    BBoxAcc acc;
    for (int it = threadIdx.x; it < items; it += ER_THREADS) {
        uint32_t w = srcw[it];
        const int lr = mdiv(it, P.wpr_magic), wx = it - lr * P.wpr;
        if (wx == P.wpr - 1) w &= keep;
        dstw[it] = w;
        acc.add(w, r0 + lr, wx);
    }
    if (bbox) acc.flush(bbox);
}

// ---------------------------------------------------------------------------
// fast path
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(ER_THREADS)
lf_erode_bulk_kernel(const __grid_constant__ ErodeParams P, const uint8_t* __restrict__ masks,
                     uint32_t* __restrict__ packed, int32_t* __restrict__ bbox) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[ER_STAGES];
    uint8_t* stage = smem;
    uint32_t* bufA = (uint32_t*)(smem + ER_STAGES * ER_CHUNK);
    uint32_t* bufB = bufA + P.buf_words;

    const int64_t mask = blockIdx.x / P.strips;
    const int strip = blockIdx.x - (int)(mask * P.strips);
    int r0, r1, ga, gb;
    strip_rows(P, strip, r0, r1, ga, gb);
    const uint8_t* src = masks + (mask * P.H + ga) * (int64_t)P.W;
    const int L = (gb - ga) * P.W;                       // bytes in the run (multiple of 32)
    const int nch = (L + ER_CHUNK - 1) / ER_CHUNK;

    if (threadIdx.x == 0) {
        for (int s = 0; s < ER_STAGES; ++s) mbar_init(&full_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 0; i < ER_STAGES && i < nch; ++i) {
            const uint32_t bytes = (uint32_t)min(ER_CHUNK, L - i * ER_CHUNK);
            mbar_expect_tx(&full_bar[i], bytes);
            bulk_g2s(stage + i * ER_CHUNK, src + (int64_t)i * ER_CHUNK, bytes, &full_bar[i]);
        }
    }
    // two stages per block barrier: halves the per-byte loop / barrier overhead
    for (int i = 0; i < nch; i += 2) {
        const int s0 = i % ER_STAGES, s1 = (i + 1) % ER_STAGES;
        const uint32_t par = (uint32_t)((i / ER_STAGES) & 1);
        const int off0 = i * ER_CHUNK + 32 * threadIdx.x;
        const int off1 = off0 + ER_CHUNK;
        mbar_wait(&full_bar[s0], par);
        uint4 a0, a1, b0, b1;
        if (off0 < L) {
            const uint4* q = (const uint4*)(stage + s0 * ER_CHUNK + 32 * threadIdx.x);
            a0 = q[0]; a1 = q[1];
        }
        if (i + 1 < nch) mbar_wait(&full_bar[s1], par);
        if (off1 < L) {
            const uint4* q = (const uint4*)(stage + s1 * ER_CHUNK + 32 * threadIdx.x);
            b0 = q[0]; b1 = q[1];
        }
        __syncthreads();                                  // both stages are in registers now
        if (threadIdx.x == 0) {
            fence_proxy_async_smem();
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int j = i + t + ER_STAGES;
                if (j < nch) {
                    const int sj = j % ER_STAGES;
                    const uint32_t bytes = (uint32_t)min(ER_CHUNK, L - j * ER_CHUNK);
                    mbar_expect_tx(&full_bar[sj], bytes);
                    bulk_g2s(stage + sj * ER_CHUNK, src + (int64_t)j * ER_CHUNK, bytes, &full_bar[sj]);
                }
            }
        }
        if (off0 < L) bufA[off0 >> 5] = pack32(a0, a1);
        if (off1 < L) bufA[off1 >> 5] = pack32(b0, b1);
    }
    erode_and_store(P, bufA, bufB, ga, gb, r0, r1, packed + mask * (int64_t)P.H * P.wpr,
                    bbox ? bbox + 4 * mask : nullptr);
}

template <class SE, int MODE>
__global__ void __launch_bounds__(ER_THREADS)
lf_erode_fast_kernel(const __grid_constant__ ErodeParams P, const uint8_t* __restrict__ masks,
                     uint32_t* __restrict__ packed, int32_t* __restrict__ bbox) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t full_bar[ER_STAGES];
    if (threadIdx.x == 0) {
        for (int s = 0; s < ER_STAGES; ++s) mbar_init(&full_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t cbase = 0;
    erode_fast_item<SE, ER_STAGES, MODE>(P, masks, packed, bbox, blockIdx.x, smem, full_bar, cbase);
}

// masks that arrive bit-packed (mode 3, whole quads per row): one CTA = (mask, strip), rows copied
// straight into the guarded layout
template <class SE>
__global__ void __launch_bounds__(ER_THREADS, 8)
lf_erode_packed_kernel(const __grid_constant__ ErodeParams P, const uint32_t* __restrict__ words,
                       uint32_t* __restrict__ packed, int32_t* __restrict__ bbox) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bar;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    uint32_t cbase = 0;
    erode_packed_item<SE>(P, words, packed, bbox, blockIdx.x, smem, &bar, cbase);
}

// Masks that arrive bit-packed, 3x3 element (cross = cv2's 3x3 ellipse, or rectangle), one iteration -- the
// reference's default erosion (cvs_erosion.py:96-106) on the packed transfer format: no shared memory, no barrier.
// A thread owns four neighbouring words of a row (one 128-bit load each for the row, the row above and the row
// below; the words left and right of the quad as two 32-bit loads that hit the neighbours' lines); a word's
// horizontal erosion is two funnel shifts and an AND, "outside the image = 1" is a constant instead of a guard
// row.  ~12 instructions per word against the ~54 of the shared-memory kernel, which stays for larger elements
// and more iterations.  CTA = (mask, strip of rows) so that the bounding box costs four atomics per warp.
constexpr int EPD_STRIPS = 4;

template <bool RECT>
__global__ void __launch_bounds__(256)
lf_erode_packed_direct_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, int32_t* __restrict__ bbox, int H,
                              int qpr) {
    const int64_t mask = blockIdx.x / EPD_STRIPS;
    const int strip = blockIdx.x - (int)(mask * EPD_STRIPS);
    const int rows = (H + EPD_STRIPS - 1) / EPD_STRIPS;
    const int r0 = strip * rows, r1 = min(H, r0 + rows);
    const uint4* min4 = in + mask * (int64_t)H * qpr;
    const uint32_t* min1 = (const uint32_t*)min4;
    uint4* mout = out + mask * (int64_t)H * qpr;
    const uint4 ONES = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
    // horizontal erosion of the four words of a row, `l` / `r` = the words left and right of them
    auto hero = [](const uint4 w, uint32_t l, uint32_t r) {
        uint4 o;
        o.x = w.x & __funnelshift_r(w.x, w.y, 1) & __funnelshift_l(l, w.x, 1);
        o.y = w.y & __funnelshift_r(w.y, w.z, 1) & __funnelshift_l(w.x, w.y, 1);
        o.z = w.z & __funnelshift_r(w.z, w.w, 1) & __funnelshift_l(w.y, w.z, 1);
        o.w = w.w & __funnelshift_r(w.w, r, 1) & __funnelshift_l(w.z, w.w, 1);
        return o;
    };
    BBoxAcc acc;
    const int nq = (r1 - r0) * qpr;                          // < 2^16 (the launch checks): q / qpr by one multiplication
    const unsigned qmagic = (unsigned)((0x100000000ULL + (unsigned)qpr - 1) / (unsigned)qpr);
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
        const int lr = mdiv(q, qmagic), qx = q - lr * qpr, y = r0 + lr;
        const int at = y * qpr + qx;                        // quad index inside the mask
        const bool up = y > 0, dn = y + 1 < H, lf = qx > 0, rt = qx + 1 < qpr;
        const uint4 c = __ldg(min4 + at);
        const uint4 u = up ? __ldg(min4 + at - qpr) : ONES;
        const uint4 d = dn ? __ldg(min4 + at + qpr) : ONES;
        const uint32_t cl = lf ? __ldg(min1 + 4 * at - 1) : 0xFFFFFFFFu, cr = rt ? __ldg(min1 + 4 * at + 4) : 0xFFFFFFFFu;
        uint4 v = hero(c, cl, cr);
        if (RECT) {
            const uint32_t ul = (up && lf) ? __ldg(min1 + 4 * (at - qpr) - 1) : 0xFFFFFFFFu;
            const uint32_t ur = (up && rt) ? __ldg(min1 + 4 * (at - qpr) + 4) : 0xFFFFFFFFu;
            const uint32_t dl = (dn && lf) ? __ldg(min1 + 4 * (at + qpr) - 1) : 0xFFFFFFFFu;
            const uint32_t dr = (dn && rt) ? __ldg(min1 + 4 * (at + qpr) + 4) : 0xFFFFFFFFu;
            const uint4 hu = hero(u, ul, ur), hd = hero(d, dl, dr);
            v.x &= hu.x & hd.x; v.y &= hu.y & hd.y; v.z &= hu.z & hd.z; v.w &= hu.w & hd.w;
        } else {
            v.x &= u.x & d.x; v.y &= u.y & d.y; v.z &= u.z & d.z; v.w &= u.w & d.w;
        }
        mout[at] = v;
        if (v.x | v.y | v.z | v.w) {
            acc.add(v.x, y, 4 * qx); acc.add(v.y, y, 4 * qx + 1);
            acc.add(v.z, y, 4 * qx + 2); acc.add(v.w, y, 4 * qx + 3);
        }
    }
    if (bbox) acc.flush(bbox + 4 * mask);
}

template <class SE, int MODE>
cudaError_t launch_fast_mode(const ErodeParams& P, const uint8_t* masks, uint32_t* packed, int32_t* bbox,
                             unsigned grid, size_t smem, cudaStream_t s) {
    cudaError_t e = cudaFuncSetAttribute(lf_erode_fast_kernel<SE, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)smem);
    if (e != cudaSuccess) return e;
    lf_erode_fast_kernel<SE, MODE><<<grid, ER_THREADS, smem, s>>>(P, masks, packed, bbox);
    return cudaGetLastError();
}

// the direct kernel's case: packed words in, 3x3 element, one pass, index ranges of its 32-bit arithmetic
static bool packed_direct_ok(const ErodeParams& P, int mode) {
    return mode == 3 && P.k == 3 && P.iterations == 1 && (int64_t)P.H * P.wpr < ((int64_t)1 << 29) &&
           (int64_t)((P.H + EPD_STRIPS - 1) / EPD_STRIPS) * (P.wpr >> 2) + 256 < 65536;
}

template <class SE>
cudaError_t launch_fast(const ErodeParams& P, int mode, const uint8_t* masks, uint32_t* packed, int32_t* bbox,
                        unsigned grid, size_t smem, cudaStream_t s) {
    if (mode == 0) return launch_fast_mode<SE, 0>(P, masks, packed, bbox, grid, smem, s);
    if (mode == 1) return launch_fast_mode<SE, 1>(P, masks, packed, bbox, grid, smem, s);
    if (SE::k == 3 && packed_direct_ok(P, mode)) {
        // 3x3, one pass: straight from global memory through registers (whole quads per row are a condition of spec >= 0)
        lf_erode_packed_direct_kernel<(SE::hw(0) == 1)><<<(unsigned)(grid / P.strips * EPD_STRIPS), 256, 0, s>>>(
            (const uint4*)masks, (uint4*)packed, bbox, P.H, P.wpr >> 2);
        return cudaGetLastError();
    }
    if (mode == 3) {
        cudaError_t e = cudaFuncSetAttribute(lf_erode_packed_kernel<SE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return e;
        // latency-bound (load -> erode -> store per CTA): as many resident CTAs as the SM takes
        e = cudaFuncSetAttribute(lf_erode_packed_kernel<SE>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (e != cudaSuccess) return e;
        lf_erode_packed_kernel<SE><<<grid, ER_THREADS, smem, s>>>(P, (const uint32_t*)masks, packed, bbox);
        return cudaGetLastError();
    }
    return launch_fast_mode<SE, 2>(P, masks, packed, bbox, grid, smem, s);
}

// ---------------------------------------------------------------------------
// generic path: MODE 0 = uint8 non-zero, 1 = float >= 1.0 (the uint8 truncation
// of cvs_erosion.py:101 / V2:197), 2 = float > 0.5 (seg_with_pointcloud.py:167)
// ---------------------------------------------------------------------------
template <int MODE> struct MaskElem { typedef uint8_t type; };
template <> struct MaskElem<1> { typedef float type; };
template <> struct MaskElem<2> { typedef float type; };
template <> struct MaskElem<3> { typedef uint32_t type; };

template <int MODE>
__global__ void __launch_bounds__(ER_THREADS)
lf_erode_generic_kernel(const __grid_constant__ ErodeParams P, const void* __restrict__ masks_v,
                        uint32_t* __restrict__ packed, int32_t* __restrict__ bbox) {
    typedef typename MaskElem<MODE>::type T;
    extern __shared__ __align__(128) uint8_t smem[];
    uint32_t* bufA = (uint32_t*)smem;
    uint32_t* bufB = bufA + P.buf_words;
    const T* masks = (const T*)masks_v;

    const int64_t mask = blockIdx.x / P.strips;
    const int strip = blockIdx.x - (int)(mask * P.strips);
    int r0, r1, ga, gb;
    strip_rows(P, strip, r0, r1, ga, gb);
    const int items = (gb - ga) * P.wpr;
    if (MODE == 3) {
        // packed words in: copy the rows, bits beyond W read as ones (erosion border)
        const uint32_t* srcw = (const uint32_t*)masks_v + (mask * P.H + ga) * (int64_t)P.wpr;
        const uint32_t padmask = (P.W & 31) ? ~((1u << (P.W & 31)) - 1u) : 0u;
        for (int it = threadIdx.x; it < items; it += ER_THREADS) {
            const int lr = mdiv(it, P.wpr_magic);
            uint32_t w = srcw[it];
            if (it - lr * P.wpr == P.wpr - 1) w |= padmask;
            bufA[it] = w;
        }
        erode_and_store(P, bufA, bufB, ga, gb, r0, r1, packed + mask * (int64_t)P.H * P.wpr,
                        bbox ? bbox + 4 * mask : nullptr);
        return;
    }
    const T* src = masks + (mask * P.H + ga) * (int64_t)P.W;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int it = warp; it < items; it += ER_THREADS / 32) {
        const int lr = it / P.wpr, wx = it - lr * P.wpr;
        const int x = 32 * wx + lane;
        bool bit = true;                                  // beyond W: ones (erosion border)
        if (x < P.W) {
            const T val = src[(int64_t)lr * P.W + x];
            if (MODE == 0) bit = val != 0;
            else if (MODE == 1) bit = (float)val >= 1.0f;
            else bit = (float)val > 0.5f;
        }
        const uint32_t w = __ballot_sync(0xFFFFFFFFu, bit);
        if (lane == 0) bufA[it] = w;
    }
    erode_and_store(P, bufA, bufB, ga, gb, r0, r1, packed + mask * (int64_t)P.H * P.wpr,
                    bbox ? bbox + 4 * mask : nullptr);
}

// Row map of packed masks that were written in full (the paths without the in-kernel map): one warp per
// (mask, row), bit y of row_map[mask][y >> 5] = some word of row y is non-zero.
__global__ void __launch_bounds__(256)
lf_row_map_kernel(const uint32_t* __restrict__ packed, int64_t rows, int H, int wpr, int rm_wpr,
                  uint32_t* __restrict__ row_map) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += nw) {
        uint32_t any = 0u;
        for (int wx = lane; wx < wpr; wx += 32) any |= __ldg(packed + r * wpr + wx);
        if (__any_sync(0xFFFFFFFFu, any != 0u) && lane == 0) {
            const int64_t mask = r / H;
            const int y = (int)(r - mask * H);
            atomicOr(row_map + mask * rm_wpr + (y >> 5), 1u << (y & 31));
        }
    }
}

// After the kernels that skip rows without a set pixel: rows INSIDE a mask's bounding-box row range that were
// skipped (a mask of two blobs, one above the other) are written as zeros, so that "every row from the first to
// the last row with a set pixel is valid" holds -- what lf_fuse_batch_rows relies on.  One warp per mask; for the
// usual single-blob mask there is nothing to do.
__global__ void __launch_bounds__(256)
lf_row_gap_fill_kernel(uint32_t* __restrict__ packed, const int32_t* __restrict__ bbox, const uint32_t* __restrict__ row_map,
                       int64_t m, int H, int wpr, int rm_wpr) {
    const int lane = threadIdx.x & 31;
    const int64_t nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t mask = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; mask < m; mask += nw) {
        const int4 r = __ldg((const int4*)bbox + mask);
        if (r.z <= 0) continue;                               // empty mask
        const int y0 = (1 << 20) - r.y, y1 = r.w - 1;
        for (int yw = y0 >> 5; yw <= (y1 >> 5); ++yw) {
            uint32_t gaps = ~__ldg(row_map + mask * rm_wpr + yw);
            if (yw == (y0 >> 5)) gaps &= 0xFFFFFFFFu << (y0 & 31);
            if (yw == (y1 >> 5)) gaps &= 0xFFFFFFFFu >> (31 - (y1 & 31));
            while (gaps) {
                const int y = 32 * yw + __ffs((int)gaps) - 1;
                gaps &= gaps - 1;
                uint32_t* row = packed + (mask * H + y) * (int64_t)wpr;
                for (int wx = lane; wx < wpr; wx += 32) row[wx] = 0u;
            }
        }
    }
}

__global__ void lf_unpack_kernel(const uint32_t* __restrict__ packed, int64_t total_px, int H, int W, int wpr,
                                 uint8_t* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t plane = (int64_t)H * W;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total_px; i += stride) {
        const int64_t m = i / plane;
        const int rem = (int)(i - m * plane);
        const int y = rem / W, x = rem - y * W;
        const uint32_t w = __ldg(packed + (m * H + y) * wpr + (x >> 5));
        out[i] = (uint8_t)((w >> (x & 31)) & 1u);
    }
}

}  // namespace

int lf_erode_plan(const lf_ctx* ctx, const void* masks, int mode, int64_t m, int H, int W, const uint8_t* se, int k,
                  int iterations, int max_rows, int stages, ErodeParams* Pout, int* spec_out, int* fast_out,
                  size_t* smem_out) {
    ErodeParams& P = *Pout;
    memset(&P, 0, sizeof(P));
    P.H = H; P.W = W; P.wpr = (W + 31) / 32;
    P.iterations = iterations;
    P.k = iterations > 0 ? k : 1;
    P.r = P.k / 2;
    bool any = false;
    if (iterations > 0) {
        for (int i = 0; i < k; ++i) {
            unsigned short bits = 0;
            for (int j = 0; j < k; ++j)
                if (se[i * k + j]) bits |= (unsigned short)(1u << j);
            P.se_rows[i] = bits;
            any |= bits != 0;
        }
        if (!any) return LF_E_ARG;                        // an empty element has no cv2 meaning
    }
    P.spans_ok = 1;
    for (int i = 0; i < P.k; ++i) {
        const unsigned bits = iterations > 0 ? P.se_rows[i] : (i == 0 ? 1u : 0u);
        if (!bits) { P.span_a[i] = -1; P.span_b[i] = 0; continue; }
        int lo = 0, hi = P.k - 1;
        while (!((bits >> lo) & 1u)) ++lo;
        while (!((bits >> hi) & 1u)) --hi;
        for (int j = lo; j <= hi; ++j)
            if (!((bits >> j) & 1u)) P.spans_ok = 0;
        if (lo > P.r || hi < P.r) P.spans_ok = 0;         // run must straddle the anchor column
        P.span_a[i] = (signed char)(P.r - lo);
        P.span_b[i] = (signed char)(hi - P.r);
    }
    P.halo = iterations * P.r;

    // strip plan: enough CTAs to fill the machine several times over, strips <= ER_MAX_ROWS rows
    int64_t want = ((int64_t)ctx->sm_count * 8 + m - 1) / m;
    int min_strips = (H + max_rows - 1) / max_rows;
    int max_strips = (H + 15) / 16;
    int strips = (int)(want < min_strips ? min_strips : (want > max_strips ? max_strips : want));
    P.rows_per_strip = (H + strips - 1) / strips;
    P.strips = (H + P.rows_per_strip - 1) / P.rows_per_strip;
    const int buf_rows = P.rows_per_strip + 2 * P.halo;
    P.buf_words = buf_rows * P.wpr;
    P.wpr_magic = (unsigned)((0x100000000ULL + P.wpr - 1) / P.wpr);
    if ((int64_t)buf_rows * P.wpr >= 65536) return LF_E_UNSUPPORTED;
    if (m * P.strips > 0x7FFFFFFF) return LF_E_UNSUPPORTED;

    // bulk-copy fast path: whole words per row, 16-byte aligned rows (uint8 or float32 pixels; packed
    // words: every strip must start and end on a 16-byte boundary, i.e. whole quads per row)
    const bool fastp = (W % 32) == 0 && (((uintptr_t)masks) & 15) == 0 && (mode != 3 || (P.wpr & 3) == 0);
    // compile-time element available?  (guarded smem layout: stride wpr+1, K/2 guard rows each side)
    int spec = -1;
    if (fastp) {
        if (iterations == 0 || se_matches<SeNone>(P)) spec = 0;
        else if (se_matches<SeCross3>(P)) spec = 1;
        else if (se_matches<SeRect3>(P)) spec = 2;
        else if (se_matches<SeEllipse5>(P)) spec = 3;
        else if (se_matches<SeRect5>(P)) spec = 4;
        else if (se_matches<SeCross5>(P)) spec = 5;
        else if (se_matches<SeEllipse7>(P)) spec = 6;
        else if (se_matches<SeRect7>(P)) spec = 7;
        else if (se_matches<SeCross7>(P)) spec = 8;
    }
    // guarded layout of the compile-time-element kernels: row stride wpr + pad, k/2 guard rows each side;
    // packed words in (mode 3) are copied there row by row (pad 4 keeps rows 16-byte aligned), no staging
    const int pad = mode == 3 ? 4 : 1;
    if (spec >= 0) P.buf_words = (buf_rows + 2 * P.r) * (P.wpr + pad) + pad;
    const size_t smem = (size_t)P.buf_words * 4 * (iterations > 0 ? 2 : 1) +
                        (fastp && mode != 3 ? stages * ER_CHUNK : 0);
    if (smem > 200 * 1024) return LF_E_UNSUPPORTED;
    *spec_out = spec; *fast_out = fastp ? 1 : 0; *smem_out = smem;
    return LF_OK;
}

extern "C" int lf_erode_pack(lf_ctx* ctx, const void* masks, int mode, int64_t m, int H, int W,
                             const uint8_t* se, int k, int iterations, uint32_t* packed, int32_t* bbox,
                             void* stream) {
    return lf_erode_pack_rows(ctx, masks, mode, m, H, W, se, k, iterations, packed, bbox, nullptr, stream);
}

extern "C" int lf_erode_pack_rows(lf_ctx* ctx, const void* masks, int mode, int64_t m, int H, int W,
                                  const uint8_t* se, int k, int iterations, uint32_t* packed, int32_t* bbox,
                                  uint32_t* row_map, void* stream) {
    if (!ctx || m < 0 || H <= 0 || W <= 0 || iterations < 0 || mode < 0 || mode > 3) return LF_E_ARG;
    if (m > 0 && (!masks || !packed)) return LF_E_ARG;
    if (mode == 3 && (const void*)masks == (const void*)packed) return LF_E_ARG;   // strips read their neighbours' rows
    if (iterations > 0 && (!se || k < 1 || (k & 1) == 0)) return LF_E_ARG;
    if (iterations > 0 && k > LF_MAX_SE) return LF_E_UNSUPPORTED;
    if (m == 0) return LF_OK;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    ErodeParams P;
    int spec, fast;
    size_t smem;
    // strip height: with a halo of one row (3x3, one pass -- the reference's default) strips of <= 64 rows, i.e. six
    // strips of a 376-row image and four CTAs per SM instead of three (measured on cfg3: 48 / 56-72 / 80 / 96 / 128
    // rows -> 1.301 / 1.274 / 1.343 / 1.360 / 1.533 ms with the row map, 1.374 against 1.448 ms without); larger
    // elements re-read more halo rows per strip and keep the 96-row strips (cfg5, 7x7: 0.368 against 0.374 ms)
    const int halo_rows = iterations > 0 ? (k / 2) * iterations : 0;
    const int strip_rows = mode == 3 ? ER_MAX_ROWS_PACKED : (halo_rows <= 1 ? ER_ROWS_SMALL_HALO : ER_MAX_ROWS);
    const int rc = lf_erode_plan(ctx, masks, mode, m, H, W, se, k, iterations, strip_rows, ER_STAGES, &P,
                                 &spec, &fast, &smem);
    if (rc != LF_OK) return rc;
    const unsigned grid = (unsigned)(m * P.strips);
    cudaStream_t s = (cudaStream_t)stream;
    // row map: the kernels that hold the eroded strip in shared memory write it themselves and skip the rows without
    // a set pixel; the others write every row, and the map is read off the packed words afterwards
    const int rm_wpr = (H + 31) / 32;
    bool map_in_kernel = false;
    if (row_map) {
        LF_CUDA(ctx, cudaMemsetAsync(row_map, 0, (size_t)m * rm_wpr * sizeof(uint32_t), s));
        map_in_kernel = spec >= 0 && !packed_direct_ok(P, mode);
        if (map_in_kernel) { P.row_map = row_map; P.rm_wpr = rm_wpr; }
    }
    auto finish = [&]() -> int {
        if (row_map && map_in_kernel && bbox) {
            int64_t blocks = (m + 7) / 8;
            const int64_t cap = (int64_t)ctx->sm_count * 16;
            if (blocks > cap) blocks = cap;
            lf_row_gap_fill_kernel<<<(unsigned)blocks, 256, 0, s>>>(packed, bbox, row_map, m, H, P.wpr, rm_wpr);
            LF_CUDA(ctx, cudaGetLastError());
        }
        if (row_map && !map_in_kernel) {
            const int64_t rows = m * H;
            int64_t blocks = (rows + 7) / 8;
            const int64_t cap = (int64_t)ctx->sm_count * 16;
            if (blocks > cap) blocks = cap;
            lf_row_map_kernel<<<(unsigned)blocks, 256, 0, s>>>(packed, rows, H, P.wpr, rm_wpr, row_map);
            LF_CUDA(ctx, cudaGetLastError());
        }
        return LF_OK;
    };
    if (spec >= 0) {
        const uint8_t* mk = (const uint8_t*)masks;
        cudaError_t e;
        switch (spec) {
            case 0: e = launch_fast<SeNone>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 1: e = launch_fast<SeCross3>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 2: e = launch_fast<SeRect3>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 3: e = launch_fast<SeEllipse5>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 4: e = launch_fast<SeRect5>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 5: e = launch_fast<SeCross5>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 6: e = launch_fast<SeEllipse7>(P, mode, mk, packed, bbox, grid, smem, s); break;
            case 7: e = launch_fast<SeRect7>(P, mode, mk, packed, bbox, grid, smem, s); break;
            default: e = launch_fast<SeCross7>(P, mode, mk, packed, bbox, grid, smem, s); break;
        }
        LF_CUDA(ctx, e);
        return finish();
    }
    if (fast && mode == 0) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_erode_bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                          (int)smem));
        lf_erode_bulk_kernel<<<grid, ER_THREADS, smem, s>>>(P, (const uint8_t*)masks, packed, bbox);
    } else if (mode == 0) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_erode_generic_kernel<0>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        lf_erode_generic_kernel<0><<<grid, ER_THREADS, smem, s>>>(P, masks, packed, bbox);
    } else if (mode == 3) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_erode_generic_kernel<3>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        lf_erode_generic_kernel<3><<<grid, ER_THREADS, smem, s>>>(P, masks, packed, bbox);
    } else if (mode == 1) {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_erode_generic_kernel<1>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        lf_erode_generic_kernel<1><<<grid, ER_THREADS, smem, s>>>(P, masks, packed, bbox);
    } else {
        LF_CUDA(ctx, cudaFuncSetAttribute(lf_erode_generic_kernel<2>,
                                          cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        lf_erode_generic_kernel<2><<<grid, ER_THREADS, smem, s>>>(P, masks, packed, bbox);
    }
    LF_CUDA(ctx, cudaGetLastError());
    return finish();
}

extern "C" int lf_unpack_masks(lf_ctx* ctx, const uint32_t* packed, int64_t m, int H, int W, uint8_t* out,
                               void* stream) {
    if (!ctx || m < 0 || H <= 0 || W <= 0 || (m > 0 && (!packed || !out))) return LF_E_ARG;
    if (m == 0) return LF_OK;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t total = m * H * W;
    int64_t blocks = (total + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 32;
    if (blocks > cap) blocks = cap;
    lf_unpack_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(packed, total, H, W, (W + 31) / 32, out);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
