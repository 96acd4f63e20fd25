// Fast path of K3 (mask binarisation + binary erosion + bit-packing) as per-item device
// functions of the erode kernels (erode.cu has the algorithm and the host side).
#pragma once
#include "common.cuh"
#include "bulk.cuh"

constexpr int ER_THREADS = 256;
constexpr int ER_STAGES = 4;          // stand-alone kernel
constexpr int ER_CHUNK = 8192;          // bytes per stage: 256 threads x 32 px
#ifndef LF_ER_MAX_ROWS
#define LF_ER_MAX_ROWS 96
#endif
constexpr int ER_MAX_ROWS = LF_ER_MAX_ROWS;   // strip height bound (smem: 2 x (96+2*halo) x wpr x 4)
constexpr int ER_ROWS_SMALL_HALO = 64;         // strip height for a halo of at most one row
constexpr int ER_MAX_ROWS_PACKED = 64;  // masks that arrive bit-packed (mode 3): 26 KB per CTA, 8 CTAs per SM

struct ErodeParams {
    int H, W, wpr;
    int rows_per_strip, strips;
    int k, r, iterations, halo;
    int buf_words;                       // words per packed smem buffer
    unsigned wpr_magic;                  // ceil(2^32 / wpr) mod 2^32: it / wpr == mdiv(it, magic) for it < 2^16
    int spans_ok;                        // every SE row is one contiguous run (RECT / CROSS / ELLIPSE are)
    unsigned short se_rows[LF_MAX_SE];   // bit j of se_rows[i] = SE[i][j]
    signed char span_a[LF_MAX_SE];       // row i covers dx in [-span_a[i], +span_b[i]]; span_a < 0: empty row
    signed char span_b[LF_MAX_SE];
    // optional row map (lf_erode_pack_rows): bit y of row_map[mask * rm_wpr + (y >> 5)] = row y of the eroded mask
    // has a set pixel; when given, rows WITHOUT a set pixel are not written to `packed` at all
    uint32_t* row_map;
    int rm_wpr;
};

// Host-side plan of one lf_erode_pack call (erode.cu): fills P, picks the compile-time
// structuring element (`spec` >= 0: index into the SeSpec list below, -1: run-time
// element), says whether the bulk-copy fast path applies and how much dynamic shared
// memory a CTA needs.  Returns LF_OK or an LF_E_* code.
int lf_erode_plan(const lf_ctx* ctx, const void* masks, int mode, int64_t m, int H, int W, const uint8_t* se, int k,
                  int iterations, int max_rows, int stages, ErodeParams* P, int* spec, int* fast, size_t* smem);

namespace {

// x / d for x < 2^16 with magic = ceil(2^32 / d) truncated to 32 bits; d == 1 wraps to 0 and means "x itself"
__device__ __forceinline__ int mdiv(int x, unsigned magic) {
    return magic ? (int)__umulhi((unsigned)x, magic) : x;
}

// Bounding box of the set pixels of the (eroded) mask, kept per thread while the
// packed words go out and merged with four atomics per CTA.  Stored so that an
// all-zero record means "empty": {2^20 - x0, 2^20 - y0, x1 + 1, y1 + 1}, all
// combined with atomicMax.  The fused point pass uses it to skip masks whose
// silhouette the points of a warp cannot touch.
struct BBoxAcc {
    int nx0, ny0, x1, y1;
    __device__ __forceinline__ BBoxAcc() : nx0(0), ny0(0), x1(0), y1(0) {}
    __device__ __forceinline__ void add(uint32_t w, int row, int wx) {
        if (!w) return;
        nx0 = max(nx0, (1 << 20) - (32 * wx + __ffs((int)w) - 1));
        x1 = max(x1, 32 * wx + 32 - __clz((int)w));
        ny0 = max(ny0, (1 << 20) - row);
        y1 = max(y1, row + 1);
    }
    __device__ __forceinline__ void flush(int32_t* bbox) {
        const int a = __reduce_max_sync(0xFFFFFFFFu, nx0), b = __reduce_max_sync(0xFFFFFFFFu, ny0);
        const int c = __reduce_max_sync(0xFFFFFFFFu, x1), d = __reduce_max_sync(0xFFFFFFFFu, y1);
        if ((threadIdx.x & 31) == 0 && c > 0) {
            atomicMax(bbox + 0, a); atomicMax(bbox + 1, b); atomicMax(bbox + 2, c); atomicMax(bbox + 3, d);
        }
    }
};

// 4 bytes -> 4 bits (bit j = byte j non-zero) in bits 0..3; the bits above are garbage.
// bit 7 of each byte of t = byte != 0; the high half of t * (2^25+2^18+2^11+2^4) lines bits
// 7, 15, 23, 31 up at 32..35 (no two partial products share a bit below, so no carries).
__device__ __forceinline__ uint32_t nz_nibble_lo(uint32_t w) {
    const uint32_t t = (((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) & 0x80808080u;
    return __umulhi(t, 0x02040810u);
}
// 32 bytes -> 32 bits, LSB = first byte.  Each step shifts the word right by four and
// drops the next nibble in at the top (one funnel shift: the garbage never gets in).
// Instance masks are mostly background: 32 zero bytes cost five instructions.
__device__ __forceinline__ uint32_t pack32(const uint4 a, const uint4 b) {
    if (((a.x | a.y | a.z) | (a.w | b.x | b.y) | (b.z | b.w)) == 0u) return 0u;
    uint32_t acc = nz_nibble_lo(a.x) << 28;
    acc = __funnelshift_r(acc, nz_nibble_lo(a.y), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(a.z), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(a.w), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(b.x), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(b.y), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(b.z), 4);
    acc = __funnelshift_r(acc, nz_nibble_lo(b.w), 4);
    return acc;
}

__device__ __forceinline__ void strip_rows(const ErodeParams& P, int strip, int& r0, int& r1, int& ga, int& gb) {
    r0 = strip * P.rows_per_strip;
    r1 = min(P.H, r0 + P.rows_per_strip);
    ga = max(0, r0 - P.halo);
    gb = min(P.H, r1 + P.halo);
}

// ---------------------------------------------------------------------------
// specialised fast path: structuring element known at compile time
// ---------------------------------------------------------------------------
// Symmetric row-span elements (every cv2 RECT / CROSS / ELLIPSE up to 7x7):
// row i and row K-1-i have the same half-width HW(i); HW < 0 = empty row.
// Packed rows live in shared memory with stride wpr+1: word 0 of every row is an
// all-ones guard that doubles as the right neighbour of the previous row, and
// K/2 all-ones guard rows sit above and below, so the erosion loop has no
// bounds checks at all (out-of-image = ones = cv2's +inf erosion border).
template <int K, int H0, int H1, int H2, int H3>
struct SeSpec {
    static constexpr int k = K;
    __host__ __device__ static constexpr int hw(int i) {
        const int j = i <= K / 2 ? i : K - 1 - i;
        return j == 0 ? H0 : (j == 1 ? H1 : (j == 2 ? H2 : H3));
    }
};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// One work item = (mask, strip) of the fast path.  `smem` = NST * ER_CHUNK bytes of
// staging followed by the two packed buffers; `full_bar` = NST initialised mbarriers
// (count 1) whose phases run on from item to item: `cbase` counts the chunks this CTA has
// streamed so far, chunk c lives in stage c % NST and completes phase c / NST.
// Executed by the ER_THREADS threads with the lowest thread indices; they meet on named
// barrier 1, so further warps of the CTA are free to do something else.
__device__ __forceinline__ void er_sync() { asm volatile("bar.sync 1, %0;" ::"n"(ER_THREADS) : "memory"); }
// the same barrier, returning the OR of `pred` over the ER_THREADS threads
__device__ __forceinline__ bool er_sync_or(bool pred) {
    uint32_t r;
    asm volatile(
        "{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %1, 0;\n\tbar.red.or.pred p, 1, %2, q;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(r) : "r"((uint32_t)pred), "n"(ER_THREADS) : "memory");
    return r != 0;
}

// 8 floats -> 8 bits (bit j = pixel j set): MODE 1 = value >= 1.0 (what the reference's uint8
// truncations of its 0.0 / 1.0 masks leave, cvs_erosion.py:101, V2:197), MODE 2 = value > 0.5
// (seg_with_pointcloud.py:167).  NaN compares false in both.  Eight +0.0 cost three instructions.
template <int MODE>
__device__ __forceinline__ uint32_t pack8f(const uint4 a, const uint4 b) {
    if (((a.x | a.y | a.z) | (a.w | b.x | b.y) | (b.z | b.w)) == 0u) return 0u;
    const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
    uint32_t r = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float f = __uint_as_float(w[j]);
        const bool set = MODE == 1 ? (f >= 1.0f) : (f > 0.5f);
        r |= (set ? 1u : 0u) << j;
    }
    return r;
}

// One erosion pass over `items` words of the guarded layout (row stride S = wpr + PAD words, PAD guard
// words of ones between rows, word (row 0, wx 0) at `data0`).  Returns the OR of the centre words the
// thread saw.
template <class SE, int PAD>
__device__ __forceinline__ uint32_t erode_pass_fast(const ErodeParams& P, const uint32_t* cur, uint32_t* nxt, int items,
                                                    int S, int data0, int tid) {
    constexpr int K = SE::k, RAD = K / 2;
    uint32_t seen = 0;
    for (int it = tid; it < items; it += ER_THREADS) {
        const int idx = it + PAD * mdiv(it, P.wpr_magic) + data0;
        const uint32_t ctr = cur[idx];
        seen |= ctr;
        if (ctr == 0u) {                                   // the anchor is part of every element here
            nxt[idx] = 0u;
            continue;
        }
        uint32_t acc = 0xFFFFFFFFu;
#pragma unroll
        for (int i = 0; i < K; ++i) {
            const int h = SE::hw(i);
            if (h < 0) continue;
            const uint32_t* rp = cur + idx + (i - RAD) * S;
            const uint32_t c = rp[0];
            acc &= c;
            if (h > 0) {
                const uint32_t l = rp[-1], r = rp[1];
#pragma unroll
                for (int d = 1; d <= h; ++d) acc &= __funnelshift_r(c, r, d) & __funnelshift_l(l, c, d);
            }
        }
        nxt[idx] = acc;
    }
    return seen;
}

// Rows [r0, r1) of the result from the guarded layout to global memory + the mask's bounding box.
// With a row map (`rowmap_rec` = the mask's words of it): a first pass over the shared-memory words notes which rows
// have a set pixel, only those rows are stored, and their bits are OR-ed into the map -- an instance mask covers a
// tenth of the image rows, so nine tenths of the packed words never travel to HBM.
template <int PAD>
__device__ __forceinline__ void store_rows_fast(const ErodeParams& P, const uint32_t* cur, uint32_t* __restrict__ dstw,
                                                int32_t* __restrict__ bbox_rec, uint32_t* __restrict__ rowmap_rec, int r0,
                                                int r1, int first, int S, int tid) {
    // `first` = word index of (row r0, wx 0) in `cur`
    __shared__ uint32_t s_rowbits[(ER_MAX_ROWS + 31) / 32 + 1];
    BBoxAcc acc;
    const bool sparse = rowmap_rec != nullptr;
    if (sparse) {
        if (tid < (int)(sizeof(s_rowbits) / 4)) s_rowbits[tid] = 0u;
        er_sync();
    }
    if ((P.wpr & 3) == 0) {
        // four words per thread: one 128-bit store, one bounding-box test per 128 pixels
        const int qpr = P.wpr >> 2;                        // quads per row
        const unsigned qmagic = (unsigned)((0x100000000ULL + qpr - 1) / qpr);
        const int nq = (r1 - r0) * qpr;
        for (int q = tid; q < nq; q += ER_THREADS) {
            const int lr = mdiv(q, qmagic);
            const int wq = q - lr * qpr;
            const uint32_t* sp = cur + lr * S + 4 * wq + first;
            const uint4 v = PAD == 4 ? *(const uint4*)sp : make_uint4(sp[0], sp[1], sp[2], sp[3]);
            if (!sparse) *(uint4*)(dstw + 4 * q) = v;
            if (v.x | v.y | v.z | v.w) {
                acc.add(v.x, r0 + lr, 4 * wq); acc.add(v.y, r0 + lr, 4 * wq + 1);
                acc.add(v.z, r0 + lr, 4 * wq + 2); acc.add(v.w, r0 + lr, 4 * wq + 3);
                if (sparse) atomicOr(&s_rowbits[lr >> 5], 1u << (lr & 31));
            }
        }
        if (sparse) {
            er_sync();
            for (int q = tid; q < nq; q += ER_THREADS) {
                const int lr = mdiv(q, qmagic);
                if (!((s_rowbits[lr >> 5] >> (lr & 31)) & 1u)) continue;
                const uint32_t* sp = cur + lr * S + 4 * (q - lr * qpr) + first;
                *(uint4*)(dstw + 4 * q) = PAD == 4 ? *(const uint4*)sp : make_uint4(sp[0], sp[1], sp[2], sp[3]);
            }
        }
    } else {
        const int out_items = (r1 - r0) * P.wpr;
        for (int it = tid; it < out_items; it += ER_THREADS) {
            const int lr = mdiv(it, P.wpr_magic);
            const uint32_t w = cur[it + PAD * lr + first];
            if (!sparse) dstw[it] = w;
            else if (w) atomicOr(&s_rowbits[lr >> 5], 1u << (lr & 31));
            acc.add(w, r0 + lr, it - lr * P.wpr);
        }
        if (sparse) {
            er_sync();
            for (int it = tid; it < out_items; it += ER_THREADS) {
                const int lr = mdiv(it, P.wpr_magic);
                if ((s_rowbits[lr >> 5] >> (lr & 31)) & 1u) dstw[it] = cur[it + PAD * lr + first];
            }
        }
    }
    if (sparse) {
        for (int lr = tid; lr < r1 - r0; lr += ER_THREADS)
            if ((s_rowbits[lr >> 5] >> (lr & 31)) & 1u) atomicOr(rowmap_rec + ((r0 + lr) >> 5), 1u << ((r0 + lr) & 31));
        er_sync();                                         // s_rowbits is free for a next item of this CTA
    }
    if (bbox_rec) acc.flush(bbox_rec);
}

__device__ __forceinline__ void store_zero_rows(const ErodeParams& P, uint32_t* __restrict__ dstw, int out_items, int tid) {
    if (P.row_map) return;                                 // no row of the strip has a set pixel: nothing to write, no bit to set
    if ((P.wpr & 3) == 0) {
        for (int q = tid; q < (out_items >> 2); q += ER_THREADS) *(uint4*)(dstw + 4 * q) = make_uint4(0u, 0u, 0u, 0u);
    } else {
        for (int it = tid; it < out_items; it += ER_THREADS) dstw[it] = 0u;
    }
}

// ones in every guard position of the layout: RAD rows above and below the window, the PAD words
// before every row (= right neighbours of the previous row) and after the last one
template <int PAD>
__device__ __forceinline__ void fill_guards(uint32_t* bufA, uint32_t* bufB, int nrows, int RAD, int S, int tid) {
    const int top = RAD * S, all = (nrows + 2 * RAD) * S + PAD;
    for (int i = tid; i < top; i += ER_THREADS) {
        bufA[i] = 0xFFFFFFFFu; bufB[i] = 0xFFFFFFFFu;
        bufA[all - 1 - i] = 0xFFFFFFFFu; bufB[all - 1 - i] = 0xFFFFFFFFu;
    }
    for (int g = tid; g < (nrows + 1) * PAD; g += ER_THREADS) {
        const int r = g / PAD, j = g - r * PAD;
        bufA[(r + RAD) * S + j] = 0xFFFFFFFFu; bufB[(r + RAD) * S + j] = 0xFFFFFFFFu;
    }
}

// MODE 0: uint8 masks (non-zero = set), 1 byte per pixel; MODE 1 / 2: float32 masks, 4 bytes per pixel.
template <class SE, int NST, int MODE = 0>
__device__ __forceinline__ void erode_fast_item(const ErodeParams& P, const uint8_t* __restrict__ masks,
                                                uint32_t* __restrict__ packed, int32_t* __restrict__ bbox,
                                                int64_t item, uint8_t* smem, uint64_t* full_bar, uint32_t& cbase) {
    constexpr int K = SE::k, RAD = K / 2;
    const uint32_t stage_a = smem_u32(smem);
    uint32_t* bufA = (uint32_t*)(smem + NST * ER_CHUNK);
    uint32_t* bufB = bufA + P.buf_words;

    const unsigned mask32 = (unsigned)item / (unsigned)P.strips;          // item < 2^31 (lf_erode_plan)
    const int64_t mask = mask32;
    const int strip = (int)((unsigned)item - mask32 * (unsigned)P.strips);
    int r0, r1, ga, gb;
    strip_rows(P, strip, r0, r1, ga, gb);
    const int row_bytes = MODE == 0 ? P.W : 4 * P.W;
    const uint8_t* src = masks + (mask * P.H + ga) * (int64_t)row_bytes;
    const int nrows = gb - ga;
    const int L = nrows * row_bytes;                       // bytes in the run (a multiple of 32)
    const int nch = (L + ER_CHUNK - 1) / ER_CHUNK;
    const int S = P.wpr + 1;                               // row stride in words
    const int tid = threadIdx.x;

    if (tid == 0) {
        fence_proxy_async_smem();                          // the previous item read these stages
        for (int i = 0; i < NST && i < nch; ++i) {
            const int si = (int)((cbase + i) % NST);
            const uint32_t bytes = (uint32_t)min(ER_CHUNK, L - i * ER_CHUNK);
            mbar_expect_tx(&full_bar[si], bytes);
            bulk_g2s(smem + si * ER_CHUNK, src + (int64_t)i * ER_CHUNK, bytes, &full_bar[si]);
        }
    }
    if (P.iterations > 0) fill_guards<1>(bufA, bufB, nrows, RAD, S, tid);
    const int data0 = RAD * S + 1;                         // word index of (row 0, wx 0)
    uint32_t nz = 0;                                       // OR of every word this thread produced

    for (int i = 0; i < nch; i += 2) {
        const uint32_t c0 = cbase + (uint32_t)i, c1 = c0 + 1u;
        const int s0 = (int)(c0 % NST), s1 = (int)(c1 % NST);
        const uint32_t par0 = (c0 / NST) & 1u, par1 = (c1 / NST) & 1u;
        const int w0 = i * (ER_CHUNK / 32) + tid;          // flat word index of this thread's 32 px
        const int w1 = w0 + ER_CHUNK / 32;
        const bool p0 = w0 * 32 < L, p1 = w1 * 32 < L;
        uint4 a0, a1, b0, b1;
        mbar_wait(&full_bar[s0], par0);
        if (p0) { a0 = lds128(stage_a + s0 * ER_CHUNK + 32 * tid); a1 = lds128(stage_a + s0 * ER_CHUNK + 32 * tid + 16); }
        if (i + 1 < nch) mbar_wait(&full_bar[s1], par1);
        if (p1) { b0 = lds128(stage_a + s1 * ER_CHUNK + 32 * tid); b1 = lds128(stage_a + s1 * ER_CHUNK + 32 * tid + 16); }
        er_sync();                                         // both stages are in registers now
        if (tid == 0) {
            fence_proxy_async_smem();
#pragma unroll
            for (int t = 0; t < 2; ++t) {
                const int j = i + t + NST;
                if (j < nch) {
                    const int sj = (int)((cbase + j) % NST);
                    const uint32_t bytes = (uint32_t)min(ER_CHUNK, L - j * ER_CHUNK);
                    mbar_expect_tx(&full_bar[sj], bytes);
                    bulk_g2s(smem + sj * ER_CHUNK, src + (int64_t)j * ER_CHUNK, bytes, &full_bar[sj]);
                }
            }
        }
        if (MODE == 0) {
            // flat word w -> row w / wpr; with the guard column the smem index is w + row + data0
            if (p0) { const uint32_t w = pack32(a0, a1); nz |= w; bufA[w0 + mdiv(w0, P.wpr_magic) + data0] = w; }
            if (p1) { const uint32_t w = pack32(b0, b1); nz |= w; bufA[w1 + mdiv(w1, P.wpr_magic) + data0] = w; }
        } else {
            // float32: the thread's 32 bytes are 8 pixels = one byte of a packed word (w0 / w1 count
            // 8-pixel groups here); little-endian bytes are in bit order, so a byte store does it
            uint8_t* bytesA = (uint8_t*)bufA;
            const int q0 = w0 >> 2, q1 = w1 >> 2;
            if (p0) {
                const uint32_t w = pack8f<MODE>(a0, a1);
                nz |= w;
                bytesA[4 * (q0 + mdiv(q0, P.wpr_magic) + data0) + (w0 & 3)] = (uint8_t)w;
            }
            if (p1) {
                const uint32_t w = pack8f<MODE>(b0, b1);
                nz |= w;
                bytesA[4 * (q1 + mdiv(q1, P.wpr_magic) + data0) + (w1 & 3)] = (uint8_t)w;
            }
        }
    }

    uint32_t* dstw = packed + (mask * (int64_t)P.H + r0) * P.wpr;
    cbase += (uint32_t)nch;
    // Instance masks are mostly background: a window without a set pixel erodes to nothing, so the
    // strip is written as zeros straight away (the barrier doubles as the one after the fill).
    if (!er_sync_or(nz != 0u)) {
        store_zero_rows(P, dstw, (r1 - r0) * P.wpr, tid);
        return;
    }
    uint32_t* cur = bufA;
    uint32_t* nxt = bufB;
    for (int itn = 0; itn < P.iterations; ++itn) {
        if (itn) er_sync();
        erode_pass_fast<SE, 1>(P, cur, nxt, nrows * P.wpr, S, data0, tid);
        uint32_t* t = cur; cur = nxt; nxt = t;
    }
    if (P.iterations > 0) er_sync();
    store_rows_fast<1>(P, cur, dstw, bbox ? bbox + 4 * mask : nullptr, P.row_map ? P.row_map + mask * P.rm_wpr : nullptr, r0,
                       r1, data0 + (r0 - ga) * S, S, tid);
}

// Masks that arrive bit-packed already (mode 3: uint32 [M][H][wpr], LSB = lowest x; whole quads per
// row).  No staging and no fill pass: one bulk copy per row drops the words straight into the guarded
// layout (row stride wpr + 4 words keeps every row 16-byte aligned), the first erosion pass notices
// an all-zero window on the way.  `smem` = the two buffers; `bar` = one mbarrier (count 1) whose
// phase runs on from item to item (`cbase` counts items).
template <class SE>
__device__ __forceinline__ void erode_packed_item(const ErodeParams& P, const uint32_t* __restrict__ words,
                                                  uint32_t* __restrict__ packed, int32_t* __restrict__ bbox,
                                                  int64_t item, uint8_t* smem, uint64_t* bar, uint32_t& cbase) {
    constexpr int K = SE::k, RAD = K / 2, PAD = 4;
    uint32_t* bufA = (uint32_t*)smem;
    uint32_t* bufB = bufA + P.buf_words;
    const unsigned mask32 = (unsigned)item / (unsigned)P.strips;
    const int64_t mask = mask32;
    const int strip = (int)((unsigned)item - mask32 * (unsigned)P.strips);
    int r0, r1, ga, gb;
    strip_rows(P, strip, r0, r1, ga, gb);
    const int nrows = gb - ga;
    const int S = P.wpr + PAD;
    const int data0 = RAD * S + PAD;
    const int tid = threadIdx.x;
    const uint32_t* src = words + (mask * P.H + ga) * (int64_t)P.wpr;
    const uint32_t row_bytes = 4u * (uint32_t)P.wpr;

    if (tid < 32) {                                        // warp 0 issues the row copies
        if (tid == 0) {
            fence_proxy_async_smem();                      // the previous item read these buffers
            mbar_expect_tx(bar, row_bytes * (uint32_t)nrows);
        }
        __syncwarp();
        for (int r = tid; r < nrows; r += 32)
            bulk_g2s(bufA + data0 + r * S, src + (int64_t)r * P.wpr, row_bytes, bar);
    }
    if (P.iterations > 0) fill_guards<PAD>(bufA, bufB, nrows, RAD, S, tid);
    uint32_t* dstw = packed + (mask * (int64_t)P.H + r0) * P.wpr;
    mbar_wait(bar, cbase & 1u);
    cbase += 1u;
    uint32_t* cur = bufA;
    uint32_t* nxt = bufB;
    if (P.iterations > 0) {
        er_sync();                                         // guards written
        uint32_t seen = 0;
        for (int itn = 0; itn < P.iterations; ++itn) {
            if (itn) er_sync();
            const uint32_t sn = erode_pass_fast<SE, PAD>(P, cur, nxt, nrows * P.wpr, S, data0, tid);
            if (itn == 0) seen = sn;
            uint32_t* t = cur; cur = nxt; nxt = t;
        }
        if (!er_sync_or(seen != 0u)) {                     // nothing set anywhere in the window
            store_zero_rows(P, dstw, (r1 - r0) * P.wpr, tid);
            return;
        }
    }
    store_rows_fast<PAD>(P, cur, dstw, bbox ? bbox + 4 * mask : nullptr, P.row_map ? P.row_map + mask * P.rm_wpr : nullptr,
                         r0, r1, data0 + (r0 - ga) * S, S, tid);
}

typedef SeSpec<3, 0, 1, 0, 0> SeCross3;     // = MORPH_ELLIPSE 3x3 (cvs_erosion.py:98 default)
typedef SeSpec<3, 1, 1, 0, 0> SeRect3;
typedef SeSpec<5, 0, 2, 2, 0> SeEllipse5;
typedef SeSpec<5, 2, 2, 2, 0> SeRect5;
typedef SeSpec<5, 0, 0, 2, 0> SeCross5;
typedef SeSpec<7, 0, 2, 3, 3> SeEllipse7;
typedef SeSpec<7, 3, 3, 3, 3> SeRect7;
typedef SeSpec<7, 0, 0, 0, 3> SeCross7;
typedef SeSpec<1, 0, 0, 0, 0> SeNone;       // pack only / 1x1

template <class SE>
bool se_matches(const ErodeParams& P) {
    if (P.k != SE::k || !P.spans_ok) return false;
    for (int i = 0; i < SE::k; ++i)
        if (P.span_a[i] != SE::hw(i) || P.span_b[i] != SE::hw(i)) return false;
    return true;
}

}  // namespace
