// lf_pack_masks_host: binarise + bit-pack instance masks on the SENDING side (plain C++,
// no device work), so that masks cross PCIe and enter HBM at one bit per pixel.  This is
// a transfer encoding, not a compute fallback: erosion, the gather and everything after
// it run on the device from the packed words (lf_erode_pack mode 3).
//
// uint8 path: 8 bytes -> 8 bits with one 64-bit SWAR step (the same non-zero test as
// pack32 in erode_body.cuh) and a multiply that gathers the eight byte MSBs; with AVX2
// (checked at run time) 32 bytes -> 32 bits by compare + movemask, with AVX-512BW 64 bytes
// -> 64 bits by one vptestmb.  The loop is bound by how many cache-line fills one core
// keeps in flight, so every path prefetches PF bytes ahead (measured on the 16-core host
// of the B200 box: AVX2 101 GB/s, AVX-512 + prefetch 157 GB/s on uint8 masks).
#include <stdint.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#if defined(__x86_64__)
#include <immintrin.h>
#endif
#include "lidarfuse.h"

namespace {

inline uint32_t nz8(uint64_t w) {
    // bit 7 of every byte = byte != 0, then gather the eight MSBs into one byte
    const uint64_t t = (((w & 0x7f7f7f7f7f7f7f7fULL) + 0x7f7f7f7f7f7f7f7fULL) | w) & 0x8080808080808080ULL;
    return (uint32_t)((t * 0x0002040810204081ULL) >> 56);
}

void pack_row_u8(const uint8_t* src, int W, uint32_t* dst) {
    const int full = W >> 5;
    for (int w = 0; w < full; ++w) {
        uint64_t q[4];
        memcpy(q, src + 32 * w, 32);
        dst[w] = nz8(q[0]) | (nz8(q[1]) << 8) | (nz8(q[2]) << 16) | (nz8(q[3]) << 24);
    }
    if (W & 31) {
        uint32_t acc = 0;
        for (int x = 32 * full; x < W; ++x) acc |= (uint32_t)(src[x] != 0) << (x & 31);
        dst[full] = acc;
    }
}

#if defined(__x86_64__)
constexpr int PF = 4096;      // prefetch distance in bytes (never dereferenced: may run past the end)

__attribute__((target("avx2"))) void pack_row_u8_avx2(const uint8_t* src, int W, uint32_t* dst) {
    const int full = W >> 5;
    const __m256i z = _mm256_setzero_si256();
    for (int w = 0; w < full; ++w) {
        if (!(w & 1)) _mm_prefetch((const char*)(src + 32 * w + PF), _MM_HINT_T0);
        const __m256i v = _mm256_loadu_si256((const __m256i*)(src + 32 * w));
        dst[w] = ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(v, z));
    }
    if (W & 31) {
        uint32_t acc = 0;
        for (int x = 32 * full; x < W; ++x) acc |= (uint32_t)(src[x] != 0) << (x & 31);
        dst[full] = acc;
    }
}
#endif

#if defined(__x86_64__)
__attribute__((target("avx512f,avx512bw"))) void pack_row_u8_avx512(const uint8_t* src, int W, uint32_t* dst) {
    const int full = W >> 6;
    const bool nt = ((uintptr_t)dst & 7) == 0;
    for (int w = 0; w < full; ++w) {
        _mm_prefetch((const char*)(src + 64 * w + PF), _MM_HINT_T0);
        const __m512i v = _mm512_loadu_si512((const void*)(src + 64 * w));
        const uint64_t m = (uint64_t)_mm512_test_epi8_mask(v, v);            // bit i = byte i != 0
        // streaming store: the words are read next by the DMA engine, not by this core, and a
        // plain store would first pull the destination line in (read-for-ownership)
        if (nt) {
            _mm_stream_si64((long long*)(dst + 2 * w), (long long)m);
        } else {
            dst[2 * w] = (uint32_t)m;
            dst[2 * w + 1] = (uint32_t)(m >> 32);
        }
    }
    for (int x0 = 64 * full; x0 < W; x0 += 32) {
        uint32_t acc = 0;
        const int n = W - x0 < 32 ? W - x0 : 32;
        for (int j = 0; j < n; ++j) acc |= (uint32_t)(src[x0 + j] != 0) << j;
        dst[x0 >> 5] = acc;
    }
}
#endif

template <int MODE>
void pack_row_f32(const float* src, int W, uint32_t* dst) {
    const int wpr = (W + 31) >> 5;
    for (int w = 0; w < wpr; ++w) {
        const int n = W - 32 * w < 32 ? W - 32 * w : 32;
        uint32_t acc = 0;
        for (int j = 0; j < n; ++j) {
            const float f = src[32 * w + j];
            acc |= (uint32_t)(MODE == 1 ? (f >= 1.0f) : (f > 0.5f)) << j;      // NaN: false in both
        }
        dst[w] = acc;
    }
}

#if defined(__x86_64__)
template <int MODE>
__attribute__((target("avx2"))) void pack_row_f32_avx2(const float* src, int W, uint32_t* dst) {
    const int full = W >> 5;
    const __m256 thr = _mm256_set1_ps(MODE == 1 ? 1.0f : 0.5f);
    for (int w = 0; w < full; ++w) {
        uint32_t acc = 0;
        _mm_prefetch((const char*)(src + 32 * w) + PF, _MM_HINT_T0);
        _mm_prefetch((const char*)(src + 32 * w) + PF + 64, _MM_HINT_T0);
        for (int g = 0; g < 4; ++g) {
            const __m256 v = _mm256_loadu_ps(src + 32 * w + 8 * g);
            // ordered compares: NaN is false, as in the scalar rule
            const __m256 c = MODE == 1 ? _mm256_cmp_ps(v, thr, _CMP_GE_OQ) : _mm256_cmp_ps(v, thr, _CMP_GT_OQ);
            acc |= (uint32_t)_mm256_movemask_ps(c) << (8 * g);
        }
        dst[w] = acc;
    }
    if (W & 31) {
        uint32_t acc = 0;
        for (int x = 32 * full; x < W; ++x) acc |= (uint32_t)(MODE == 1 ? (src[x] >= 1.0f) : (src[x] > 0.5f)) << (x & 31);
        dst[full] = acc;
    }
}
#endif

#if defined(__x86_64__)
template <int MODE>
__attribute__((target("avx512f,avx512bw"))) void pack_row_f32_avx512(const float* src, int W, uint32_t* dst) {
    const int full = W >> 5;
    const __m512 thr = _mm512_set1_ps(MODE == 1 ? 1.0f : 0.5f);
    for (int w = 0; w < full; ++w) {
        _mm_prefetch((const char*)(src + 32 * w) + PF, _MM_HINT_T0);
        _mm_prefetch((const char*)(src + 32 * w) + PF + 64, _MM_HINT_T0);
        const __m512 a = _mm512_loadu_ps(src + 32 * w), b = _mm512_loadu_ps(src + 32 * w + 16);
        // ordered compares: NaN is false, as in the scalar rule
        const uint32_t lo = MODE == 1 ? _mm512_cmp_ps_mask(a, thr, _CMP_GE_OQ) : _mm512_cmp_ps_mask(a, thr, _CMP_GT_OQ);
        const uint32_t hi = MODE == 1 ? _mm512_cmp_ps_mask(b, thr, _CMP_GE_OQ) : _mm512_cmp_ps_mask(b, thr, _CMP_GT_OQ);
        dst[w] = lo | (hi << 16);
    }
    if (W & 31) {
        uint32_t acc = 0;
        for (int x = 32 * full; x < W; ++x) acc |= (uint32_t)(MODE == 1 ? (src[x] >= 1.0f) : (src[x] > 0.5f)) << (x & 31);
        dst[full] = acc;
    }
}
#endif

void pack_rows(const void* masks, int mode, int64_t row0, int64_t row1, int W, uint32_t* packed, int simd) {
    const int wpr = (W + 31) >> 5;
    for (int64_t r = row0; r < row1; ++r) {
        uint32_t* dst = packed + r * wpr;
        if (mode == 0) {
            const uint8_t* src = (const uint8_t*)masks + r * W;
#if defined(__x86_64__)
            if (simd == 2) { pack_row_u8_avx512(src, W, dst); continue; }
            if (simd == 1) { pack_row_u8_avx2(src, W, dst); continue; }
#endif
            pack_row_u8(src, W, dst);
        } else {
            const float* src = (const float*)masks + r * W;
#if defined(__x86_64__)
            if (simd == 2) {
                if (mode == 1) pack_row_f32_avx512<1>(src, W, dst);
                else pack_row_f32_avx512<2>(src, W, dst);
                continue;
            }
            if (simd == 1) {
                if (mode == 1) pack_row_f32_avx2<1>(src, W, dst);
                else pack_row_f32_avx2<2>(src, W, dst);
                continue;
            }
#endif
            if (mode == 1) pack_row_f32<1>(src, W, dst);
            else pack_row_f32<2>(src, W, dst);
        }
    }
#if defined(__x86_64__)
    _mm_sfence();                                       // streaming stores of the AVX-512 path
#endif
}

std::atomic<int> g_simd_cap{2};

}  // namespace

extern "C" int lf_host_simd(int cap) {
    if (cap >= 0) g_simd_cap.store(cap > 2 ? 2 : cap);
    int simd = 0;                                       // 0 scalar, 1 AVX2, 2 AVX-512BW (checked at run time)
#if defined(__x86_64__)
    if (__builtin_cpu_supports("avx2")) simd = 1;
    if (__builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw")) simd = 2;
#endif
    const int c = g_simd_cap.load();
    return simd < c ? simd : c;
}

extern "C" int lf_pack_masks_host(const void* masks, int mode, int64_t m, int H, int W, uint32_t* packed,
                                  int threads) {
    if (m < 0 || H <= 0 || W <= 0 || mode < 0 || mode > 2) return LF_E_ARG;
    if (m == 0) return LF_OK;
    if (!masks || !packed) return LF_E_ARG;
    const int simd = lf_host_simd(-1);
    const int64_t rows = m * H;
    int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 256) nt = 256;
    if ((int64_t)nt > rows / 64 + 1) nt = (int)(rows / 64 + 1);            // not worth a thread below 64 rows
    if (nt == 1) {
        pack_rows(masks, mode, 0, rows, W, packed, simd);
        return LF_OK;
    }
    std::vector<std::thread> pool;
    try {
        for (int t = 0; t < nt; ++t) {
            const int64_t a = rows * t / nt, b = rows * (t + 1) / nt;
            pool.emplace_back(pack_rows, masks, mode, a, b, W, packed, simd);
        }
    } catch (...) {
        for (auto& th : pool) th.join();
        return LF_E_UNSUPPORTED;                                            // could not start the threads
    }
    for (auto& th : pool) th.join();
    return LF_OK;
}
