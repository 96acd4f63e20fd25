// the only exchange of the path.  Every rank owns the per-car rows of its
// frames; instead of zero-filling a global table and all-reducing it, the rank's rows are
// STORED straight into every peer's copy of the table over NVLink (peer-mapped symmetric
// memory: one 16-byte store per car and peer), followed by one cross-rank barrier on the
// signal pads.  One launch + one barrier per batch; no staging, no reduction arithmetic.
#include "common.cuh"

namespace {

__global__ void __launch_bounds__(256)
lf_publish_table_kernel(const int32_t* __restrict__ total, const int32_t* __restrict__ best_idx,
                        const int32_t* __restrict__ best_cnt, const int32_t* __restrict__ mask_off,
                        const int32_t* __restrict__ box_off, const long long* __restrict__ agg, int F, int max_cars,
                        int64_t row0, void* const* __restrict__ peers, int world, int rank, int64_t table_off,
                        int64_t agg_off) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (int64_t)F * max_cars) {
        const int f = (int)(i / max_cars), c = (int)(i - (int64_t)f * max_cars);
        const int m0 = mask_off[f], M = mask_off[f + 1] - m0;
        int4 row = make_int4(0, 0, 0, 0);                   // slots past the frame's last car stay empty
        if (c < M) {
            const int g = m0 + c;
            row = make_int4(total[g], best_cnt[g], best_idx[g] + 1, box_off[f + 1] > box_off[f] ? 1 : 0);
        }
        const int64_t slot = (row0 + f) * max_cars + c;
        for (int p = 0; p < world; ++p) {
            const int q = (rank + p) % world;               // start with the own copy, spread the peers
            ((int4*)((char*)peers[q] + table_off))[slot] = row;
        }
    }
    if (agg && blockIdx.x == 0 && threadIdx.x < 8) {
        const long long a = agg[threadIdx.x];
        for (int p = 0; p < world; ++p)
            ((long long*)((char*)peers[p] + agg_off))[rank * 8 + threadIdx.x] = a;
    }
}

}  // namespace

extern "C" int lf_publish_table(lf_ctx* ctx, const int32_t* total, const int32_t* best_idx, const int32_t* best_cnt,
                                const int32_t* mask_off, const int32_t* box_off, const int64_t* agg, int F,
                                int max_cars, int64_t row0, void* const* peers, int world, int rank,
                                int64_t table_off_bytes, int64_t agg_off_bytes, void* stream) {
    if (!ctx || F < 0 || max_cars <= 0 || world <= 0 || rank < 0 || rank >= world || row0 < 0) return LF_E_ARG;
    if (F == 0 && !agg) return LF_OK;
    if (!peers || (F > 0 && (!total || !best_idx || !best_cnt || !mask_off || !box_off))) return LF_E_ARG;
    if ((table_off_bytes & 15) != 0 || (agg_off_bytes & 7) != 0) return LF_E_ARG;
    LF_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t n = (int64_t)F * max_cars;
    const int64_t blocks = n > 0 ? (n + 255) / 256 : 1;
    if (blocks > 0x7FFFFFFF) return LF_E_UNSUPPORTED;
    lf_publish_table_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        total, best_idx, best_cnt, mask_off, box_off, (const long long*)agg, F, max_cars, row0, peers, world, rank,
        table_off_bytes, agg_off_bytes);
    LF_CUDA(ctx, cudaGetLastError());
    return LF_OK;
}
