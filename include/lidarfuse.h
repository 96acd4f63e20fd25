/*
 * lidarfuse.h -- C ABI of the B200-native LiDAR-camera fusion / evaluation path.
 *
 * The reference (KaranSankla/Lidar_Object_Detection) is 100 % Python and has no
 * FFI of its own; its "operator interface" for this path is the set of
 * module-level functions of Coding_testes/cvs_erosion.py, V1..V5 and
 * seg_with_pointcloud.py.  Each entry point below names the reference lines it
 * replaces (paths relative to the reference root).  The ctypes binding a
 * maintainer would add to those scripts is shown in INTEGRATION.md.
 *
 * Conventions
 *   - every pointer marked DEVICE is a device pointer owned by the caller
 *     (e.g. torch.Tensor.data_ptr()); the library owns only the opaque lf_ctx;
 *   - every call returns 0 on success or a negative LF_E_* code; nothing
 *     throws, aborts or prints across this boundary;
 *   - work is enqueued asynchronously on the caller's cudaStream_t (passed as
 *     void*); no call synchronises or allocates, so every call is CUDA-graph
 *     capturable;
 *   - a ctx is bound to one device and is not thread-safe.
 *
 * Layouts
 *   points        float [N][4]  (x, y, z, reflectance), 16-byte aligned
 *   masks (raw)   uint8 [M][H][W], non-zero = set      (or float [M][H][W], >= 1.0 = set)
 *   masks (packed) uint32 [M][H][wpr], wpr = ceil(W/32), bit (x & 31) of word x>>5, LSB = lowest x
 *   boxes         lf_box (below), filled by lf_prepare_boxes or by the host binding
 *   CSR offsets   pt_off int64 [F+1], mask_off / box_off int32 [F+1], cnt_off int64 [F+1]
 *                 with cnt_off[f+1]-cnt_off[f] = M_f * B_f  (inside[] is row-major [M_f][B_f])
 */
#ifndef LIDARFUSE_H
#define LIDARFUSE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LF_VERSION 200

#define LF_OK            0
#define LF_E_ARG        -1   /* bad argument (null pointer, negative size, unsupported shape) */
#define LF_E_CUDA       -2   /* a CUDA call failed; lf_last_cuda_error() has the cudaError_t */
#define LF_E_UNSUPPORTED -3  /* size beyond what the kernels are built for */
#define LF_E_NOCAMERA   -4   /* lf_set_camera has not been called */
#define LF_E_LIMIT      -5   /* beyond the fixed capacity of a no-allocation entry point (lf_depth_maps_noalloc) */

#define LF_MAX_SE 15         /* structuring elements up to 15x15 */
#define LF_MAX_CARS 64       /* masks per frame handled in one pass of lf_fuse_batch */

/* `use_oriented` of lf_fuse_batch / lf_process_batch is a flag word */
#define LF_FUSE_ORIENTED 1   /* oriented slab test (cvs_erosion.py:114-145); clear = axis-aligned (V1:118-139) */
#define LF_FUSE_EXACT    2   /* verification mode: no point is classified in fp32, every point that may pass the
                                gate takes the float64 chain (same results by construction; the tests compare) */

#define LF_NOALLOC_MAX_PIXELS (1 << 20)   /* W * H limit of lf_depth_maps_noalloc: 8 MiB of static device memory */

#define LF_POLICY_NEAREST 0  /* z-buffer: smallest depth per pixel (north-star) */
#define LF_POLICY_LAST    1  /* last point in index order wins (seg_with_pointcloud.py:162-168) */

typedef struct lf_ctx lf_ctx;

/* One oriented box in the form the containment test consumes
 * (cvs_erosion.py:114-138): c0 = corner 0, e1 = c1-c0, e2 = c3-c0, e3 = c4-c0,
 * ee[i] = dot(e_i, e_i); lo/hi = conservative float32 bounds of the region the
 * three closed slabs intersect in, united with [amin, amax] (early-out only). */
typedef struct lf_box {
    double c0[3];
    double e1[3];
    double e2[3];
    double e3[3];
    double ee[3];
    double amin[3];           /* min / max over the 8 corners: the axis-aligned */
    double amax[3];           /* variant point_in_bbox (V1:118-139)             */
    float lo[3];
    float hi[3];
} lf_box;                     /* 192 bytes */

int lf_version(void);
const char* lf_strerror(int code);

int lf_create(lf_ctx** out, int device);
int lf_destroy(lf_ctx* ctx);
int lf_last_cuda_error(const lf_ctx* ctx);

/* Calibration: TrVeloToRect (4x4 row-major, f64) and K (3x4 row-major, f64) as
 * built by cvs_erosion.py:303-314 and CameraPerspective.{K,width,height}. */
int lf_set_camera(lf_ctx* ctx, const double TrVeloToRect[16], const double K[12], int W, int H);

/* A3 + A4  projection and validity gate.
 * Replaces cvs_erosion.py:342-346, 359 (V1:340-344, V4:258-262, V5:467-471,
 * seg_with_pointcloud.py:132-139).  u, v: int64 exactly as numpy's
 * round(...).astype(int); depth f64; pix = v*W+u for points passing
 * 0<=u<W, 0<=v<H, 0<depth<max_depth, else -1.  Any output pointer may be NULL. */
int lf_project(lf_ctx* ctx, const float* pts /*DEVICE [n][4]*/, int64_t n, double max_depth,
               int64_t* u /*DEVICE*/, int64_t* v /*DEVICE*/, double* depth /*DEVICE*/,
               int32_t* pix /*DEVICE*/, void* stream);

/* A5  mask binarisation + erosion + bit-packing.
 * Replaces cvs_erosion.py:96-106 (V3:81-97): cv2.erode with the k x k
 * structuring element `se` (row-major uint8 [k][k], HOST pointer, anchor at the
 * centre, border = +inf), `iterations` times; iterations == 0 packs only (the
 * V2 no-erosion variant, V2:197-200).
 * mode: 0 = uint8 masks, non-zero = set; 1 = float32 masks, >= 1.0 = set (the
 * uint8 truncation of cvs_erosion.py:101 and V2:197); 2 = float32 masks,
 * > 0.5 = set (seg_with_pointcloud.py:167); 3 = masks that are bit-packed
 * already (`masks` = uint32 [m][H][wpr] as `packed` below, e.g. written by
 * lf_pack_masks_host on the sending side: 1/8 of the bytes over PCIe and from
 * HBM); `masks` and `packed` must not be the same buffer then.
 * bbox (optional): per mask the bounding box of the set pixels of the result,
 * encoded {2^20 - x0, 2^20 - y0, x1 + 1, y1 + 1} (all zero = empty mask); it
 * lets lf_fuse_batch skip masks a group of points cannot touch. */
int lf_erode_pack(lf_ctx* ctx, const void* masks /*DEVICE [m][H][W], mode 3: [m][H][wpr] words*/, int mode, int64_t m,
                  int H, int W, const uint8_t* se /*HOST [k][k]*/, int k, int iterations,
                  uint32_t* packed /*DEVICE [m][H][wpr]*/,
                  int32_t* bbox /*DEVICE [m][4], zeroed by the caller, or NULL*/, void* stream);

/* lf_erode_pack with a ROW MAP (same reference lines: cvs_erosion.py:96-106): bit y of
 * row_map[mask * ceil(H/32) + (y >> 5)] = row y of the eroded mask has a set pixel.  Rows OUTSIDE the row range
 * of the mask's bounding box (before its first / after its last row with a set pixel) may be left unwritten
 * in `packed` -- the kernels that hold the eroded strip in shared memory skip every row without a set pixel,
 * and the skipped rows inside the range are zero-filled afterwards: an instance mask covers a tenth of the
 * image rows, so nine tenths of the packed words never travel to HBM.  Unwritten words are unspecified.
 * For consumers that read a packed word only inside the bounding box's rows -- lf_fuse_batch_rows.
 * row_map: DEVICE uint32 [m][ceil(H/32)], zeroed by the call; bbox must be given for the guarantee about
 * the rows inside the range.  row_map == NULL is lf_erode_pack. */
int lf_erode_pack_rows(lf_ctx* ctx, const void* masks /*DEVICE*/, int mode, int64_t m, int H, int W,
                       const uint8_t* se /*HOST [k][k]*/, int k, int iterations, uint32_t* packed /*DEVICE [m][H][wpr]*/,
                       int32_t* bbox /*DEVICE [m][4], zeroed by the caller, or NULL*/,
                       uint32_t* row_map /*DEVICE [m][ceil(H/32)] or NULL*/, void* stream);

/* Host-side transfer encoding of instance masks (plain C++, no device work): binarise
 * with the rule of `mode` (0, 1, 2 as lf_erode_pack) and pack to one bit per pixel,
 * uint32 [m][H][wpr], LSB = lowest x, pad bits zero -- the layout lf_erode_pack (mode 3)
 * and lf_fuse_batch read.  The reference hands masks over as full-resolution arrays
 * (cvs_erosion.py:93 `masks.cpu().numpy()`); packed they are 1/8 (uint8) or 1/32
 * (float32) of the bytes.  `threads` <= 0: one per hardware thread.  Erosion is NOT
 * done here: it stays on the device. */
int lf_pack_masks_host(const void* masks /*HOST [m][H][W]*/, int mode, int64_t m, int H, int W,
                       uint32_t* packed /*HOST [m][H][wpr]*/, int threads);

/* Instruction set lf_pack_masks_host uses: 0 = scalar, 1 = AVX2, 2 = AVX-512BW (detected at
 * run time).  cap >= 0 lowers the ceiling first (the tests run every path); cap < 0 only
 * queries.  Returns the level in use. */
int lf_host_simd(int cap);

/* packed -> uint8 0/1 images (the eroded masks the reference keeps as float32,
 * cvs_erosion.py:103). */
int lf_unpack_masks(lf_ctx* ctx, const uint32_t* packed /*DEVICE*/, int64_t m, int H, int W,
                    uint8_t* out /*DEVICE [m][H][W]*/, void* stream);

/* A7  box preparation on the host (plain C, no device work): corners_velo
 * [n][8][3] f64 -> lf_box[n].  Replaces the per-call edge set-up of
 * cvs_erosion.py:120-123, 129-131. */
int lf_prepare_boxes(const double* corners_velo /*HOST [n][8][3]*/, int64_t n, lf_box* out /*HOST*/);

/* A3 + A4 + A6 + A8 + A9 counters, fused, for a batch of F frames.
 * Replaces the per-frame body cvs_erosion.py:342-372: projection, gate,
 * extract_car_points_by_mask (:148-162), oriented_point_in_bbox (:114-145, or
 * the AABB variant V1:118-139 when use_oriented == 0) and the counting half of
 * calculate_car_point_statistics (:175-195).  Outputs must be zeroed by the
 * caller (lf_zero_counters does it in one launch).
 *   total[mask]        points of the car (gated, in mask)
 *   inside[cnt_off[f] + c*B_f + b]   points of car c inside box b
 *   n_in_box[box]      gated points inside box b (extension metrics)
 *   n_valid[f]         points passing the gate
 *   car_bits           optional DEVICE uint64 [total points]: bit c = point in mask c of its frame
 *                      (0 for gated-out points); may be NULL
 * use_oriented: LF_FUSE_ORIENTED | LF_FUSE_EXACT flags (0 / 1 keep their old meaning).  Any F: batches
 * beyond 65535 frames are issued as several launches.  Camera images up to 65535 x 65535.         */
int lf_fuse_batch(lf_ctx* ctx, const float* pts /*DEVICE*/, const int64_t* pt_off /*DEVICE [F+1]*/,
                  const uint32_t* packed /*DEVICE*/,
                  const int32_t* mask_bbox /*DEVICE [sum M][4] from lf_erode_pack, 16-B aligned, or NULL*/,
                  const int32_t* mask_off /*DEVICE [F+1]*/,
                  const lf_box* boxes /*DEVICE*/, const int32_t* box_off /*DEVICE [F+1]*/,
                  const int64_t* cnt_off /*DEVICE [F+1]*/, int F, int64_t max_points_per_frame,
                  int max_masks_per_frame, int max_boxes_per_frame, double max_depth, int use_oriented,
                  int32_t* total /*DEVICE*/, int32_t* inside /*DEVICE*/, int32_t* n_in_box /*DEVICE*/,
                  int32_t* n_valid /*DEVICE*/, uint64_t* car_bits /*DEVICE or NULL*/, void* stream);

/* lf_fuse_batch for packed masks written by lf_erode_pack_rows (row_map != NULL says so; mask_bbox must be the
 * one of that call): a packed word is read only inside its mask's bounding-box rows (same reference lines:
 * cvs_erosion.py:342-372 without the erosion).  row_map == NULL is lf_fuse_batch. */
int lf_fuse_batch_rows(lf_ctx* ctx, const float* pts /*DEVICE*/, const int64_t* pt_off /*DEVICE [F+1]*/,
                  const uint32_t* packed /*DEVICE*/,
                  const int32_t* mask_bbox /*DEVICE [sum M][4] from lf_erode_pack, 16-B aligned, or NULL*/,
                  const int32_t* mask_off /*DEVICE [F+1]*/,
                  const lf_box* boxes /*DEVICE*/, const int32_t* box_off /*DEVICE [F+1]*/,
                  const int64_t* cnt_off /*DEVICE [F+1]*/, int F, int64_t max_points_per_frame,
                  int max_masks_per_frame, int max_boxes_per_frame, double max_depth, int use_oriented,
                  int32_t* total /*DEVICE*/, int32_t* inside /*DEVICE*/, int32_t* n_in_box /*DEVICE*/,
                  int32_t* n_valid /*DEVICE*/, uint64_t* car_bits /*DEVICE or NULL*/,
                  const uint32_t* row_map /*DEVICE [sum M][ceil(H/32)] from lf_erode_pack_rows, or NULL*/, void* stream);

/* The whole per-batch body of cvs_erosion.process_frames (:342-372) in one call:
 * lf_erode_pack + lf_fuse_batch + lf_car_stats with the same arguments and the same
 * results (three launches on `stream`).  masks == NULL: `packed` and `mask_bbox` are
 * inputs (already packed masks).  best_idx / best_cnt / agg as in lf_car_stats (NULL
 * best_idx skips the decision).  `sched` is reserved and ignored (pass NULL): the
 * persistent erode+fuse kernel it selected in version 100 was slower than the launches. */
int lf_process_batch(lf_ctx* ctx, const float* pts /*DEVICE*/, const int64_t* pt_off /*DEVICE [F+1]*/,
                     const void* masks /*DEVICE [n_masks][H][W] or NULL*/, int mask_mode, int64_t n_masks,
                     const int32_t* mask_off /*DEVICE [F+1]*/, const uint8_t* se /*HOST [k][k]*/, int k,
                     int iterations, uint32_t* packed /*DEVICE*/, int32_t* mask_bbox /*DEVICE [n_masks][4]*/,
                     const lf_box* boxes /*DEVICE*/, const int32_t* box_off /*DEVICE [F+1]*/,
                     const int64_t* cnt_off /*DEVICE [F+1]*/, int F, int64_t max_points_per_frame,
                     int max_masks_per_frame, int max_boxes_per_frame, double max_depth, int use_oriented,
                     int min_points, int32_t* total, int32_t* inside, int32_t* n_in_box, int32_t* n_valid,
                     int32_t* best_idx, int32_t* best_cnt, int64_t* agg, int32_t* sched /*DEVICE or NULL*/,
                     void* stream);

/* A6 stand-alone: bit c of bits[i] = packed mask c is set at (v[i], u[i])
 * (extract_car_points_by_mask, cvs_erosion.py:148-162).  m <= LF_MAX_CARS. */
int lf_mask_gather(lf_ctx* ctx, const uint32_t* packed /*DEVICE [m][H][wpr]*/, int m, int H, int W,
                   const int64_t* u /*DEVICE*/, const int64_t* v /*DEVICE*/, int64_t n,
                   uint64_t* bits /*DEVICE [n]*/, void* stream);

/* A4 stand-alone: valid[i] = 0 <= u < W, 0 <= v < H, 0 < depth < max_depth on already projected arrays
 * (the boolean gate of cvs_erosion.py:359; 30 m in V4:275, V5:508, seg_with_pointcloud.py:154-158). */
int lf_filter_valid(lf_ctx* ctx, const int64_t* u /*DEVICE*/, const int64_t* v /*DEVICE*/, const double* depth /*DEVICE*/,
                    int64_t n, int W, int H, double max_depth, uint8_t* valid /*DEVICE [n]*/, void* stream);

/* A6 compaction: car_points = points_valid[mask_indices] for ALL masks of a frame at once
 * (extract_car_points_by_mask, cvs_erosion.py:156-160), order-preserving, from the membership words of
 * lf_mask_gather.  Two calls around one small read-back:
 *   lf_car_counts   counts[c] = points mask c selects (DEVICE int32 [m]); fills `scratch` (DEVICE int32
 *                   [lf_car_scratch_elems(n, m)], 8-byte aligned) with the offsets lf_car_compact needs
 *   lf_car_compact  out_xyz (DEVICE [sum counts][3], float32 or float64 like xyz) holds car 0's points, then
 *                   car 1's, ... each in point order; out_idx (optional) the point indices. */
int64_t lf_car_scratch_elems(int64_t n, int m);
int lf_car_counts(lf_ctx* ctx, const uint64_t* bits /*DEVICE [n]*/, int64_t n, int m, int32_t* scratch /*DEVICE*/,
                  int32_t* counts /*DEVICE [m]*/, void* stream);
int lf_car_compact(lf_ctx* ctx, const uint64_t* bits /*DEVICE [n]*/, int64_t n, int m, const void* xyz /*DEVICE*/,
                   int is_f64, int stride_elems, const int32_t* scratch /*DEVICE, from lf_car_counts*/,
                   void* out_xyz /*DEVICE*/, int32_t* out_idx /*DEVICE or NULL*/, void* stream);

/* A8 stand-alone: out[i] = point i inside the box (oriented_point_in_bbox,
 * cvs_erosion.py:114-145; point_in_bbox, V1:118-139 when use_oriented == 0).
 * xyz: DEVICE float32 or float64 rows of stride_elems elements (x, y, z first). */
int lf_points_in_box(lf_ctx* ctx, const void* xyz /*DEVICE*/, int is_f64, int64_t n, int stride_elems,
                     const lf_box* box /*DEVICE*/, int use_oriented, uint8_t* out /*DEVICE [n]*/, void* stream);

/* A9 stand-alone counting: inside[set_id[i]*B + b] += point i in box b
 * (the double loop of calculate_car_point_statistics, cvs_erosion.py:175-195). */
int lf_count_in_boxes(lf_ctx* ctx, const void* xyz /*DEVICE*/, int is_f64, int64_t n, int stride_elems,
                      const int32_t* set_id /*DEVICE [n]*/, const lf_box* boxes /*DEVICE [B]*/, int B,
                      int use_oriented, int32_t* inside /*DEVICE, zeroed by the caller*/, void* stream);

/* A9  per-car decision: first strict maximum over boxes starting from 0
 * (cvs_erosion.py:181-195), matched iff best >= min_points (:198).
 *   best_idx[mask] (-1 when no box holds a point), best_cnt[mask]
 *   agg int64[8]: {frames with rows, rows, matched rows, sum total (matched),
 *                  sum inside (matched), sum total (all rows), 0, 0} ADDED to what is there */
int lf_car_stats(lf_ctx* ctx, const int32_t* total, const int32_t* inside, const int32_t* mask_off,
                 const int32_t* box_off, const int64_t* cnt_off, int F, int min_points,
                 int32_t* best_idx /*DEVICE*/, int32_t* best_cnt /*DEVICE*/, int64_t* agg /*DEVICE*/,
                 void* stream);

int lf_zero_counters(lf_ctx* ctx, void* buf /*DEVICE*/, int64_t bytes, void* stream);

/* A10  depth maps.  Replaces seg_with_pointcloud.py:153-170.  One frame.
 * scratch: DEVICE uint64 [H*W].  full: optional DEVICE f64 [H][W] full-frame
 * map (0 where no point).  per_car: optional DEVICE f64 [m][H][W], pixel kept
 * where packed mask c is set.  policy: LF_POLICY_*. */
int lf_depth_maps(lf_ctx* ctx, const float* pts /*DEVICE [n][4]*/, int64_t n, double max_depth,
                  int policy, const uint32_t* packed /*DEVICE [m][H][wpr] or NULL*/, int m,
                  uint64_t* scratch /*DEVICE*/, double* full /*DEVICE or NULL*/,
                  double* per_car /*DEVICE or NULL*/, void* stream);

/* 8(e)  the counter exchange between the ranks of one node, as stores into peer memory.
 * Replaces the table all-reduce of distributed.CounterExchange (zero-filled table, sum over
 * disjoint supports): each rank writes the rows of ITS frames -- {total_points, best_count,
 * best_idx + 1, frame has boxes} as int32[4] per (frame, car slot), slots past a frame's
 * last car zero -- at row0 + f into the table copy of EVERY rank, and its eight aggregates
 * (lf_car_stats `agg`) into slot `rank` of every copy's int64 [world][8] block.
 * peers: DEVICE array [world] of the ranks' buffer base addresses, all mapped into this
 * process (torch symmetric memory `buffer_ptrs_dev`); the table lives table_off_bytes
 * (16-aligned) and the aggregates agg_off_bytes into each buffer.  The caller orders the
 * ranks afterwards (one signal-pad barrier) and double-buffers the table between batches. */
int lf_publish_table(lf_ctx* ctx, const int32_t* total /*DEVICE*/, const int32_t* best_idx /*DEVICE*/,
                     const int32_t* best_cnt /*DEVICE*/, const int32_t* mask_off /*DEVICE [F+1]*/,
                     const int32_t* box_off /*DEVICE [F+1]*/, const int64_t* agg /*DEVICE [8] or NULL*/, int F,
                     int max_cars, int64_t row0, void* const* peers /*DEVICE [world]*/, int world, int rank,
                     int64_t table_off_bytes, int64_t agg_off_bytes, void* stream);

/* lf_depth_maps that allocates nothing and takes no scratch from the caller.  NO COUNTERPART IN THE
 * REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The 64-bit key image
 * lives in a file-scope static device array of LF_NOALLOC_MAX_PIXELS keys (8 MiB, held by every GPU
 * that loads the library until exit); W * H beyond that returns LF_E_LIMIT.  Inputs and outputs are
 * still the caller's; results are those of lf_depth_maps.  Calls no cudaMalloc / cudaFree / allocating
 * routine.  Every call on a device shares the array and starts by clearing it, so calls that overlap
 * on the GPU (two streams, or two contexts on one device) race: serialise them. */
int64_t lf_depth_maps_noalloc_max_pixels(void);
int lf_depth_maps_noalloc(lf_ctx* ctx, const float* pts /*DEVICE [n][4]*/, int64_t n, double max_depth, int policy,
                          const uint32_t* packed /*DEVICE [m][H][wpr] or NULL*/, int m,
                          double* full /*DEVICE [H][W] or NULL*/, double* per_car /*DEVICE [m][H][W] or NULL*/,
                          void* stream);

/* A10 for a batch of frames: the loop of seg_with_pointcloud.py:153-170 over F frames in
 * three launches (key images cleared, one atomic per gated point, resolve).  Frames are
 * CSR-concatenated as in lf_fuse_batch: pt_off DEVICE int64 [F+1], mask_off DEVICE int32
 * [F+1] (masks of frame f = packed[mask_off[f] .. mask_off[f+1])).  scratch: DEVICE uint64
 * [F][H*W].  full: optional DEVICE f64 [F][H][W]; per_car: optional DEVICE f64 [sum M][H][W].
 * Keys of policy `last` are point indices inside the frame, so the result of every frame
 * is the one lf_depth_maps gives for it alone.  Needs an even image width. */
int lf_depth_maps_batch(lf_ctx* ctx, const float* pts /*DEVICE [sum N][4]*/, const int64_t* pt_off /*DEVICE*/, int F,
                        int64_t max_points_per_frame, double max_depth, int policy,
                        const uint32_t* packed /*DEVICE [sum M][H][wpr] or NULL*/, const int32_t* mask_off /*DEVICE or NULL*/,
                        uint64_t* scratch /*DEVICE*/, double* full /*DEVICE or NULL*/, double* per_car /*DEVICE or NULL*/,
                        void* stream);

/* ---- next rows: mask resize to camera resolution ------------------ */

/* cv2.resize(mask.astype(np.uint8), (W, H)) with the default INTER_LINEAR
 * (cvs_erosion.py:153; V4:293; V5:531), bit-exact with OpenCV's uint8 path: copy when the
 * sizes agree, the 2x2 average OpenCV substitutes for an exact 2x downscale, otherwise
 * the 11-bit fixed-point bilinear.  src uint8 [m][h][w] -> dst uint8 [m][H][W]; feed dst
 * to lf_erode_pack (mode 0, iterations 0) to get the "> 0.5" mask the gather reads. */
int lf_resize_masks(lf_ctx* ctx, const uint8_t* src /*DEVICE*/, int64_t m, int h, int w, uint8_t* dst /*DEVICE*/,
                    int H, int W, void* stream);

/* The same resize for masks that are bit-packed already (0/1 images, e.g. the result of
 * lf_erode_pack at the masks' own resolution): src uint32 [m][h][ceil(w/32)] ->
 * dst uint32 [m][H][ceil(W/32)] with dst bit = (cv2.resize(src01, (W, H)) != 0), i.e. exactly what
 * the gather reads after cvs_erosion.py:153-154; bbox (optional, zeroed by the caller) as
 * lf_erode_pack writes it.  1/8 of the bytes of lf_resize_masks and no work for empty regions. */
int lf_resize_pack(lf_ctx* ctx, const uint32_t* src /*DEVICE*/, int64_t m, int h, int w, uint32_t* dst /*DEVICE*/,
                   int32_t* bbox /*DEVICE [m][4] or NULL*/, int H, int W, void* stream);

/* ---- next rows: box projection and 2D matching matrices ---------- */

#define LF_VIEW_VALID           0   /* secondtest.py:350 'valid' */
#define LF_VIEW_ALL_BEHIND      1   /* :304-305 no corner inside the depth range */
#define LF_VIEW_NO_INTERSECTION 2   /* :326-334 */
#define LF_VIEW_TOO_SMALL       3   /* :337-348 projected area < 100 px^2 */

/* What the three reference filters derive from camera.cam2image(corners_cam0.T). */
typedef struct lf_box2d {
    int64_t x0, y0, x1, y1;     /* V5 project_3d_bbox_to_2d: min / max pixel over corners with depth > 0 */
    double avg_depth;           /* np.mean(depth[depth > 0]) (V5:243); 0 when n_front == 0 */
    double range_avg_depth;     /* np.mean(depth[in range]) (secondtest.py:354) */
    int32_t n_front;            /* corners with depth > 0; 0 = project_3d_bbox_to_2d returns None */
    int32_t n_visible;          /* corners with depth > 0.1 inside the image: filter_visible_bboxes keeps >= 2 */
    int32_t n_in_range;         /* secondtest corners_with_valid_depth */
    int32_t corners_in_view;    /* secondtest corners_in_view */
    int32_t view_code;          /* LF_VIEW_* : is_bbox_in_camera_view keeps LF_VIEW_VALID */
    int32_t pad_;
} lf_box2d;                     /* 72 bytes */

/* Projects the 8 cam0 corners of n boxes with K (no rectification, as the reference does:
 * cvs_erosion.py:64, V5:222, secondtest.py:298) and evaluates filter_visible_bboxes
 * (cvs_erosion.py:55-74), project_3d_bbox_to_2d (V5:215-252) and is_bbox_in_camera_view
 * (secondtest.py:277-359; depth range and min_points_in_view as there).  u, v, depth
 * (optional, all or none): the raw per-corner projections, int64 / f64 [n][8]. */
int lf_project_boxes(lf_ctx* ctx, const double* corners_cam0 /*DEVICE [n][8][3]*/, int64_t n, double depth_min,
                     double depth_max, int min_points_in_view, lf_box2d* out /*DEVICE [n]*/,
                     int64_t* u /*DEVICE or NULL*/, int64_t* v /*DEVICE or NULL*/, double* depth /*DEVICE or NULL*/,
                     void* stream);

/* Per frame f the D_f x B_f matrices of calculate_iou_2d (V4:118-137) and
 * calculate_matching_score (V5:277-304; weights as there: 0.5, 0.3, 0.2) between the 2D
 * detections [x1,y1,x2,y2] (f64) and the projected rectangles.  Boxes with n_front == 0
 * get iou 0 and score -inf.  out_off[f] = start of frame f's row-major [D_f][B_f] block. */
int lf_match_matrices(lf_ctx* ctx, const double* dets /*DEVICE [sum D][4]*/, const int32_t* det_off /*DEVICE [F+1]*/,
                      const lf_box2d* rects /*DEVICE [sum B]*/, const int32_t* box_off /*DEVICE [F+1]*/,
                      const int64_t* out_off /*DEVICE [F+1]*/, int F, int max_pairs_per_frame, double w_iou,
                      double w_center, double w_size, double* iou /*DEVICE*/, double* score /*DEVICE*/,
                      void* stream);

/* ---- next rows: headless colour buffers --------------------------- */

/* Per-point RGB of the V4 / V5 coloured clouds (V4:290-307; V5:524-550): colors_rgb[c]
 * (f64 [m][3], already BGR->RGB and / 255) of the last mask whose bit is set in
 * car_bits[i] (from lf_mask_gather / lf_fuse_batch), bg_rgb (HOST [3]) otherwise.
 * assigned (optional) = the scripts' bg_assigned flag.
 * policy: 0 = the LAST mask containing the point wins (V4 / V5 / vis.py:175-205 build one cloud
 * per mask and draw them in mask order); 1 = the FIRST one wins (the per-point loop with
 * `break` of Same_color.py:112-127). */
int lf_point_colors(lf_ctx* ctx, const uint64_t* car_bits /*DEVICE [n]*/, int64_t n,
                    const double* colors_rgb /*DEVICE [m][3]*/, int m, const double bg_rgb[3] /*HOST*/,
                    double* rgb /*DEVICE [n][3]*/, uint8_t* assigned /*DEVICE [n] or NULL*/, int policy,
                    void* stream);

/* Depth map overlay of seg_with_pointcloud.py:173-180: where depth_map > 0 the pixel is
 * lut[int(depth / *depth_max * lut_n)] (matplotlib Colormap.__call__), elsewhere
 * image_rgb / 255; then np.uint8(x * 255) and RGB -> BGR.  depth_max: DEVICE pointer to
 * np.max(depthMap). */
int lf_depth_overlay(lf_ctx* ctx, const double* depth_map /*DEVICE [H][W]*/, const double* depth_max /*DEVICE [1]*/,
                     const uint8_t* image_rgb /*DEVICE [H][W][3]*/, const double* lut_rgb /*DEVICE [lut_n][3]*/,
                     int lut_n, int H, int W, uint8_t* out_bgr /*DEVICE [H][W][3]*/, void* stream);

/* The same for n depth maps in one launch (the loop over cars of seg_with_pointcloud.py:173-180,
 * or over frames): depth_maps [n][H][W], depth_max [n]; image_stride = bytes between the images
 * of consecutive maps, 0 = every map is laid over the same image. */
int lf_depth_overlay_batch(lf_ctx* ctx, const double* depth_maps /*DEVICE [n][H][W]*/,
                           const double* depth_max /*DEVICE [n]*/, const uint8_t* image_rgb /*DEVICE*/,
                           int64_t image_stride, const double* lut_rgb /*DEVICE [lut_n][3]*/, int lut_n, int n,
                           int H, int W, uint8_t* out_bgr /*DEVICE [n][H][W][3]*/, void* stream);

/* ---- next rows: I/O formats ----------------------------------------- */

/* Text of n master_car_statistics.csv rows (HOST arrays, plain C, no device work),
 * byte-identical with append_to_master_csv (cvs_erosion.py:232-265): percentages are
 * (inside / total) * 100 and its complement, rounded like Python's round(x, 2) and
 * printed like repr(); is_matched = matched_bbox_id >= 0; `timestamp` is written as
 * given.  with_header: start with the column line.  Returns the number of bytes the text
 * needs (negative LF_E_*); it is written to `out` only if it fits `cap` (call with
 * out = NULL to size the buffer). */
int64_t lf_format_csv_rows(const int64_t* frame, const int32_t* car_id, const int32_t* matched_bbox_id,
                           const int64_t* total, const int64_t* inside, int64_t n, const char* timestamp,
                           int with_header, char* out, int64_t cap);

/* Synthetic HDL-64-shaped workload (bench / tests only).
 * cars: DEVICE f64 [F][ncars][8] = {cx, cy, cz, yaw, l, w, h, pad} in the velodyne frame. */
int lf_synth_points(lf_ctx* ctx, const double* cars /*DEVICE*/, int ncars, int F, int rings, int azimuths,
                    uint64_t seed, float* pts /*DEVICE [F][rings*azimuths][4]*/, void* stream);
/* masks rendered by casting the pixel ray against each (inflated) car box;
 * Tinv = inverse of TrVeloToRect (4x4 row-major f64, HOST). */
int lf_synth_masks(lf_ctx* ctx, const double* cars /*DEVICE*/, int ncars, int F, const double Tinv[16],
                   double inflate, uint8_t* masks /*DEVICE [F][ncars][H][W]*/, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* LIDARFUSE_H */
