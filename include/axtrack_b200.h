/* axtrack_b200 -- C ABI of the B200-native AxTrack association hot path.
 *
 * This is the drop-in boundary: a plain C shared library (libaxtrack_b200.so), no torch or
 * Python types in any signature.  Every entry point names the reference interface it replaces
 * (paths relative to the reference checkout, LoaloaF/axtrack @ 47bfa98).
 *
 * Conventions
 *   - all `*_dev` pointers are CUDA device pointers owned by the caller (e.g. torch tensors);
 *     `*_host` pointers are host memory (pinned for best speed); nothing is allocated for the
 *     caller and nothing is kept after the call returns except ctx-owned scratch;
 *   - `stream` is a cudaStream_t passed as void* (NULL = the legacy default stream) and must belong
 *     to the ctx's device; device entry points are asynchronous on that stream.  The ctx owns the
 *     scratch of the calls in flight, so ONE ctx serves ONE stream at a time: issue the calls of a
 *     ctx on one stream (or order them yourself), and create one ctx per thread / per concurrent
 *     stream.  Every entry point selects the ctx's device and restores the calling thread's
 *     current device before it returns.  axb_cost_matrix_batch with two or more pairs forks part of its
 *     work (the all-levels searches of off-mask detections) onto a ctx-owned side stream and joins it back
 *     into `stream` before its last launch: the caller still sees one call ordered on `stream`;
 *   - the device-side status is sticky PER CTX until axb_status() reads it: a flag raised by one
 *     call (e.g. AXB_ERR_CAPACITY of axb_decode_nms) is reported by the next axb_status() or
 *     *_host entry point on that ctx, whichever comes first;
 *   - return value: AXB_OK or a negative AXB_ERR_* code.  Input errors that the reference
 *     raises from inside pyastar2d (out-of-grid start / goal, astar_wrapper.py:39-45) are
 *     detected on the device and reported by axb_status() after the stream has drained;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *     AXB_ERR_CUDA.
 */
#ifndef AXTRACK_B200_H
#define AXTRACK_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AXB_OK                 0
#define AXB_ERR_CUDA          -1   /* CUDA runtime error / no device; see axb_last_error()      */
#define AXB_ERR_OOB_START     -2   /* "Start of (y, x) lies outside grid."  astar_wrapper.py:39-42 */
#define AXB_ERR_OOB_GOAL      -3   /* "Goal of (y, x) lies outside grid."   astar_wrapper.py:43-45 */
#define AXB_ERR_BAD_WEIGHT    -4   /* "Minimum cost to move must be 1"      astar_wrapper.py:35-37 */
#define AXB_ERR_BAD_ARG       -5   /* NULL pointer, negative size, max_px outside [1, AXB_MAX_PX] */
#define AXB_ERR_CAPACITY      -6   /* more detections than `cap` in axb_decode_nms              */
#define AXB_ERR_UNSUPPORTED_WEIGHT -7 /* weight grid holds values other than {1, 65536}        */

/* Largest max_px (the reference's max_px_assoc_dist / max_path_length; 500 in
 * AxonDetections.py:77, an attribute the reference honours at :623-625,736).  Up to 511 the
 * search window lives in shared memory (tile-word kernels); 512..1023 run on the all-levels
 * kernel with its window bitmaps in global memory (several times slower per source). */
#define AXB_MAX_PX 1023

typedef struct axb_ctx axb_ctx;

/* One frame pair = one key f'{name}_t:{t:0>3}-t:{t_bef:0>3}' of astar_dets_paths
 * (AxonDetections.py:563).  Sources are the detections of the predecessor frame t_bef,
 * targets those of frame t, the mask is frame t's (AxonDetections.py:556-566). */
typedef struct axb_pair_desc {
    int32_t mask_idx;   /* which packed mask of `mask_bits_dev` (frame t's)                    */
    int32_t src_off;    /* first source detection in dets_yx_dev (rows of (y, x))             */
    int32_t n_src;      /* n_pred                                                              */
    int32_t dst_off;    /* first target detection in dets_yx_dev                              */
    int32_t n_dst;      /* n_cur                                                               */
    int32_t time_gap;   /* t - t_bef; read by axb_transition_cost_batch only (0 = not given)      */
    int64_t out_off;    /* element offset of this pair's [n_src, n_dst] block in the outputs  */
} axb_pair_desc;

/* ---- lifetime ------------------------------------------------------------------------- */
int  axb_create(int device, axb_ctx** out);
void axb_destroy(axb_ctx* ctx);
const char* axb_last_error(const axb_ctx* ctx);
const char* axb_version(void);
/* number of kernels this ctx has launched since creation (bench.py's gpu_launches) */
int64_t axb_launch_count(const axb_ctx* ctx);

/* ---- mask ------------------------------------------------------------------------------
 * Replaces AxonDetections._get_maskweights (AxonDetections.py:587-598), which expands the
 * boolean mask to float32 weights where(mask==1, 1, 65536).  Here the mask is bit-packed
 * once per frame: bit (x & 31) of word [y][x >> 5] = (mask[y][x] == 1).
 * axb_mask_words(W) = words per packed row. */
int axb_mask_words(int W);
int axb_pack_mask(axb_ctx* ctx, const uint8_t* mask_dev, int n_masks, int H, int W,
                  uint32_t* mask_bits_dev, void* stream);
/* Same from a float32 weight grid as pyastar2d.astar_path receives it
 * (astar_wrapper.py:25-37).  Sets the sticky status AXB_ERR_BAD_WEIGHT if any weight < 1 and
 * AXB_ERR_UNSUPPORTED_WEIGHT if any weight is neither 1 nor 65536. */
int axb_pack_weights(axb_ctx* ctx, const float* weights_dev, int n_masks, int H, int W,
                     uint32_t* mask_bits_dev, void* stream);

/* ---- A* path-length cost matrix --------------------------------------------------------
 * Replaces, for all (source, target) pairs of the given frame pairs, the reference's
 *   AxonDetections._compute_detections_astar_paths     AxonDetections.py:526-585
 *     -> _get_path_between_dets                        AxonDetections.py:600-629
 *     -> utils._compute_astar_path                     utils.py:351-390
 *     -> pyastar2d.astar_path / astar()                astar_wrapper.py:25-54, astar.cpp:46-172
 *   followed by _get_astar_path_distances              AxonDetections.py:717-752
 * Outputs, for pair p, source i, target j at element out_off + i*n_dst + j:
 *   path_len    int32  number of pixels on the A* path (both ends included), or max_px when the
 *               reference leaf is None (AxonDetections.py:736)
 *   unreachable uint8  1 where the reference leaf is None (Euclid distance >= max_px, or the
 *               search hit the max_path_length cut-off, astar.cpp:88-92), else 0
 *   wcost       float32 (optional, may be NULL) cost of the returned path = sum of weights over
 *               path[1:] (astar.cpp:122), +inf where unreachable
 * dets_yx_dev: int32 [n_dets, 2] rows (y, x).  pairs_dev: n_pairs descriptors in DEVICE memory.
 * total_src = sum of n_src over the pairs (host value, sizes the work list); out_capacity = elements
 * each output buffer holds.  The descriptors are validated on the device before any search runs
 * (the pair loop of AxonDetections.py:553-566 cannot index outside its own lists; a C caller can):
 * a pair with a negative count, detections outside [0, n_dets], mask_idx outside [0, n_masks) or
 * an output block outside [0, out_capacity) gets no work and sets AXB_ERR_BAD_ARG in axb_status,
 * as does a total_src smaller than the pairs' sources (then nothing runs). */
/* Optional: what axb_cost_matrix_batch derives from the masks alone (8x4-pixel tile words, mask components for the
 * cover) computed AHEAD of it, e.g. while the host still waits for the detection counts it builds the pair
 * descriptors from (AxonDetections.py:553-566 reads len(dets) per frame).  One shot: the NEXT axb_cost_matrix_batch
 * on this ctx with the same pointer and sizes uses the products and skips that work; any other call in between
 * discards them, and so does axb_prepare_masks(ctx, NULL, ...) (do that when the batch call is not going to
 * happen: the products are keyed by the pointer, not by the mask's contents).  Same stream as that call; the bits
 * must not change in between.  No reference counterpart (the reference has no such stage); results are identical
 * with and without it. */
int axb_prepare_masks(axb_ctx* ctx, const uint32_t* mask_bits_dev, int n_masks, int H, int W, void* stream);

int axb_cost_matrix_batch(axb_ctx* ctx,
                          const uint32_t* mask_bits_dev, int n_masks, int H, int W,
                          const int32_t* dets_yx_dev, int n_dets,
                          const axb_pair_desc* pairs_dev, int n_pairs, int total_src,
                          int max_px,
                          int32_t* path_len_dev, uint8_t* unreachable_dev, float* wcost_dev,
                          int64_t out_capacity, void* stream);

/* Single frame pair, device buffers (a 1-element batch). */
int axb_cost_matrix(axb_ctx* ctx, const uint32_t* mask_bits_dev, int H, int W,
                    const int32_t* src_yx_dev, int n_src, const int32_t* dst_yx_dev, int n_dst,
                    int max_px,
                    int32_t* path_len_dev, uint8_t* unreachable_dev, float* wcost_dev,
                    void* stream);

/* Many frame pairs, HOST buffers: uint8 masks [n_masks, H, W], host detections [n_dets, 2],
 * host descriptors, host outputs; does the host->device copies, the mask packing, the search
 * and the device->host copies on the ctx's own stream and synchronises.  This is the call a
 * non-torch binding of _compute_detections_astar_paths (AxonDetections.py:526-585) makes, and
 * what bench.py times as `e2e`. */
int axb_cost_matrix_batch_host(axb_ctx* ctx, const uint8_t* masks_host, int n_masks, int H, int W,
                               const int32_t* dets_yx_host, int n_dets,
                               const axb_pair_desc* pairs_host, int n_pairs, int max_px,
                               int32_t* path_len_host, uint8_t* unreachable_host, float* wcost_host);

/* Single frame pair, HOST buffers: uint8 mask [H, W], host detections, host outputs; does the
 * host<->device copies and synchronises (a 1-element axb_cost_matrix_batch_host). */
int axb_cost_matrix_host(axb_ctx* ctx, const uint8_t* mask_host, int H, int W,
                         const int32_t* src_yx_host, int n_src,
                         const int32_t* dst_yx_host, int n_dst, int max_px,
                         int32_t* path_len_host, uint8_t* unreachable_host, float* wcost_host);

/* ---- decode + confidence threshold + NMS ------------------------------------------------
 * Replaces, per frame,
 *   AxonDetections._yolo_Y2pandas_det       AxonDetections.py:178-248
 *   Timelapse.stitch_tiles (coordinates)    Timelapse.py:166-197 (+ :141-148)
 *   AxonDetections._non_max_supression      AxonDetections.py:250-278
 * Y_dev      float32 [F, ntiles, Sx, Sy, 3] detector output (conf, x, y) per grid cell
 *            (model.py:119-125); dim 1 of a tile is the x cell, dim 2 the y cell
 * tile_yx_dev int32 [ntiles, 2] (tile_y, tile_x) of each compact tile (Timelapse.py:141-148)
 * Outputs (cap = row pitch; AXB_ERR_CAPACITY is set if a frame has more candidates):
 *   pre_*     the stitched pre-NMS frame detections in the reference's concat order
 *             (tile-major; inside a tile conf ascending): conf [F,cap], frame coords
 *             [F,cap,2] as (y, x) -- the row layout axb_cost_matrix_batch reads --, cell [F,cap] = the
 *             grid cell the candidate came from, tile*Sx*Sy + ix*Sy + iy (the reference names a
 *             tile's detections Axon_### by their rank in this order, AxonDetections.py:222-235)
 *   n_pre     int32 [F]
 *   conf/yx   NMS survivors in conf-descending order ([F,cap], [F,cap,2] rows (y, x));
 *             keep_idx = their pre-NMS indices ("survivor indices")
 *   n_out     int32 [F]
 * Any of the pre_* pointers may be NULL.  The yx output can be passed as dets_yx_dev to
 * axb_cost_matrix_batch unchanged (frame f's detections start at row f*cap). */
int axb_decode_nms(axb_ctx* ctx, const float* Y_dev, const int32_t* tile_yx_dev,
                   int F, int ntiles, int Sx, int Sy, int tilesize, float conf_thr, int nms_dist,
                   int cap,
                   float* pre_conf_dev, int32_t* pre_yx_dev, int32_t* pre_cell_dev, int32_t* n_pre_dev,
                   float* conf_dev, int32_t* yx_dev, int32_t* keep_idx_dev, int32_t* n_out_dev,
                   void* stream);

/* Same with HOST buffers (Y, tile table and every output in host memory; pre_* may be NULL):
 * host->device copy of the head tensor, the fused kernel, device->host copies of the results on
 * the ctx's own stream, then synchronises and returns the sticky status (AXB_ERR_CAPACITY).
 * This is the call a non-torch binding of the per-frame loop body AxonDetections.py:122-128
 * makes for all frames at once. */
int axb_decode_nms_host(axb_ctx* ctx, const float* Y_host, const int32_t* tile_yx_host,
                        int F, int ntiles, int Sx, int Sy, int tilesize, float conf_thr, int nms_dist,
                        int cap,
                        float* pre_conf_host, int32_t* pre_yx_host, int32_t* pre_cell_host, int32_t* n_pre_host,
                        float* conf_host, int32_t* yx_host, int32_t* keep_idx_host, int32_t* n_out_host);

/* ---- path pixels -------------------------------------------------------
 * Replaces utils._compute_astar_path (utils.py:351-390) -> pyastar2d.astar_path for the pairs
 * whose PIXELS are wanted (the reference wraps them as a bool coo_matrix, utils.py:381-382; its
 * own consumers only read getnnz()).  One request = (mask_idx, source y, source x, target y,
 * target x), host int32 [n_req][5].  For every request the all-levels search records the
 * direction each window cell was entered from and walks it back from the target:
 *   path_len_dev  int32 [n_req]: pixels on the path L (start and goal included), 0 = None (the
 *                 reference returns None: no path within max_px / Euclid >= max_px)
 *   path_yx_dev   int32 [n_req][max_px][2]: (row, col) of the L pixels, start -> goal
 * The path is A minimum-cost path under the reference's weights with exactly the reference's
 * length and cost; which of the equal-cost paths is heap-tie dependent in the reference and is
 * not reproduced.  Errors (start / goal outside the grid) arrive through axb_status. */
int axb_trace_paths(axb_ctx* ctx, const uint32_t* mask_bits_dev, int n_masks, int H, int W,
                    const int32_t* req_host, int n_req, int max_px,
                    int32_t* path_yx_dev, int32_t* path_len_dev, void* stream);

/* ---- appearance features ------------------------------------------------
 * Replaces mincostflow_models.feature_model (mincostflow_models.py:30-64): per box (x, y, w, h)
 * of boxes_dev int32 [n][4], the histogram of the image crop image[max(y,0) : max(y0+1, y0+h),
 * max(x,0) : max(x0+1, x0+w)] (clipped at the image border like the numpy slice) over `bins`
 * uniform bins of [lo, hi) as cv2.calcHist computes it, min-max normalised to [0, 1] as
 * cv2.normalize(..., NORM_MINMAX) does (OpenCV 4.x float arithmetic, fused multiply-add).
 * image_dev float32 [H][W] (one frame, Timelapse.get_frame_and_truedets(t)[0][0]);
 * out_dev float32 [n][bins] (the reference's (n, bins, 1) array).  bins <= AXB_MAX_BINS. */
#define AXB_MAX_BINS 1024
int axb_feature_hist(axb_ctx* ctx, const float* image_dev, int H, int W,
                     const int32_t* boxes_dev, int n_boxes, int bins, float lo, float hi,
                     float* out_dev, void* stream);
/* The same for the boxes of a BLOCK of frames in one launch (libmot calls feature_model once per frame,
 * offline_mincostflow.py:130; the frames of a timelapse are known up front): images_dev float32
 * [n_frames][H][W], box_frame_dev int32 [n_boxes] = frame of each box (NULL: all on frame 0). */
int axb_feature_hist_frames(axb_ctx* ctx, const float* images_dev, int n_frames, int H, int W,
                            const int32_t* boxes_dev, const int32_t* box_frame_dev, int n_boxes,
                            int bins, float lo, float hi, float* out_dev, void* stream);

/* ---- visual similarity (the vis_sim term of transition_model) ------------------
 * Replaces the double loop mincostflow_models.py:106-113 over (predecessor, current) detections:
 *   vis[i][j] = nan_to_num(1 - cv2.compareHist(f_src[i], f_dst[j], cv2.HISTCMP_BHATTACHARYYA))
 * feats_*_dev float32 [n][bins] (axb_feature_hist output); vis_dev float64 [n_src][n_dst].  The sums
 * follow OpenCV 4.x's baseline (SSE) order in double, so the result is bit-identical to cv2 4.13.
 * path_len_dev (optional int32 [n_src][n_dst], the A* distances of the same frame pair): entries equal
 * to max_px are skipped and written as 0 -- transition_model overwrites their cost with inf
 * (mincostflow_models.py:104,118), so the similarity never reaches the solver. */
int axb_vis_sim(axb_ctx* ctx, const float* feats_src_dev, int n_src, const float* feats_dst_dev, int n_dst,
                int bins, const int32_t* path_len_dev, int max_px, double* vis_dev, void* stream);

/* ---- tiling front end of the detector --------------------------------------
 * X_dev float32 [T][C][H][W]: the dense timelapse (Timelapse.X.to_dense(), Timelapse.py:426-433:
 * channel 0 image, 1 / 2 positive / negative motion), resident in HBM; tiles are tilesize x tilesize,
 * ytiles = ceil(H / tilesize), xtiles = ceil(W / tilesize), zero padded at the right and bottom
 * border (Timelapse.construct_tiles, Timelapse.py:508-525).
 *
 * axb_tile_nonempty replaces tile_info[..., 0] (Timelapse.py:551: (X[:, :, :, 0] > 0).any(-1).any(-1)):
 * nonempty_dev uint8 [ytiles * xtiles][T].  The tiles the reference keeps are those with a flag set at
 * any t (Timelapse.py:558), in row-major order -- the compact tile list tile_yx of axb_decode_nms.
 *
 * axb_gather_frametiles replaces Timelapse.get_frametiles_stack (Timelapse.py:150-157, __getitem__
 * :111-125) for n_frames frames at once: out_dev float32
 * [n_frames][ntiles][(2 * temporal_context + 1) * chan_n][tilesize][tilesize], plane
 * (dt, ch) = X[frames_dev[b] + dt - temporal_context][chan_lo + ch] cut at tile tile_yx_dev[i] =
 * (tile_y, tile_x).  USE_MOTION_DATA 'exclude' / 'temp_context' = channels (0, 1), 'only' = (1, 2),
 * 'include' = (0, 3).  A frame whose temporal context leaves [0, T) sets AXB_ERR_BAD_ARG in axb_status
 * (its planes are written as zeros). */
int axb_tile_nonempty(axb_ctx* ctx, const float* X_dev, int T, int C, int H, int W, int tilesize,
                      uint8_t* nonempty_dev, void* stream);
/* axb_tile_nonempty without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and result as axb_tile_nonempty; the per-(tile, frame) vote flags
 * live in a static device array of the library instead of the ctx's grown scratch, so
 *   - ytiles * xtiles * T must be <= AXB_NOALLOC_MAX_TILE_FRAMES, else AXB_ERR_BAD_ARG (axb_last_error says so);
 *   - every GPU that loads the library holds that array (AXB_NOALLOC_MAX_TILE_FRAMES * 32 B = 2 MiB) until exit;
 *   - all calls on one device share it (any ctx, any stream): each call zeroes its flags first, and two calls
 *     that overlap on the GPU race -- order them on one stream or with events;
 *   - the call does no cudaMalloc / cudaFree / allocating library routine; the first call on a device looks the
 *     array's device address up (cudaGetSymbolAddress), so make one call per device before a graph capture. */
#define AXB_NOALLOC_MAX_TILE_FRAMES 65536
int axb_tile_nonempty_noalloc(axb_ctx* ctx, const float* X_dev, int T, int C, int H, int W, int tilesize,
                              uint8_t* nonempty_dev, void* stream);
int axb_gather_frametiles(axb_ctx* ctx, const float* X_dev, int T, int C, int H, int W, int tilesize,
                          const int32_t* tile_yx_dev, int ntiles, const int32_t* frames_dev, int n_frames,
                          int temporal_context, int chan_lo, int chan_n, float* out_dev, void* stream);

/* The frame a tracker crops its feature boxes from: AxonDetections.get_frame_and_truedets(t)[0]
 * (AxonDetections.py:355-376), which the reference obtains by tiling the frame and stitching the non-empty
 * tiles back (Timelapse.stitch_tiles image half, Timelapse.py:173-188) = the dense frame with the tiles that
 * are empty at every timepoint zeroed.  tile_keep_dev uint8 [ytiles * xtiles] (1 = kept: non-empty at any t,
 * Timelapse.py:558); frame = raw frame index; out_dev float32 [chan_n][H][W]. */
int axb_stitched_frame(axb_ctx* ctx, const float* X_dev, int T, int C, int H, int W, int tilesize,
                       const uint8_t* tile_keep_dev, int frame, int chan_lo, int chan_n,
                       float* out_dev, void* stream);

/* ---- transition costs ------------------------------------------------------------------
 * Replaces the arithmetic of mincostflow_models.transition_model (mincostflow_models.py:100-119)
 * and the arc cost int(multi_factor * cost) of libmot's graph builder
 * (libmot/libmot/data_association/offline_mincostflow.py:192-198) for vis_sim_weight = 0.
 * cost_tab_dev  float64 [n_gaps][max_px + 1]: cost of path length d at time gap g + 1, evaluated
 *               by the caller with the reference's own numpy expression (so every value is
 *               bit-identical by construction; axtrack_b200/mincostflow_models.py does it)
 * arc_tab_dev   int64 [n_gaps][max_px + 1] or NULL: int(1e5 * cost), or the caller's "no arc"
 *               sentinel where cost > cost_threshold (offline_mincostflow.py:194)
 * pairs_dev[i].time_gap selects the table row; path_len_dev is the output of
 * axb_cost_matrix_batch with the same pair descriptors.  cost_dev float64 / arc_dev int64 (NULL
 * when arc_tab_dev is NULL) use the same [n_src, n_dst] blocks at out_off.  A path length
 * outside [0, max_px] or a time gap outside [1, n_gaps] sets AXB_ERR_BAD_ARG in axb_status. */
int axb_transition_cost_batch(axb_ctx* ctx, const int32_t* path_len_dev,
                              const axb_pair_desc* pairs_dev, int n_pairs, int max_elems_per_pair,
                              const double* cost_tab_dev, const int64_t* arc_tab_dev, int n_gaps, int max_px,
                              double* cost_dev, int64_t* arc_dev, void* stream);

/* ---- status ----------------------------------------------------------------------------
 * Synchronises `stream` and returns the sticky device-side status of the calls issued on this
 * ctx since the last axb_status(): AXB_OK, or the FIRST error in the reference's iteration
 * order (pairs in the order given, sources ascending, targets ascending;
 * AxonDetections.py:553-576).  For AXB_ERR_OOB_*: info[0] = pair index, info[1] = source
 * index, info[2] = target index.  The status is cleared. */
int axb_status(axb_ctx* ctx, void* stream, int64_t info[3]);

#ifdef __cplusplus
}
#endif
#endif /* AXTRACK_B200_H */
