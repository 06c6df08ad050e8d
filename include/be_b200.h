/*
 * be_b200.h — C-ABI of the B200-native tiled bioimage inference kernels.
 *
 * This is the drop-in boundary: every entry point is `extern "C"`, takes plain device
 * pointers + sizes + a cudaStream_t (passed as void*), returns 0 on success or a negative BE_ERR_* code,
 * and never allocates device memory — scratch is a caller-provided workspace whose size comes from the
 * matching *_workspace_bytes() query.  All pointers are DEVICE pointers unless a comment says "host".
 *
 * The reference (aicell-lab/bioengine-worker) is pure Python; the arithmetic each entry point replaces
 * lives in its apps/ directory or in the third-party packages those apps pin.  Every declaration cites
 * the reference call site (file:line under the reference tree) it stands in for.
 *
 * Layout conventions: images are (N, C, Z, Y, X) C-contiguous fp32 (2-D images have Z = 1); tile tensors
 * are (N*T, C, Bz, By, Bx) with tile index t = ((n*Tz + kz)*Ty + ky)*Tx + kx — i.e. exactly the batched
 * NCHW / NCDHW tensor the torch UNet forward consumes; masks are uint8, labels int32.
 */
#ifndef BE_B200_H_
#define BE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BE_OK 0
#define BE_ERR_INVALID_ARG (-1) /* bad size / null pointer / unsupported parameter */
#define BE_ERR_CUDA (-2)        /* a CUDA runtime call or launch failed; see be_last_error_string() */
#define BE_ERR_TOO_LARGE (-3)   /* size exceeds what the entry point supports */

typedef void* be_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------------------------ */

/* ABI version (major*1000 + minor). */
int be_version(void);
/* Last error message of the calling thread ("" if none). Host pointer, valid until the next failing call. */
const char* be_last_error_string(void);

/* ---- pre-processing statistics ------------------------------------------------------------------ */

/* Per-segment mean / population variance, fp64 accumulation in one pass.
 * Replaces bioimageio.core MeanVarStdCalculator used by `zero_mean_unit_variance`
 * (pinned at apps/model-runner/runtime_deployment.py:17, invoked through :278/:280).
 *   x        [n_seg * seg_len] fp32, segment s = x[s*seg_len .. (s+1)*seg_len)
 *   stats    [n_seg * 2] fp64 out: {mean, var(ddof=0)}
 *   norm     [n_seg * 2] fp64 out or NULL: {sub = mean, div = sqrt(var) + eps} for be_normalize / be_tile_gather
 *   ws       be_meanvar_workspace_bytes(n_seg, seg_len) bytes */
size_t be_meanvar_workspace_bytes(int64_t n_seg, int64_t seg_len);
int be_meanvar_f32(const float* x, int64_t n_seg, int64_t seg_len, double eps, double* stats, double* norm,
                   void* ws, be_stream_t stream);

/* Exact per-segment quantiles, numpy `nanquantile(..., method="linear")` semantics (NaNs dropped, fp64
 * virtual index q*(n-1), numpy's two-sided lerp).  Segments of >= 2^18 elements: one streaming pass that brackets
 * each wanted rank between two pivots taken from a sorted 2048-element sample, then an 11/11/10-bit radix select on
 * the ~2 % of elements inside the bracket (on the whole segment if the bracket missed — the result is exact either
 * way); smaller segments: the radix select directly, on the order-preserving uint32 key of fp32.  Replaces bioimageio.core SamplePercentilesCalculator / `scale_range`
 * (runtime_deployment.py:17,:278; Appendix A.2/A.3) and `np.percentile(image_np, [1, 99])`
 * (apps/fibsem-mito-analysis/analysis_deployment.py:241).
 *   q        [n_q] fp64 in [0,1], DEVICE pointer, 1 <= n_q <= 2
 *   quant    [n_seg * n_q] fp64 out
 *   norm     [n_seg * 2] fp64 out or NULL (requires n_q == 2): {sub = q0, div = q1 - q0 + eps}
 *   ws       be_quantile_workspace_bytes(n_seg, seg_len) bytes */
size_t be_quantile_workspace_bytes(int64_t n_seg, int64_t seg_len);
int be_quantile_f32(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                    double* quant, double* norm, void* ws, be_stream_t stream);

/* be_quantile_f32 for callers that cannot provide a workspace because they cannot allocate device memory at all:
 * e.g. a live acquisition loop that must keep clear of cudaMalloc/cudaFree (both stall every stream on the device),
 * or a host runtime that owns all device memory through its own pool and gives a plugin only the input and output
 * pointers.  Same arguments and results as be_quantile_f32 minus `ws`: the scratch is a fixed 336 KiB array inside
 * the library, one copy per GPU for as long as the library stays loaded there.  No counterpart in the reference
 * (it never manages device memory itself); this entry point exists only for such embedders.
 *   - at most 8 segments per call (n_seg > 8 → BE_ERR_TOO_LARGE); seg_len as for be_quantile_f32;
 *   - never calls cudaMalloc, cudaFree or any allocating routine;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress): make that first
 *     call with the device current and before starting any stream capture on it;
 *   - all calls on one GPU share the array: calls whose GPU work can overlap (different streams, different host
 *     threads) race on it — issue them on one stream or order them with events. */
int be_quantile_f32_noalloc(const float* x, int64_t n_seg, int64_t seg_len, const double* q, int n_q, double eps,
                            double* quant, double* norm, be_stream_t stream);

/* ---- elementwise -------------------------------------------------------------------------------- */

/* y = (float)(((double)x - norm[2s]) / norm[2s+1]), optionally clipped to [clip_lo, clip_hi].
 * The apply step of `scale_range` / `zero_mean_unit_variance` / `fixed_zero_mean_unit_variance`
 * (Appendix A.2) and of analysis_deployment.py:242 (clip 0..1).  x and y may alias. */
int be_normalize_f32(const float* x, float* y, int64_t n_seg, int64_t seg_len, const double* norm, int clip,
                     float clip_lo, float clip_hi, be_stream_t stream);

#define BE_OP_CLIP 0         /* y = min(max(x, a), b)              bioimageio `clip`          */
#define BE_OP_SCALE_LINEAR 1 /* y = x*a + b (two fp32 roundings)   bioimageio `scale_linear`  */
#define BE_OP_SIGMOID 2      /* y = 1/(1+exp(-x)), fp32 semantics  bioimageio `sigmoid`       */
int be_elementwise_f32(const float* x, float* y, int64_t n, int op, float a, float b, be_stream_t stream);

/* y = x > thr (strict), 0/1 uint8.  bioimageio `binarize` (Appendix A.2); `prob > 0.5`
 * at analysis_deployment.py:166. */
int be_binarize_f32_u8(const float* x, uint8_t* y, int64_t n, float thr, be_stream_t stream);

/* y = x * gain + offset in fp32 with per-segment constants on the device: offset_gain[2*seg] = offset, [2*seg + 1] = gain
 * (fp64, converted to fp32) — the second half of bioimageio `scale_mean_variance`
 * ((x - mean) / (std + eps) * (ref_std + eps) + ref_mean, Appendix A.2), fed with the `norm` output be_meanvar_f32 wrote
 * for the reference tensor, so the op needs no host round trip. */
int be_scale_linear_seg_f32(const float* x, float* y, int64_t n_seg, int64_t seg_len, const double* offset_gain,
                            be_stream_t stream);

/* y = (float)x for the integer / half / double element types a microscope or a caller hands over — the leading
 * `EnsureDtype(float32)` bioimageio.core inserts before the pre-processing (Appendix A.2) and
 * `np.array(image, dtype=np.float32)` of analysis_deployment.py:236, on the device, so that a uint8 / uint16 frame crosses
 * PCIe at its own size.  Rounding as numpy's astype (to nearest even). */
#define BE_DTYPE_U8 0
#define BE_DTYPE_I8 1
#define BE_DTYPE_U16 2
#define BE_DTYPE_I16 3
#define BE_DTYPE_I32 4
#define BE_DTYPE_U32 5
#define BE_DTYPE_F16 6
#define BE_DTYPE_F64 7
int be_cast_f32(const void* x, int dtype, float* y, int64_t n, be_stream_t stream);

/* ---- halo tiles --------------------------------------------------------------------------------- */

/* Gather halo tiles (any padding rule) into the batched tile tensor, fused with the normalise-apply so
 * the image is read once.  Source coordinates come from three small per-axis tables built by the host
 * (zsrc[Tz*Bz], ysrc[Ty*By], xsrc[Tx*Bx]: image coordinate feeding in-tile offset i of tile k along that
 * axis), which expresses both the bioimageio.core convention (halo inside the block, last block shifted,
 * reflect pad; Appendix A.4, `predict_sample_with_blocking` at runtime_deployment.py:278) and the fibsem
 * convention (stride tiles, bottom/right reflect pad; analysis_deployment.py:132-138).
 *   norm_mode 0: copy only; 1: norm[2*(n*C+c)]; 2: norm[2*n] (statistics joint over channels)
 *   clip      as be_normalize_f32
 * `xsrc` and `tiles` must be 16-byte aligned (BE_ERR_INVALID_ARG otherwise); `img` may have any float alignment. */
int be_tile_gather_f32(const float* img, float* tiles, int64_t N, int64_t C, int64_t Z, int64_t Y, int64_t X,
                       int Tz, int Ty, int Tx, int Bz, int By, int Bx, const int32_t* zsrc, const int32_t* ysrc,
                       const int32_t* xsrc, const double* norm, int norm_mode, int clip, float clip_lo,
                       float clip_hi, be_stream_t stream);

#define BE_EPI_NONE 0
#define BE_EPI_SIGMOID 1

/* Inner-region reassembly (bioimageio `Sample.from_blocks`: out[inner_slice] = block[local_inner],
 * last writer wins; Appendix A.4, runtime_deployment.py:278).  Per-axis tables zsel[Z], ysel[Y], xsel[X]
 * hold k*B + i of the winning block for every output coordinate.  Optional fused epilogue: sigmoid on the
 * stitched value, and/or a mask = value > thr (the `sigmoid` / `binarize` post-processing ops) as uint8 (`mask`) and / or
 * bit-packed (`bits`: per (n, c, z, y) row ceil(X / 64) 64-bit words, lowest bit = first pixel, tail bits zero — the
 * input of be_ccl_bits_i32).  Any of out / mask / bits may be NULL (not all). */
int be_stitch_crop_f32(const float* tiles, float* out, uint8_t* mask, uint64_t* bits, int64_t N, int64_t C, int64_t Z,
                       int64_t Y, int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx, const int32_t* zsel,
                       const int32_t* ysel, const int32_t* xsel, int epilogue, float thr, be_stream_t stream);

/* Gaussian-weighted overlap blend, gather form of analysis_deployment.py:125-156: tiles sit at
 * k*stride, window w = (float)(wz[iz]*wy[iy]*wx[ix]) from host-computed fp64 1-D tables
 * (exp(-2*linspace(-1,1,B)^2); wz = {1.0} for 2-D), acc += (double)(tile*w), wacc += (double)w in
 * ascending tile order, value = (float)(acc/wacc).  Epilogue / mask / bits as be_stitch_crop_f32.
 * `wx` must be 16-byte aligned (BE_ERR_INVALID_ARG otherwise).
 * `window` (optional, may be NULL): the reference's fp32 window table `weight_win` (analysis_deployment.py:125-127) as
 * written by be_blend_window_f32 for the same wz / wy / wx.  With it a 2-D call that wants only the mask decides each
 * pixel from fp32 sums and an error bound, and redoes the few pixels next to the threshold exactly; results are the
 * same with and without it. */
int be_stitch_blend_f32(const float* tiles, float* out, uint8_t* mask, uint64_t* bits, int64_t N, int64_t C, int64_t Z, int64_t Y,
                        int64_t X, int Tz, int Ty, int Tx, int Bz, int By, int Bx, int Sz, int Sy, int Sx,
                        const double* wz, const double* wy, const double* wx, const float* window, int epilogue,
                        float thr, be_stream_t stream);

/* window[iy*Bx + ix] = (float)((wz ? wz[0] : 1.0) * wy[iy] * wx[ix]) — np.outer(...).astype(np.float32) of
 * analysis_deployment.py:127, the table the reference multiplies every tile with.  Written once per tile geometry
 * (By*Bx floats, 16-byte aligned) and handed to be_stitch_blend_f32.  Reads wz[0] back to the host (synchronises). */
int be_blend_window_f32(const double* wz, const double* wy, const double* wx, int By, int Bx, float* window,
                        be_stream_t stream);

/* ---- connected components ----------------------------------------------------------------------- */

/* Connected-component labelling of N independent (Z,Y,X) uint8 masks (non-zero = foreground): background 0,
 * components numbered 1..n in raster order of their first voxel — the numbering of
 * `skimage.measure.label` / `scipy.ndimage.label` (analysis_deployment.py:175; Appendix C).
 * connectivity = maximum number of orthogonal steps to a neighbour (skimage's meaning):
 * 2-D (Z == 1): 1 → 4-conn, 2 → 8-conn; 3-D: 1 → 6, 2 → 18, 3 → 26.
 *   labels    [N*Z*Y*X] int32 out      n_labels [N] int32 out (device)
 *   ws        be_ccl_workspace_bytes(N, Z, Y, X) bytes */
size_t be_ccl_workspace_bytes(int64_t N, int64_t Z, int64_t Y, int64_t X);
int be_ccl_u8_i32(const uint8_t* mask, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                  int connectivity, int32_t* n_labels, void* ws, be_stream_t stream);

/* The same labelling from a bit-packed mask: bits[((n*Z + z)*Y + y)*W + k] holds pixels 64k .. 64k+63 of row (n, z, y),
 * lowest bit first, W = ceil(X / 64) words per row (bits past the end of a row are ignored) — the layout
 * be_stitch_blend_f32 / be_stitch_crop_f32 write when asked for a packed mask, so `binarize` → `label`
 * (analysis_deployment.py:166,175) moves one bit per pixel between the two kernels.  8-byte aligned. */
int be_ccl_bits_i32(const uint64_t* bits, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                    int connectivity, int32_t* n_labels, void* ws, be_stream_t stream);

/* The same labelling straight from fp32 values: foreground = value > thr after the optional sigmoid (epilogue as in
 * be_stitch_crop_f32) — bioimageio `sigmoid` → `binarize` → label of an untiled prediction
 * (apps/model-runner/runtime_deployment.py:280 followed by the label call sites above) without writing a mask at all. */
int be_ccl_f32_i32(const float* x, int epilogue, float thr, int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X,
                   int connectivity, int32_t* n_labels, void* ws, be_stream_t stream);

/* ---- screening metric --------------------------------------------------------------------------- */

/* Sparse label-pair overlap histogram of N (true, pred) label-map pairs of n_px pixels each — the
 * `_label_overlap` / `_intersection_over_union` step of cellpose.metrics.average_precision
 * (apps/cellpose-finetuning/main.py:2002; Appendix B.1).  Per image an open-addressing table of
 * `capacity` (power of two) slots holding the pairs with BOTH labels > 0 (the only ones that can match):
 * keys[slot] = (true << 32) | pred, or ~0 if empty; counts[slot] = #pixels.  The per-label pixel counts (row / column
 * sums of the overlap matrix, overlaps with the background included; index 0 = the background's own row / column sum)
 * and the max labels are filled as well.  All outputs are (re)initialised by the call.
 *   row_len   pixels per image row (n_px % row_len == 0): work is cut into 128 x 128 px tiles; 0 = unknown
 *   keys      [N * capacity] uint64 out     counts [N * capacity] uint32 out
 *   area_true [N * cap_true], area_pred [N * cap_pred] int32 out (index = label)
 *   max_lbl   [N * 2] int32 out {max true, max pred}
 *   status    [N] int32 out: 0 ok, 1 table full (retry with a larger capacity), 2 label negative or >= cap
 *   label_bounds  NULL, or [N * 2] int32 on the device: per image an upper bound of the true / pred labels (e.g. the
 *                 n_labels be_ccl_u8_i32 wrote).  Then only the part of the tables those bounds need is cleared and used
 *                 (pair table: the power of two >= 4 (bt + bp) + 64, at most `capacity`); a label above its bound gives
 *                 status 2.  be_match_counts must be given the same bounds. */
int be_label_overlap_i32(const int32_t* lab_true, const int32_t* lab_pred, int64_t N, int64_t n_px, int64_t row_len,
                         uint64_t* keys, uint32_t* counts, int64_t capacity, int32_t* area_true, int64_t cap_true,
                         int32_t* area_pred, int64_t cap_pred, int32_t* max_lbl, int32_t* status,
                         const int32_t* label_bounds, be_stream_t stream);

/* Threshold matching on the sparse table: for every image and threshold, the number of (true, pred) pairs
 * (both > 0) with (double)ov / (double)(a_t + a_p - ov) >= thr, and `extra` = the number of times a label
 * received a second (third, ...) such partner.  tp = pairs - extra is exact for thr >= 0.5 (Appendix B.2/B.3;
 * replaces the Hungarian assignment in cellpose `_true_positive`); for thr < 0.5 a non-zero `extra` means the
 * caller must solve the assignment for that image.
 *   thr        [n_thr] fp64, device          counts_out [N * n_thr * 2] int64 out {pairs, extra}
 *   ws         be_match_workspace_bytes(N, cap_true, cap_pred, n_thr) bytes
 *   label_bounds  as given to be_label_overlap_i32 (NULL if none) */
size_t be_match_workspace_bytes(int64_t N, int64_t cap_true, int64_t cap_pred, int n_thr);
int be_match_counts(const uint64_t* keys, const uint32_t* counts, int64_t N, int64_t capacity,
                    const int32_t* area_true, int64_t cap_true, const int32_t* area_pred, int64_t cap_pred,
                    const double* thr, int n_thr, int64_t* counts_out, void* ws, const int32_t* label_bounds,
                    be_stream_t stream);

/* Rows of the screening table for one batch of images and one model, written on the device: for image i of the batch,
 * table[rows[i]*(n_thr+2) + k] = tp at threshold k (= pairs - extra of be_match_counts), then n_true, n_pred
 * (= max labels, what cellpose uses for fp = n_pred - tp, fn = n_true - tp) — the integers the ranks all-reduce before
 * `ap = tp/(tp+fp+fn)` and `np.nanmean` (apps/cellpose-finetuning/main.py:2002-2008).  `flags` (2 int32, zeroed by the
 * caller once per shard, sticky): flags[0] = largest status word seen (1 pair table full, 2 label out of range),
 * flags[1] = 1 if a threshold below 0.5 met a label with several partners (that image needs the host assignment).
 *   counts [N*n_thr*2] int64, max_lbl [N*2], status [N] as written by be_match_counts / be_label_overlap_i32
 *   rows   [N] int64 image index of each batch entry      table: the model's [n_images][n_thr+2] int64 block */
int be_screen_rows_i64(const int64_t* counts, const int32_t* max_lbl, const int32_t* status, int64_t N, int n_thr,
                       const double* thr, const int64_t* rows, int64_t* table, int32_t* flags, be_stream_t stream);

/* Pixel-level binary confusion counts of a prediction against a target — `_binary_threshold` +
 * `_compute_binary_metrics` of apps/cellpose-finetuning/main.py:1225-1270.  adaptive != 0 picks each threshold on the
 * device the way the reference does (0.5 if that tensor's values lie within [0, 1], else 0.0); otherwise thr_pred /
 * thr_target are used.  foreground = value > threshold.
 *   minmax [4] fp32 out {pred min, pred max, target min, target max}     counts [4] int64 out {tp, fp, fn, tn} */
int be_binary_confusion_f32(const float* pred, const float* target, int64_t n, int adaptive, float thr_pred,
                            float thr_target, float* minmax, int64_t* counts, be_stream_t stream);

/* Per-label integer raw moments of N label maps (Z,Y,X): for label l in 1..cap-1,
 * sums[((n*cap + l)*10 + j)] = {area, Σz, Σy, Σx, Σzz, Σyy, Σxx, Σzy, Σzx, Σyx} (int64, exact) and
 * bbox[((n*cap + l)*6 + j)] = {zmin, ymin, xmin, zmax, ymax, xmax} (int32, inclusive).  The integer core of
 * skimage `regionprops` area / centroid / inertia-tensor properties (analysis_deployment.py:258-267).
 * sums/bbox are (re)initialised by the call. */
int be_label_stats_i32(const int32_t* labels, int64_t N, int64_t Z, int64_t Y, int64_t X, int64_t cap,
                       int64_t* sums, int32_t* bbox, int32_t* status, be_stream_t stream);

/* ---- mask clean-up and distance transform (first part) ------------------------ */

/* Pixel count of every label: area[n*cap + l] for l in 0..cap-1 (area[.., 0] stays 0).  With be_ccl_u8_i32 at
 * connectivity 1 and be_keep_large_u8 this is `skimage.morphology.remove_small_objects(mask, min_size)`
 * (apps/fibsem-mito-analysis/analysis_deployment.py:166): components with fewer than min_size pixels are dropped.
 *   status [N] int32 out: 0 ok, 2 a label >= cap was met */
int be_label_area_i32(const int32_t* labels, int64_t N, int64_t n_px, int64_t cap, int32_t* area, int32_t* status,
                      be_stream_t stream);
int be_keep_large_u8(const int32_t* labels, const int32_t* area, int64_t N, int64_t n_px, int64_t cap, int32_t min_size,
                     uint8_t* out, be_stream_t stream);

/* `scipy.ndimage.binary_closing(mask, structure)` for N independent (Y, X) masks and a structuring element given by
 * its rows: half_widths[dy + radius] = half width of row dy (-1: empty row), symmetric in dy, HOST pointer —
 * e.g. skimage `disk(4)` = {0, 2, 3, 3, 4, 3, 3, 2, 0} (analysis_deployment.py:169).  Dilation then erosion, both with
 * scipy's border_value = 0.  mask and out may alias.
 *   ws   be_binary_closing_workspace_bytes(N, Y, X) bytes */
size_t be_binary_closing_workspace_bytes(int64_t N, int64_t Y, int64_t X);
int be_binary_closing_u8(const uint8_t* mask, uint8_t* out, int64_t N, int64_t Y, int64_t X, const int32_t* half_widths,
                         int radius, void* ws, be_stream_t stream);

/* Exact Euclidean distance of every pixel to the nearest zero pixel, fp64 — `scipy.ndimage.distance_transform_edt`
 * (analysis_deployment.py:170) for N independent (Y, X) masks.  X <= 16384.  A mask without any zero pixel has no
 * defined answer (scipy's is implementation-specific); very large values are returned.
 *   ws   be_edt_workspace_bytes(N, Y, X) bytes */
size_t be_edt_workspace_bytes(int64_t N, int64_t Y, int64_t X);
int be_edt_u8_f64(const uint8_t* mask, double* dist, int64_t N, int64_t Y, int64_t X, void* ws, be_stream_t stream);

/* The same transform as its exact integer square (saturated at INT32_MAX): the order and the ties of the distance
 * without a square root — what the seed detection and the flood below work on. */
int be_edt_sq_u8_i32(const uint8_t* mask, int32_t* dist2, int64_t N, int64_t Y, int64_t X, void* ws, be_stream_t stream);

/* ---- seeds and marker flood (second part) ------------------------------------- */

/* `skimage.feature.peak_local_max(dist, min_distance=d, labels=mask)` with every other argument at its default
 * (analysis_deployment.py:171), on dist2 = the squared distance (any int32 image whose order is the wanted one):
 * exclude_border = d, threshold = image minimum, (2d+1)^2 window maximum clipped to the bounding box of the
 * border-stripped mask, skimage's special case for an image in which every masked pixel is a maximum, and the greedy
 * thinning of the candidates (value descending, raster order among equals) to Chebyshev distance >= d.
 *   peaks  [Y*X] uint8 out: 1 at the selected coordinates (the reference's boolean image `m`), 0 elsewhere
 *   ws     be_peak_local_max_workspace_bytes(Y, X) bytes
 * No host synchronisation; plateaus longer than the two grid-wide rounds settle are finished by one block. */
size_t be_peak_local_max_workspace_bytes(int64_t Y, int64_t X);
int be_peak_local_max_i32(const int32_t* dist2, const uint8_t* mask, uint8_t* peaks, int64_t Y, int64_t X,
                          int min_distance, void* ws, be_stream_t stream);

/* `skimage.segmentation.watershed(image, markers, mask=mask)` with connectivity 1, no compactness, no watershed line
 * (analysis_deployment.py:175) for one (Y, X) image, Y*X < 2^27.
 *   height     [Y*X] int32: the flood order (lower first); negate != 0 floods by -height (pass the squared distance)
 *   markers    [Y*X] int32: seeds (0 = none); must lie inside the mask to flood
 *   regions    [Y*X] int32: the mask as its 4-connected components, labelled 1..*n_regions (be_ccl_u8_i32 with
 *              connectivity 1), 0 outside the mask; regions flood independently, one thread block each
 *   n_regions  device pointer to the number of regions (ids beyond it count as outside the mask)
 *   out        [Y*X] int32: markers, grown over the mask; mask pixels no seed reaches stay 0
 *   ws         be_watershed_workspace_bytes(Y, X) bytes
 * A region's seeds are put in raster order by its block with a rank sort: quadratic in the seeds of ONE region (fine for
 * the handful a distance map yields, slow for a region with many thousands of seed pixels).
 * Pixels leave the queue by (height, push order).  Seeds are pushed in raster order and seeds of equal height leave in
 * that order; skimage's binary heap orders such complete ties by its array layout instead, so results can differ from
 * skimage's on the boundary between two basins whose seeds have exactly the same height. */
size_t be_watershed_workspace_bytes(int64_t Y, int64_t X);
int be_watershed_i32(const int32_t* height, int negate, const int32_t* markers, const int32_t* regions,
                     const int32_t* n_regions, int32_t* out, int64_t Y, int64_t X, void* ws, be_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* BE_B200_H_ */
