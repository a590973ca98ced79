/*
 * aiko_b200.h -- C ABI of libaiko_b200.so: the B200 (sm_100a) implementation of
 * aiko_services' per-frame media hot path.
 *
 * The reference is pure Python; it has no FFI of its own.  Each entry point
 * below replaces the third-party library call made at the cited line of the
 * reference element (R/ = src/aiko_services/ in rskew/aiko_services), batched
 * over the frames of one process_frame() call.  INTEGRATION.md shows the ctypes
 * binding a maintainer would add to those elements.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in _host;
 *   - images are uint8, interleaved channels (HWC), batched as N frames with an
 *     explicit row stride and frame stride in bytes (tight: row = W*C,
 *     frame = H*W*C);
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream); all
 *     work is enqueued asynchronously on it;
 *   - nothing here allocates or frees device memory: the caller passes outputs
 *     and, where stated, a workspace;
 *   - the return value is 0 (AIKO_B200_OK) or a negative AIKO_B200_ERR_* code;
 *     aiko_b200_last_error() gives a thread-local message.  Errors never
 *     abort() or throw across the ABI, so an element can map them to
 *     StreamEvent.ERROR (R/main/pipeline.py:1057-1066).
 */
#ifndef AIKO_B200_H_
#define AIKO_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AIKO_B200_OK 0
#define AIKO_B200_ERR_INVALID_ARGUMENT (-1)
#define AIKO_B200_ERR_UNSUPPORTED (-2)
#define AIKO_B200_ERR_CUDA (-3)
#define AIKO_B200_ERR_WORKSPACE (-4)

#define AIKO_B200_ABI_VERSION 2

int aiko_b200_abi_version(void);
const char* aiko_b200_last_error(void);
/* Name of the device the library would launch on, compute capability major*10+minor
 * in *cc (must be 100 for the kernels to load).  Returns ERR_CUDA without a GPU. */
int aiko_b200_device_info(char* name_host, size_t name_len, int* cc_host, int* sm_count_host);

/* ---------------------------------------------------------------------- */
/* Overlay draw list (ImageOverlay, R/elements/media/image_io.py:107-163). */
/* One op = one cv2.rectangle or one rendered glyph of a cv2.putText label; */
/* ops are applied in array order, per frame.                               */

#define AIKO_B200_OP_RECT 0
#define AIKO_B200_OP_GLYPH 1

typedef struct aiko_b200_overlay_op {
  int32_t kind;           /* AIKO_B200_OP_* */
  int32_t x0, y0, x1, y1; /* RECT: corner points as given to cv2.rectangle (x0<=x1, y0<=y1);
                             GLYPH: inclusive bounding box of touched pixels, image coordinates */
  uint32_t p0;            /* RECT: colour, byte c = value stored to memory channel c;
                             GLYPH: patch index for channel 0 | channel 1 << 16 */
  uint32_t p1;            /* GLYPH: patch index for channel 2 */
  uint32_t p2;            /* GLYPH: (x0 - patch_origin_x) | (y0 - patch_origin_y) << 16 */
} aiko_b200_overlay_op;

/* Glyph atlas: patch_ids[patch][patch_h][patch_w] are uint16 indices into
 * luts[index][256] (index 0 = identity); a touched pixel becomes
 * luts[id][old_value], which replays cv2.putText(..., LINE_AA) exactly. */
typedef struct aiko_b200_overlay {
  const aiko_b200_overlay_op* ops; /* device */
  const int32_t* op_offsets;       /* device, [n_lists + 1]: list i is ops[op_offsets[i] .. op_offsets[i+1]) */
  const int32_t* list_of_frame;    /* device, [n_frames]: the list frame f is drawn with, or NULL = list f.
                                      The reference applies ONE overlay to all images of a call
                                      (image_io.py:111-127), so a batch usually shares one list. */
  const uint16_t* patch_ids;       /* device, may be NULL when no GLYPH ops */
  const uint8_t* luts;             /* device */
  int32_t patch_w, patch_h;
  int32_t n_patches;
  int32_t n_lists;                 /* lists described by op_offsets */
  int32_t n_frames;                /* frames the overlay was made for; a call with more frames is refused */
  int32_t max_ops_per_list;        /* largest op count of any list, or 0 if unknown */
  /* From aiko_b200_overlay_plan_host / aiko_b200_overlay_cells_host, for frames of
   * frame_width x frame_height: the largest number of marks (blocks of 4-pixel groups an op can
   * touch) and of cells (groups in those blocks) of any list.  With them the draw list runs as
   * one thread per cell (overlay_cells_kernel); 0 = unknown, the kernels then cull the list
   * per 64 x 16 tile. */
  int32_t max_marks_per_list;
  int32_t max_cells_per_list;
  int32_t frame_width, frame_height;
  /* Optional: the planned buffer (ops | op_offsets | list_of_frame, contiguous from `ops`) still sits
   * in page-locked HOST memory; the call that draws the overlay first copies upload_bytes from here to
   * `ops` on its stream, so a list that changes every call costs no extra call.  Up to 16 KB travel as
   * a kernel argument (the host bytes are read before the call returns; the device buffer must hold
   * upload_bytes rounded up to 16), more as one cudaMemcpyAsync (the host buffer then has to be
   * page-locked and stay untouched until the copy has run).  NULL = the device buffers are filled. */
  const void* upload_from_host;
  uint64_t upload_bytes;
} aiko_b200_overlay;

/* ---------------------------------------------------------------------- */
/* Host-side planning of ImageOverlay's drawing (image_io.py:119-155): no   */
/* device work, no OpenCV.  The caller has already paired objects with      */
/* rectangles, dropped objects at or below the confidence threshold          */
/* (image_io.py:131) and formatted the label text (image_io.py:139).         */

typedef struct aiko_b200_overlay_plan_info {
  int32_t n_ops;               /* ops produced (may exceed the capacity given: ERR_WORKSPACE) */
  int32_t max_ops_per_list;
  int32_t max_marks_per_list;
  int32_t max_cells_per_list;
  uint64_t atlas_version;      /* glyph atlas state the ops refer to (see aiko_b200_overlay_atlas_host) */
  uint64_t atlas_epoch;
} aiko_b200_overlay_plan_info;

/* n_lists draw lists; list i covers objects object_offsets_host[i] .. [i+1).  rects_host: x, y, w, h per
 * object (the ints of image_io.py:132-135).  labels_host: one NUL-terminated string per object, in
 * order; an empty string = no label (NULL = no labels at all).  layout: 0 = the frames stay RGB
 * (image_io.py:117,160), 1 = stored BGR (fused RGB->BGR, or the gray round trip image_io.py:115).
 * Per object: label size as cv2.getTextSize(FONT_HERSHEY_SIMPLEX, 0.75, 2), placement rule
 * image_io.py:143-150, one GLYPH op per character (cv2.putText LINE_AA, image_io.py:151-153), then
 * the RECT op (cv2.rectangle thickness 2, image_io.py:154-155).  May grow the glyph atlas (labels
 * clipped by the image border). */
int aiko_b200_overlay_plan_host(int n_lists, const int32_t* object_offsets_host, const int32_t* rects_host,
                                const char* labels_host, int image_width, int image_height, int layout,
                                aiko_b200_overlay_op* ops_out_host, int ops_capacity,
                                int32_t* op_offsets_out_host, aiko_b200_overlay_plan_info* info_host);

/* The same planning, written straight into a caller-owned buffer pair and into the aiko_b200_overlay
 * the drawing call takes: buffer_host (page-locked, 16-byte aligned, buffer_bytes long) receives
 * ops | op_offsets | list_of_frame; overlay_host gets its device pointers set to the same offsets
 * inside buffer_dev, its counts, bounds and frame size, and upload_from_host / upload_bytes so that
 * the drawing call ships the bytes itself.  list_of_frame_host: n_frames list numbers, or NULL = every
 * frame uses list 0.  The atlas fields of overlay_host (patch_ids, luts, patch_w, patch_h, n_patches)
 * are the caller's and are left alone.  One call per new overlay on the per-frame path. */
int aiko_b200_overlay_plan_struct_host(int n_lists, const int32_t* object_offsets_host, const int32_t* rects_host,
                                       const char* labels_host, int image_width, int image_height, int layout,
                                       int n_frames, const int32_t* list_of_frame_host, void* buffer_host,
                                       size_t buffer_bytes, void* buffer_dev, aiko_b200_overlay* overlay_host,
                                       aiko_b200_overlay_plan_info* info_host);

/* Ships a planned buffer (ops | op_offsets | list_of_frame, as laid out by the caller around
 * aiko_b200_overlay_plan_host) to the device: one cudaMemcpyAsync on `stream`.  plan_host should be
 * page-locked and stay untouched until the copy has run. */
int aiko_b200_overlay_upload(void* plan_dev, const void* plan_host, size_t bytes, void* stream);

/* cv2.getTextSize(text, FONT_HERSHEY_SIMPLEX, 0.75, 2)[0] (image_io.py:141-142). */
int aiko_b200_overlay_text_size_host(const char* text_host, int* width_host, int* height_host);

/* Marks and cells of a draw list built by other means (see aiko_b200_overlay). */
int aiko_b200_overlay_cells_host(const aiko_b200_overlay_op* ops_host, int n_ops, int image_width,
                                 int image_height, int* n_marks_host, long long* n_cells_host);

/* The marks of one op for frames image_width x image_height (what overlay_cells_kernel enumerates):
 * up to 4 blocks of 4-pixel groups, written as int16 {first row, last row, first group, last group}
 * (group g = pixels 4g .. 4g+3).  Every pixel the op can change lies in one of them.  Returns the
 * number of marks (0..4) or a negative error code. */
int aiko_b200_overlay_marks_host(const aiko_b200_overlay_op* op_host, int image_width, int image_height,
                                 int16_t* marks_out_host);

/* The host copy of the glyph atlas, to be mirrored on each GPU by the caller (the library never
 * allocates device memory).  Rows only ever get appended while `epoch` stays the same -- copy rows
 * [old n .. new n) when `version` moved; a new epoch means the border-clipped glyph variants were
 * dropped to make room and indices beyond the first 761 patches were re-used: upload afresh, into
 * new device buffers if kernels using the old ones may still be queued.  Pointers stay valid for
 * the life of the process (capacity max_patches / max_luts is reserved up front). */
typedef struct aiko_b200_atlas_view {
  const uint16_t* patch_ids_host; /* [n_patches][patch_h][patch_w] */
  const uint8_t* luts_host;       /* [n_luts][256] */
  int32_t patch_w, patch_h, n_patches, n_luts, max_patches, max_luts;
  uint64_t version, epoch;
} aiko_b200_atlas_view;
int aiko_b200_overlay_atlas_host(aiko_b200_atlas_view* view_host);

/* What a planner needs to know about the 95 x 4 interior glyphs of the atlas (used by the device-side planner
 * below): advance per character 32..126 in glyph units (pixels = units x 0.75), per (character, pen phase) the
 * inclusive box x0, y0, x1, y1 of the touched pixels inside the patch window (x1 < x0: blank), and the
 * planner's constants {patch_w, patch_h, pad_x, pad_y, border_margin, text_height, thickness, 0}. */
int aiko_b200_overlay_glyph_metrics_host(uint8_t* advance_out_host, int16_t* boxes_out_host, int32_t* constants_out_host);

/* ImageOverlay for detections that never leave the GPU (a detector element's output tensors instead of the
 * overlay dict that R/examples/yolo/yolo.py:56-87 builds on the host with six .item() calls per box): one
 * kernel plans the draw lists on the device -- confidence threshold (image_io.py:131), label
 * f"{name}: {confidence:0.2f}" (image_io.py:139), label size and placement (image_io.py:140-150), one op per
 * glyph, the rectangle (image_io.py:154-155) -- and the lists are drawn on the RGB frames in place.
 * boxes: int32 [n_lists][max_objects][4] = x, y, w, h; class_ids: int32 [n_lists][max_objects], an index into
 * names (outside 0..n_names-1 or an empty name: no label); confidences: float32 [n_lists][max_objects],
 * slots with confidence <= 0 (or NaN) are skipped, as the reference skips them; n_lists = 1 (the reference: one
 * overlay for all images of a call) or n_frames (one list per frame).  names: char [n_names][name_stride], NUL
 * padded; patch_ids / luts / n_patches: the caller's device copy of the glyph atlas (aiko_b200_overlay_atlas_host);
 * variant_*: its border variants (aiko_b200_overlay_variants_host, below; n_variants may be 0).  All of these are
 * DEVICE pointers.  layout as in aiko_b200_overlay_plan_host.  workspace: 16-byte aligned device
 * memory of aiko_b200_overlay_boxes_workspace_bytes.  *status (device int32) receives 0 = drawn, or -- with
 * nothing drawn at all -- 1 = a label comes within the border margin of the frame and needs a glyph variant that
 * is not in the table yet (take the host path for this call: it rasterises the variant, refresh the table after), 2 = more ops than the workspace holds,
 * 3 = a confidence that is not finite or >= 1e9. */
size_t aiko_b200_overlay_boxes_workspace_bytes(int n_lists, int max_objects, int n_frames);
int aiko_b200_overlay_boxes_u8(uint8_t* frames, int n_frames, int height, int width, int64_t frame_stride, int layout,
                               const int32_t* boxes, const int32_t* class_ids, const float* confidences, int n_lists,
                               int max_objects, const char* names, int n_names, int name_stride,
                               const uint16_t* patch_ids, const uint8_t* luts, int n_patches,
                               const uint64_t* variant_keys, const int32_t* variant_firsts, const int16_t* variant_boxes,
                               int n_variants, void* workspace, size_t workspace_bytes, int32_t* status, void* stream);
/* The border variants of the glyph atlas rasterised so far, as the table aiko_b200_overlay_boxes_u8 takes (device
 * copies of the three arrays): keys sorted ascending, first patch per variant (0 = the glyph does not reach the
 * image), inclusive box of the touched pixels inside the patch window.  -> number of variants (only the first
 * `capacity` are written; the arrays may be NULL to just count); version / epoch: the atlas state they belong to. */
int aiko_b200_overlay_variants_host(uint64_t* keys_out_host, int32_t* firsts_out_host, int16_t* boxes_out_host,
                                    int capacity, uint64_t* version_out_host, uint64_t* epoch_out_host);

/* ---------------------------------------------------------------------- */
/* ImageResize, ndarray branch: cv2.resize(image, (W, H)), INTER_LINEAR,   */
/* bit-exact fixed-point arithmetic -- R/elements/media/image_io.py:216.   */
/* Optionally fused with what the reference runs next on the same frames:  */
/*   overlay  != NULL : ImageOverlay's drawing (image_io.py:107-163)       */
/*   swap_rb  != 0    : cv2.cvtColor RGB<->BGR (video_io.py:129,299)       */
/*                      applied to the stored pixels (op colours refer to   */
/*                      the stored channel order).                          */
/* channels in {1, 3, 4}.                                                   */
int aiko_b200_resize_linear_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                               int64_t src_row_stride, int64_t src_frame_stride, uint8_t* dst, int dst_h,
                               int dst_w, int64_t dst_row_stride, int64_t dst_frame_stride, int swap_rb,
                               const aiko_b200_overlay* overlay_host, void* stream);

/* The per-frame call of a detection pipeline, ImageResize -> ImageOverlay (-> RGB2BGR) with an overlay that is
 * new on every call (image_io.py:119-155 runs for every frame: boxes move, the label text carries the
 * confidence): aiko_b200_overlay_plan_struct_host followed by aiko_b200_resize_linear_u8 as ONE call, so the
 * host pays one crossing of the ABI per frame batch.  Arguments as in those two; the draw lists are planned for
 * the dst_w x dst_h frames.  overlay_host->patch_ids / luts are the caller's device copy of the glyph atlas in
 * the state (atlas_version, atlas_epoch); when planning had to grow the atlas past that state nothing is
 * launched and AIKO_B200_ATLAS_GREW is returned: overlay_host is planned, the caller brings its device copy
 * up to date (aiko_b200_overlay_atlas_host) and calls aiko_b200_resize_linear_u8 itself. */
#define AIKO_B200_ATLAS_GREW 1
int aiko_b200_resize_linear_objects_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                       int64_t src_row_stride, int64_t src_frame_stride, uint8_t* dst, int dst_h,
                                       int dst_w, int64_t dst_row_stride, int64_t dst_frame_stride, int swap_rb,
                                       int n_lists, const int32_t* object_offsets_host, const int32_t* rects_host,
                                       const char* labels_host, int layout, const int32_t* list_of_frame_host,
                                       void* buffer_host, size_t buffer_bytes, void* buffer_dev,
                                       aiko_b200_overlay* overlay_host, aiko_b200_overlay_plan_info* info_host,
                                       uint64_t atlas_version, uint64_t atlas_epoch, void* stream);

/* ImageResize, PIL branch: PIL.Image.resize((W, H)) -- default BICUBIC filter   */
/* with antialiasing, 8 bits per channel, horizontal pass then vertical pass on a  */
/* uint8 intermediate -- R/elements/media/image_io.py:218.  Bit-exact.             */
/* The two fixed-point weight tables live in a caller-provided device workspace:   */
/*   bytes   = aiko_b200_resize_bicubic_workspace_bytes(...)  (0 for bad sizes)    */
/*   prepare = fills it for one (src, dst) size pair; reusable for any number of   */
/*             resize calls with that size pair, on any stream ordered after it.   */
/* Frames are tight: [n, h, w, channels], channels in {1, 3, 4}.  The workspace      */
/* holds the 22-bit coefficients and, for reductions up to ~5x, the same weights as  */
/* integer-MMA operand tables (237 KB in all for 1080p -> 640x480).  After prepare,   */
/* int32 word 1 of the workspace is non-zero if a table overflowed (not for any size  */
/* the library admits); a resize with such a workspace leaves dst untouched.  Source  */
/* rows that do not start on 16-byte boundaries are fetched from the boundary below   */
/* to the boundary above: src must be readable for 15 bytes before its first and      */
/* after its last byte (any cudaMalloc / torch allocation is).                        */
size_t aiko_b200_resize_bicubic_workspace_bytes(int src_h, int src_w, int dst_h, int dst_w);
int aiko_b200_resize_bicubic_prepare(void* workspace, size_t workspace_bytes, int src_h, int src_w, int dst_h,
                                     int dst_w, void* stream);
int aiko_b200_resize_bicubic_aa_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                   uint8_t* dst, int dst_h, int dst_w, const void* workspace,
                                   size_t workspace_bytes, void* stream);

/* The source rows aiko_b200_resize_linear_u8 reads for src_h -> dst_h (needed_host[y] = 1): a 2.25x   */
/* vertical reduction touches 8 rows of every 9.  With aiko_b200_copy_rows_h2d a caller that stages    */
/* frames for a resize ships only the rows that carry weight over PCIe: the frames of a batch are      */
/* contiguous (frame_bytes = rows x row_bytes), runs of equal length at equal distance travel as one   */
/* cudaMemcpy2DAsync (1080 -> 480: two copies per batch), patterns needing more than 8 copies are      */
/* copied whole.                                                                                        */
int aiko_b200_resize_linear_rows_host(int src_h, int dst_h, uint8_t* needed_host);
int aiko_b200_copy_rows_h2d(uint8_t* dst, const uint8_t* src_host, int n_frames, int64_t frame_bytes, int64_t row_bytes,
                            const int32_t* runs_host, int n_runs, void* stream);

/* What PIL.Image.resize does around the filter for other modes (image_io.py:218 resizes   */
/* whatever mode the file had):                                                             */
/*   premultiply_u8 : RGBA (channels 4) / LA (channels 2) -> 4-byte premultiplied pixels     */
/*                    (R, G, B, A) / (L, L, L, A) before the filter (inverse = 0), and back  */
/*                    after it (inverse = 1); Pillow's MULDIV255 and CLIP8(255 * v / a).     */
/*                    Run aiko_b200_resize_bicubic_aa_u8 with channels = 4 in between.       */
/*   gather_nearest_u8 : modes P and 1 are resized with NEAREST: dst[y][x] = src[ytab[y]][xtab[x]]; */
/*                    xtab / ytab (device int32, dst_w / dst_h entries; -1 = outside) are      */
/*                    the source positions Pillow pretabulates: (int)(0.5 * s + k * s summed   */
/*                    in double), s = src / dst.                                               */
int aiko_b200_premultiply_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int channels, int inverse,
                             void* stream);
int aiko_b200_gather_nearest_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels, uint8_t* dst,
                                int dst_h, int dst_w, const int32_t* xtab, const int32_t* ytab, void* stream);

/* Same resize for callers that cannot allocate device memory where they call it    */
/* (an element running on the pipeline's one event-loop thread, where cudaMalloc     */
/* would stall every stream; a process whose device memory is all owned by a pool;   */
/* a minimal binding that only has the frame buffers).  No counterpart in the        */
/* reference.  The weight tables go into a 256 KiB array built into the library      */
/* (resident on every GPU that loads it); size pairs needing more are refused with   */
/* AIKO_B200_ERR_UNSUPPORTED (the limit applies to the coefficient tables: 73 KB for */
/* 1080p -> 640x480; the tensor-core operand tables, 164 KB more, are used when they */
/* fit as well, else the slower scalar kernel runs).  The first call on a            */
/* device looks up that array's device address (cudaGetSymbolAddress): make it       */
/* before any stream capture.  All calls share the one array and each rewrites it:   */
/* calls that can overlap on the GPU (different streams, no ordering) race -- give   */
/* such callers the workspace variant above.  Allocates and frees nothing.           */
int aiko_b200_resize_bicubic_aa_u8_noalloc(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                           uint8_t* dst, int dst_h, int dst_w, void* stream);

/* ImageOverlay alone on already-resident frames (image_io.py:107-163).     */
/* channels == 3: frames are drawn in place of a copy src -> dst (src may   */
/* equal dst).  channels == 1: the GRAY -> BGR -> draw -> BGR2GRAY round     */
/* trip of image_io.py:115,158.                                             */
int aiko_b200_overlay_u8(const uint8_t* src, uint8_t* dst, int n_frames, int h, int w, int channels,
                         int64_t row_stride, int64_t frame_stride, const aiko_b200_overlay* overlay_host,
                         void* stream);

/* BASELINE config 4 (extension, no reference element: the closest consumer is  */
/* R/examples/yolo/yolo.py:62-64 handing resized frames to a torch model):        */
/* cv2-exact bilinear resize of RGB frames [n, src_h, src_w, 3], then             */
/* out[n][c][y][x] = ((float)u8 / 255 - mean[c]) / std[c] as fp32 NCHW.           */
/* mean_host / std_host: 3 floats each in HOST memory.                            */
int aiko_b200_resize_norm_f32_nchw(const uint8_t* src, int n_frames, int src_h, int src_w, float* dst, int dst_h,
                                   int dst_w, const float* mean_host, const float* std_host, void* stream);

/* ---------------------------------------------------------------------- */
/* Baseline JPEG decoding on the GPU: the decode half of PIL.Image.open(path) ... image.resize(...)  */
/* (R/elements/media/image_io.py:184,218 -- Pillow decodes lazily, inside the resize call), so that   */
/* frames are born device-resident and only the compressed bytes cross PCIe.  The kernels restate      */
/* libjpeg-turbo's default decoder (Huffman decoding, accurate integer IDCT, "fancy" chroma             */
/* upsampling, fixed-point YCbCr -> RGB): the decoded pixels are BIT-IDENTICAL to Pillow's.             */
/* Subset: SOF0, 8 bit, one interleaved scan, gray or YCbCr (4:4:4, 4:2:2, 4:2:0), with or without      */
/* restart markers -- entropy decoding runs one thread per subsequence of the scan and finds the         */
/* decoder state at the subsequence boundaries by iteration (images with more than 512 restart           */
/* segments: one thread per segment).  Progressive / arithmetic / 12-bit files, RGB- or CMYK-coded       */
/* files and other sampling layouts return AIKO_B200_ERR_UNSUPPORTED (use another decoder for those,     */
/* e.g. libaiko_b200_jpeg.so).                                                                           */
/* All images of a call have one geometry.  dst: uint8 [n][height][width][3] RGB at frame_stride.        */
/*   info_host       : width, height, components (1 / 3), luma sampling, restart interval of one file    */
/*   staging_bytes   : page-locked host buffer the call fills with descriptors and tables                */
/*   workspace_bytes : device workspace for n images of that geometry holding up to compressed_capacity  */
/*                     bytes of entropy-coded data in all and max_segments restart segments per image    */
/*                     (quantised coefficients and component planes live there: ~9.4 MB per 1080p        */
/*                     4:2:0 frame, plus two copies of the compressed data); 0 for an unsupported geometry;                                      */
/*   decode_rgb_u8   : everything is enqueued on `stream`; the compressed scans are copied straight      */
/*                     from jpegs_host (asynchronously when that memory is page-locked; they and the     */
/*                     staging buffer must stay untouched until the copies have run)                     */
/*   status          : synchronises `stream` and reports corrupt entropy-coded data met by the last      */
/*                     decode that used the workspace (the kernels never read or write out of bounds)    */
int aiko_b200_jpeg_gpu_info_host(const uint8_t* jpeg_host, size_t length, int* width_host, int* height_host,
                                 int* components_host, int* hmax_host, int* vmax_host, int* restart_interval_host);
size_t aiko_b200_jpeg_gpu_staging_bytes(int n_images);
size_t aiko_b200_jpeg_gpu_workspace_bytes(int n_images, int width, int height, int components, int hmax, int vmax,
                                          size_t compressed_capacity, int max_segments);
int aiko_b200_jpeg_gpu_decode_rgb_u8(const uint8_t* const* jpegs_host, const size_t* lengths_host, int n_images,
                                     uint8_t* dst, int height, int width, int64_t frame_stride, void* staging_host,
                                     size_t staging_bytes, void* workspace, size_t workspace_bytes,
                                     size_t compressed_capacity, int max_segments, void* stream);
int aiko_b200_jpeg_gpu_status(const void* workspace, int n_images, int width, int height, int components, int hmax,
                              int vmax, size_t compressed_capacity, int max_segments, void* stream);

/* ImageWriteFile: image.save(path) for a JPEG path -- R/elements/media/image_io.py:250 -- without the frames
 * leaving the GPU uncompressed.  Baseline JPEG files byte-identical to what Pillow (libjpeg-turbo, standard Huffman
 * tables, JFIF 1.01 header) writes for the same pixels: quality as Pillow's `quality` (the reference's default
 * call: 75), luma sampling hs x vs = 2x2 (4:2:0, Pillow's default for RGB), 2x1 (4:2:2) or 1x1 (4:4:4);
 * channels 3 = RGB frames, 1 = gray frames (mode L).  src: n_images frames of height x width x channels bytes on
 * the device.  out (device): n_images slots of out_stride bytes, slot i receives file i; sizes_out (device,
 * int32 per image): the file's length, or -1 when its scan did not fit scan_capacity bytes or the file did not
 * fit the slot (call again with more room: a scan never needs more than twice the raw frame).  workspace:
 * 256-byte aligned, aiko_b200_jpeg_gpu_encode_workspace_bytes for the same arguments (0 = unsupported).
 * Asynchronous on `stream`. */
size_t aiko_b200_jpeg_gpu_encode_workspace_bytes(int n_images, int width, int height, int channels, int hs, int vs,
                                                 size_t scan_capacity);
int aiko_b200_jpeg_gpu_encode_u8(const uint8_t* src, int n_images, int height, int width, int channels,
                                 int64_t frame_stride, int quality, int hs, int vs, void* workspace,
                                 size_t workspace_bytes, size_t scan_capacity, uint8_t* out, size_t out_stride,
                                 int32_t* sizes_out, void* stream);

/* ---------------------------------------------------------------------- */
/* Colour conversion / flip / crop call sites.                              */

/* cv2.cvtColor(.., COLOR_RGB2BGR / COLOR_BGR2RGB): image_io.py:117,160;   */
/* video_io.py:129,214-216,299; webcam_io.py:122.  n_pixels 3-byte pixels.  */
int aiko_b200_swap_rb_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, void* stream);
/* cv2.cvtColor(.., COLOR_BGR2GRAY) (rgb_order=0) / COLOR_RGB2GRAY (1):      */
/* webcam_io.py:124, image_io.py:158.                                        */
int aiko_b200_bgr2gray_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int rgb_order, void* stream);
/* RGB / BGR <-> YUV, 8 bits per channel (extension: BASELINE.json's north_star names YUV     */
/* conversion, the reference has no call site for it).  Arithmetic: cv2.cvtColor(..,          */
/* COLOR_RGB2YUV / COLOR_BGR2YUV) for inverse = 0 and COLOR_YUV2RGB / COLOR_YUV2BGR for       */
/* inverse = 1 (bgr_order picks the BGR variants), 14-bit fixed point, bit-exact with OpenCV   */
/* over all 2^24 colours.  n_pixels 3-byte pixels.                                            */
int aiko_b200_yuv_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int bgr_order, int inverse, void* stream);
/* Alpha blending (extension: BASELINE.json's north_star names it, the reference has no call   */
/* site): the RGBA image rgba[image_h][image_w][4] (device, 4-byte aligned) composited onto every  */
/* RGB frame at (x, y), in place, with the arithmetic of PIL's frame.paste(image, (x, y), image):  */
/* t = dst * (255 - a) + src * a + 128, out = ((t >> 8) + t) >> 8; the box is clipped to the frame. */
int aiko_b200_alpha_paste_u8(uint8_t* frames, int n_frames, int height, int width, int64_t frame_stride,
                             const uint8_t* rgba, int image_h, int image_w, int x, int y, void* stream);
/* cv2.cvtColor(.., COLOR_GRAY2BGR): image_io.py:115.                        */
int aiko_b200_gray2bgr_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, void* stream);
/* cv2.flip(image, flip_code): webcam_io.py:127 (1, horizontal), :129 (0,    */
/* vertical); flip_code < 0 flips both.  Tight frames [n, h, w, channels].   */
int aiko_b200_flip_u8(const uint8_t* src, uint8_t* dst, int n_frames, int h, int w, int channels,
                      int flip_code, void* stream);
/* image[y0:y0+ch, x0:x0+cw] (numpy slicing as used on frames; north_star).  */
int aiko_b200_crop_u8(const uint8_t* src, int n_frames, int h, int w, int channels, uint8_t* dst, int x0,
                      int y0, int crop_w, int crop_h, void* stream);

/* ---------------------------------------------------------------------- */
/* Audio chain, batched over windows (the reference code is dead -- inside a    */
/* string literal, R/elements/media/audio_io.py:162-642 -- so parity is against */
/* a numpy restatement, oracle/audio.py):                                        */
/*   PE_FFT            :306-320  amplitudes = |fft(window)| / 1e6, all n bins    */
/*   PE_AudioResampler :250-277  sums of 8 bands of the first half of the bins   */
/*   PE_AudioFilter    :213-230  0.1 <= amplitude <= 12, 10 <= |f| <= 9000 Hz,   */
/*                               100 largest by amplitude (stable order)         */
/* audio: [n_windows, n_samples] int16 (dtype 0) or float32 (dtype 1);           */
/* n_samples even, n_samples / 2 = 2^a 3^b 5^c <= 8192.  The workspace holds the */
/* twiddle tables (generated in double precision by _prepare).  band_ranges_host */
/* (16 ints: lo, hi per band) and gate_lo / gate_hi (folded bin range passing    */
/* the frequency gate) are evaluated by the caller in float64 as numpy would.    */
/* band_sums and the three top_* outputs may be NULL to skip those stages.       */
size_t aiko_b200_audio_workspace_bytes(int n_samples);
int aiko_b200_audio_prepare(void* workspace, size_t workspace_bytes, int n_samples, void* stream);
int aiko_b200_audio_chain(const void* audio, int dtype, int n_windows, int n_samples, double sample_rate,
                          float* amplitudes, float* band_sums, const int* band_ranges_host, float* top_freq,
                          float* top_amp, int* top_count, int gate_lo, int gate_hi, const void* workspace,
                          size_t workspace_bytes, void* stream);

/* The same two elements for (amplitudes, frequencies) arrays that did not come from         */
/* aiko_b200_audio_chain -- the reference's elements accept any float64 array pair                */
/* (audio_io.py:213-216, 250-251).  float64 in, float64 out, one CTA per window:                  */
/*   audio_filter_f64 : audio_io.py:216-227; outputs [n_windows][100] (|f|, a) sorted by a          */
/*                      descending (ties in original order) and the number of valid entries.        */
/*   audio_bands_f64  : audio_io.py:253-277; outputs [n_windows][8] band centres and band sums.     */
/* frequency_stride: elements between the frequency rows of consecutive windows; 0 = one row        */
/* shared by all windows (np.fft.fftfreq is the same for every window, audio_io.py:312-313).       */
int aiko_b200_audio_filter_f64(const double* amplitudes, const double* frequencies, int n_windows, int n,
                               int64_t frequency_stride, double* top_frequencies, double* top_amplitudes,
                               int* top_count, void* stream);
int aiko_b200_audio_bands_f64(const double* amplitudes, const double* frequencies, int n_windows, int n,
                              int64_t frequency_stride, double* band_frequencies, double* band_amplitudes,
                              void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AIKO_B200_H_ */
