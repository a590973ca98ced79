// ImageOverlay for detections that never leave the GPU (second half).  The reference's
// detector element reads every box, class and confidence back to the host (six .item() calls per box,
// R/examples/yolo/yolo.py:66-80), builds the overlay dict there, and ImageOverlay draws it
// (R/elements/media/image_io.py:119-155).  Here the model's output tensors are the input: one small kernel does
// what the host planner does per object (csrc/overlay_plan.cpp plan_pass) -- confidence threshold, label text
// f"{name}: {confidence:0.2f}", label size and placement rule, one GLYPH op per character, the RECT op -- and
// writes the 32-byte op records straight into device memory, where the drawing kernels read them.
//
// A glyph whose strokes come within the border margin of the frame needs a variant rasterised against that border
// geometry, which only the host rasteriser makes.  The variants made so far travel to the device as a sorted key
// table (aiko_b200_overlay_variants_host); when a label of the call needs one that is not there yet, the planner
// draws NOTHING and reports status 1: the caller takes the host path for that call (boxes read back,
// aiko_b200_overlay_plan_host -- which rasterises the missing variants, so the next call finds them).
#include <math.h>
#include <string.h>

#include <mutex>

#include "overlay.cuh"

extern "C" int aiko_b200_overlay_glyph_metrics_host(uint8_t* advance_out_host, int16_t* boxes_out_host,
                                                    int32_t* constants_out_host);

namespace aiko {

namespace {

constexpr int kPlanThreads = 1024;
constexpr int kMaxLabel = 64;                 // characters of a label on the device path

struct GlyphMetrics {                         // by value as a kernel argument
  uint8_t advance[96];
  short4 box[95 * 4];                         // per (character, pen phase): x0, y0, x1, y1 inside the patch window
  int patch_w, patch_h, pad_x, pad_y, margin, text_height, thickness;
};

struct BoxesParams {
  const int32_t* boxes;                       // [n_lists][max_objects][4]: x, y, w, h
  const int32_t* class_ids;                   // [n_lists][max_objects]; < 0 or >= n_names: no label
  const float* confidences;                   // [n_lists][max_objects]
  const char* names;                          // [n_names][name_stride], NUL padded
  int n_lists, max_objects, n_names, name_stride;
  int W, H, layout, n_frames;
  const unsigned long long* variant_keys;     // the border variants the host has rasterised so far, sorted by key
  const int32_t* variant_firsts;
  const short4* variant_boxes;
  int n_variants;
  aiko_b200_overlay_op* ops;
  int ops_capacity;
  int32_t* op_offsets;                        // [n_lists + 1]
  int32_t* list_of_frame;                     // [n_frames]
  int32_t* status;                            // 0 drawn, 1 a label needs a border variant, 2 ops do not fit, 3 unsupported value
};

__device__ __forceinline__ int glyph_slot(unsigned char c) { return (c < 32 || c > 126) ? '?' - 32 : c - 32; }

// f"{name}: {confidence:0.2f}": correctly rounded (ties to even on the exact binary value, as Python's float
// formatting), for a float32 confidence that was widened to double without change.  -> length, or -1 when the
// value is not finite or needs more than 9 integer digits.
__device__ int format_label(const char* name, int name_stride, float confidence, unsigned char* out) {
  int n = 0;
  while (n < name_stride && name[n] && n < kMaxLabel - 16) { out[n] = (unsigned char)name[n]; n++; }
  out[n++] = ':'; out[n++] = ' ';
  if (!(confidence >= 0.0f) || isinf(confidence) || confidence >= 1.0e9f) return -1;
  const uint32_t bits = __float_as_uint(confidence);
  const int exponent = (int)((bits >> 23) & 0xffu);
  const unsigned long long mantissa = exponent ? ((bits & 0x7fffffu) | 0x800000u) : (bits & 0x7fffffu);
  const int e = (exponent ? exponent : 1) - 150;          // value = mantissa * 2^e
  unsigned long long hundredths;                          // round_half_even(value * 100)
  const unsigned long long scaled = mantissa * 100ull;    // < 2^31
  if (e >= 0) {
    hundredths = scaled << e;                             // (value < 1e9: e <= 6)
  } else if (-e >= 40) {
    hundredths = 0;                                       // value * 100 < 2^-9
  } else {
    const int s = -e;
    const unsigned long long q = scaled >> s, r = scaled & ((1ull << s) - 1ull), half = 1ull << (s - 1);
    hundredths = q + ((r > half || (r == half && (q & 1ull))) ? 1ull : 0ull);
  }
  unsigned long long whole = hundredths / 100ull;
  const int cents = (int)(hundredths - whole * 100ull);
  unsigned char digits[12];
  int nd = 0;
  do { digits[nd++] = (unsigned char)('0' + (int)(whole % 10ull)); whole /= 10ull; } while (whole);
  while (nd) out[n++] = digits[--nd];
  out[n++] = '.';
  out[n++] = (unsigned char)('0' + cents / 10);
  out[n++] = (unsigned char)('0' + cents % 10);
  return n;
}

// The ops of one object, in the order the reference draws (label glyphs, then the rectangle): the device-side
// twin of plan_pass in csrc/overlay_plan.cpp.  EMIT = false only counts.  -> number of ops; *needs_host is set
// when a glyph would need a border variant or a value cannot be formatted here.
template <bool EMIT>
__device__ int plan_object(const BoxesParams& p, const GlyphMetrics& gm, int object, aiko_b200_overlay_op* out, int* needs_host) {
  const float confidence = p.confidences[object];
  if (!(confidence > 0.0f)) return 0;                     // image_io.py:131 (NaN: skipped too)
  const int x = p.boxes[4 * object], y = p.boxes[4 * object + 1], w = p.boxes[4 * object + 2], h = p.boxes[4 * object + 3];
  const int cls = p.class_ids[object];
  const int W = p.W, H = p.H;
  const int comp[3] = {p.layout == 0 ? 1 : 0, 1, p.layout == 0 ? 0 : 1};
  int n_ops = 0;
  if (cls >= 0 && cls < p.n_names && p.names[(size_t)cls * p.name_stride]) {
    unsigned char text[kMaxLabel];
    const int len = format_label(p.names + (size_t)cls * p.name_stride, p.name_stride, confidence, text);
    if (len < 0) { *needs_host = 3; return 0; }
    long units = 0;
    for (int i = 0; i < len; i++) units += gm.advance[glyph_slot(text[i])];
    const int tw = __double2int_rn((double)units * 0.75 + (double)gm.thickness), th = gm.text_height;   // image_io.py:140-150
    int tx = x, ty;
    if ((double)y > (double)th * 1.5) ty = (int)((double)y - (double)th / 2.0 - 2.0);
    else ty = y + h + th + 2;
    if (x + tw > W) { tx = x - tw; ty = y + th; }
    units = 0;
    for (int i = 0; i < len; i++) {
      const int slot = glyph_slot(text[i]);
      const long pen_q = 3 * units;                       // pen position in quarter pixels
      units += gm.advance[slot];
      const int q = (int)(pen_q & 3);
      const long long pxl = (long long)tx + (pen_q >> 2) - gm.pad_x, pyl = (long long)ty - gm.pad_y;
      if (pxl > W || pxl + gm.patch_w < 0 || pyl > H || pyl + gm.patch_h < 0) continue;      // window off the image
      const int px = (int)pxl, py = (int)pyl;
      const short4 u = gm.box[slot * 4 + q];
      if (u.z < u.x) continue;                            // blank
      int first = 1 + (slot * 4 + q) * 2;
      short4 box = u;
      if (!(px + u.x >= gm.margin && py + u.y >= gm.margin && px + u.z < W - gm.margin && py + u.w < H - gm.margin)) {
        // near a border: the variant rasterised against this border geometry, if the host has made it already
        // (csrc/overlay_plan.cpp glyph_patches: the same key)
        const int far = 1 << 20;
        const int left = (px + u.x < gm.margin) ? px : -far, top = (py + u.y < gm.margin) ? py : -far;
        const int right = (px + u.z >= W - gm.margin) ? W - px : far, bottom = (py + u.w >= H - gm.margin) ? H - py : far;
        const bool ok = (left == -far || (left >= -127 && left <= 126)) && (top == -far || (top >= -127 && top <= 126)) &&
                        (right == far || (right >= -127 && right <= 126)) && (bottom == far || (bottom >= -127 && bottom <= 126));
        int found = -1;
        if (ok) {
          const unsigned long long key = ((unsigned long long)slot << 34) | ((unsigned long long)q << 32) |
                                         ((unsigned long long)(left == -far ? 0 : left + 128) << 24) |
                                         ((unsigned long long)(top == -far ? 0 : top + 128) << 16) |
                                         ((unsigned long long)(right == far ? 255 : right + 128) << 8) |
                                         (unsigned long long)(bottom == far ? 255 : bottom + 128);
          int lo = 0, hi = p.n_variants - 1;
          while (lo <= hi) {
            const int mid = (lo + hi) >> 1;
            const unsigned long long k = p.variant_keys[mid];
            if (k == key) { found = mid; break; }
            if (k < key) lo = mid + 1; else hi = mid - 1;
          }
        }
        if (found < 0) {
          *needs_host = 1;                                // not rasterised yet: only the host rasteriser can
          continue;
        }
        first = p.variant_firsts[found];
        if (first == 0) continue;                         // nothing of the glyph lands on the image
        box = p.variant_boxes[found];
        if (box.z < box.x) continue;
      }
      const int cx0 = max(px + box.x, 0), cy0 = max(py + box.y, 0), cx1 = min(px + box.z, W - 1), cy1 = min(py + box.w, H - 1);
      if (cx1 < cx0 || cy1 < cy0) continue;
      if (EMIT) {
        aiko_b200_overlay_op op;
        op.kind = AIKO_B200_OP_GLYPH;
        op.x0 = cx0; op.y0 = cy0; op.x1 = cx1; op.y1 = cy1;
        op.p0 = (uint32_t)(first + comp[0]) | ((uint32_t)(first + comp[1]) << 16);
        op.p1 = (uint32_t)(first + comp[2]);
        op.p2 = (uint32_t)(cx0 - px) | ((uint32_t)(cy0 - py) << 16);
        out[n_ops] = op;
      }
      n_ops++;
    }
  }
  const long long x1l = (long long)x + w, y1l = (long long)y + h;                            // image_io.py:154-155
  long long xa = x < x1l ? x : x1l, xb = x < x1l ? x1l : x, ya = y < y1l ? y : y1l, yb = y < y1l ? y1l : y;
  if (xb + 1 < 0 || yb + 1 < 0 || xa - 1 >= W || ya - 1 >= H) return n_ops;
  const long long lim = 1 << 24;
  xa = max(-lim, min(lim, xa)); xb = max(-lim, min(lim, xb)); ya = max(-lim, min(lim, ya)); yb = max(-lim, min(lim, yb));
  if (EMIT) {
    aiko_b200_overlay_op op;
    op.kind = AIKO_B200_OP_RECT;
    op.x0 = (int)xa; op.y0 = (int)ya; op.x1 = (int)xb; op.y1 = (int)yb;
    op.p0 = p.layout == 0 ? 0x00ffffu : 0xffff00u; op.p1 = 0; op.p2 = 0;
    out[n_ops] = op;
  }
  return n_ops + 1;
}

// One CTA plans the whole call: a thread per object slot, two passes (count, block scan, write).
__global__ void __launch_bounds__(kPlanThreads) overlay_boxes_plan_kernel(const BoxesParams p, const __grid_constant__ GlyphMetrics gm) {
  __shared__ int warp_sums[kPlanThreads / 32];
  __shared__ int s_carry;
  const int tid = (int)threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int total = p.n_lists * p.max_objects;
  if (tid == 0) s_carry = 0;
  const int nthreads = (int)blockDim.x;
  for (int i = tid; i < p.n_frames; i += nthreads) p.list_of_frame[i] = p.n_lists == 1 ? 0 : i;
  __syncthreads();
  // pass 1: does any label need the host?  do the ops fit?
  int flag = 0, mine = 0;
  for (int base = 0; base < total; base += nthreads) {
    const int object = base + tid;
    if (object < total) mine += plan_object<false>(p, gm, object, nullptr, &flag);
  }
  flag = __syncthreads_or(flag == 1) ? 1 : (__syncthreads_or(flag == 3) ? 3 : 0);
  int sum = mine;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  if (lane == 0) warp_sums[warp] = sum;
  __syncthreads();
  int all = 0;
  for (int w = 0; w < nthreads / 32; w++) all += warp_sums[w];
  if (flag == 0 && all > p.ops_capacity) flag = 2;
  __syncthreads();
  if (flag != 0) {                                         // (uniform) nothing is drawn: empty lists
    for (int i = tid; i <= p.n_lists; i += nthreads) p.op_offsets[i] = 0;
    if (tid == 0) *p.status = flag;
    return;
  }
  // pass 2: objects in order, kPlanThreads at a time
  for (int base = 0; base < total; base += nthreads) {
    const int object = base + tid;
    int unused = 0;
    // (a call that fits one round of the CTA counted its objects in pass 1 already)
    const int count = total <= nthreads ? mine : (object < total ? plan_object<false>(p, gm, object, nullptr, &unused) : 0);
    int incl = count;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    int before = s_carry, chunk = 0;
    for (int w = 0; w < nthreads / 32; w++) {
      const int v = warp_sums[w];
      before += w < warp ? v : 0;
      chunk += v;
    }
    const int at = before + incl - count;
    if (object < total) {
      if (object % p.max_objects == 0) p.op_offsets[object / p.max_objects] = at;
      if (count) plan_object<true>(p, gm, object, p.ops + at, &unused);
    }
    __syncthreads();
    if (tid == 0) s_carry += chunk;
    __syncthreads();
  }
  if (tid == 0) {
    p.op_offsets[p.n_lists] = s_carry;
    *p.status = 0;
  }
}

struct BoxesLayout { size_t ops, offsets, frames, total; int ops_capacity; };

BoxesLayout boxes_layout(int n_lists, int max_objects, int n_frames) {
  BoxesLayout L{};
  const long long cap = (long long)n_lists * max_objects * (kMaxLabel + 1);
  L.ops_capacity = (int)(cap < (1 << 22) ? cap : (1 << 22));
  size_t at = 0;
  L.ops = at; at += (size_t)L.ops_capacity * sizeof(aiko_b200_overlay_op);
  L.offsets = at; at += (((size_t)n_lists + 1) * 4 + 15) & ~(size_t)15;
  L.frames = at; at += ((size_t)n_frames * 4 + 15) & ~(size_t)15;
  L.total = at;
  return L;
}

}  // namespace

}  // namespace aiko

using namespace aiko;

extern "C" size_t aiko_b200_overlay_boxes_workspace_bytes(int n_lists, int max_objects, int n_frames) {
  if (n_lists <= 0 || max_objects <= 0 || n_frames <= 0) return 0;
  return boxes_layout(n_lists, max_objects, n_frames).total;
}

extern "C" int aiko_b200_overlay_boxes_u8(uint8_t* frames, int n_frames, int height, int width, int64_t frame_stride,
                                          int layout, const int32_t* boxes, const int32_t* class_ids,
                                          const float* confidences, int n_lists, int max_objects, const char* names,
                                          int n_names, int name_stride, const uint16_t* patch_ids, const uint8_t* luts,
                                          int n_patches, const uint64_t* variant_keys, const int32_t* variant_firsts,
                                          const int16_t* variant_boxes, int n_variants, void* workspace,
                                          size_t workspace_bytes, int32_t* status, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && height > 0 && width > 0 && (layout == 0 || layout == 1), "overlay_boxes: bad size / layout");
  if (n_frames == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(n_lists == 1 || n_lists == n_frames, "overlay_boxes: %d lists for %d frames (1 or one per frame)", n_lists,
                 n_frames);
  AIKO_CHECK_ARG(max_objects > 0 && n_names >= 0 && name_stride > 0, "overlay_boxes: bad object / name counts");
  AIKO_CHECK_ARG(frames && boxes && class_ids && confidences && workspace && status && patch_ids && luts && (names || n_names == 0),
                 "overlay_boxes: null pointer");
  AIKO_CHECK_ARG(frame_stride >= (int64_t)height * width * 3, "overlay_boxes: frame stride smaller than a frame");
  AIKO_CHECK_ARG(n_variants == 0 || (variant_keys && variant_firsts && variant_boxes), "overlay_boxes: null variant table");
  AIKO_CHECK_ARG(n_patches >= 1 + 95 * 4 * 2 && n_variants >= 0, "overlay_boxes: the atlas must hold the interior glyphs");
  AIKO_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "overlay_boxes: workspace must be 16-byte aligned");
  const BoxesLayout L = boxes_layout(n_lists, max_objects, n_frames);
  if (workspace_bytes < L.total) {
    set_last_error("overlay_boxes: workspace of %zu bytes, need %zu", workspace_bytes, L.total);
    return AIKO_B200_ERR_WORKSPACE;
  }
  static GlyphMetrics metrics;                           // the interior glyphs never change: fetched once per process
  static std::once_flag metrics_once;
  static int metrics_rc = AIKO_B200_OK;
  std::call_once(metrics_once, [] {
    uint8_t advance[95];
    int16_t glyph_boxes[95 * 4 * 4];
    int32_t constants[8];
    metrics_rc = aiko_b200_overlay_glyph_metrics_host(advance, glyph_boxes, constants);
    if (metrics_rc != AIKO_B200_OK) return;
    memset(&metrics, 0, sizeof(metrics));
    memcpy(metrics.advance, advance, 95);
    for (int i = 0; i < 95 * 4; i++)
      metrics.box[i] = make_short4(glyph_boxes[4 * i], glyph_boxes[4 * i + 1], glyph_boxes[4 * i + 2], glyph_boxes[4 * i + 3]);
    metrics.patch_w = constants[0]; metrics.patch_h = constants[1]; metrics.pad_x = constants[2]; metrics.pad_y = constants[3];
    metrics.margin = constants[4]; metrics.text_height = constants[5]; metrics.thickness = constants[6];
  });
  if (metrics_rc != AIKO_B200_OK) return metrics_rc;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  BoxesParams p{};
  p.boxes = boxes; p.class_ids = class_ids; p.confidences = confidences; p.names = names;
  p.n_lists = n_lists; p.max_objects = max_objects; p.n_names = n_names; p.name_stride = name_stride;
  p.W = width; p.H = height; p.layout = layout; p.n_frames = n_frames;
  p.variant_keys = reinterpret_cast<const unsigned long long*>(variant_keys);
  p.variant_firsts = variant_firsts;
  p.variant_boxes = reinterpret_cast<const short4*>(variant_boxes);
  p.n_variants = n_variants;
  p.ops = reinterpret_cast<aiko_b200_overlay_op*>(ws + L.ops);
  p.ops_capacity = L.ops_capacity;
  p.op_offsets = reinterpret_cast<int32_t*>(ws + L.offsets);
  p.list_of_frame = reinterpret_cast<int32_t*>(ws + L.frames);
  p.status = status;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int total_objects = n_lists * max_objects;
  const int threads = total_objects <= 32 ? 32 : total_objects <= 128 ? 128 : kPlanThreads;     // one CTA; small calls: fewer warps at the barriers
  overlay_boxes_plan_kernel<<<1, threads, 0, st>>>(p, metrics);
  AIKO_CUDA(cudaGetLastError());
  aiko_b200_overlay ov;
  memset(&ov, 0, sizeof(ov));
  ov.ops = p.ops; ov.op_offsets = p.op_offsets; ov.list_of_frame = p.list_of_frame;
  ov.patch_ids = patch_ids; ov.luts = luts;
  ov.patch_w = metrics.patch_w; ov.patch_h = metrics.patch_h;
  ov.n_patches = n_patches;
  ov.n_lists = n_lists; ov.n_frames = n_frames;
  ov.frame_width = width; ov.frame_height = height;      // op counts are only known on the device: the tile kernel culls the lists
  return aiko_b200_overlay_u8(frames, frames, n_frames, height, width, 3, (int64_t)width * 3, frame_stride, &ov, stream);
}
