// Host side of ImageOverlay (R/elements/media/image_io.py:119-155): turns the rectangles and label
// strings of a call into the draw list the overlay kernels replay, and owns the glyph atlas.
//
//   * label size and placement: cv2.getTextSize arithmetic + the reference's rule (image_io.py:140-150);
//   * one GLYPH op per character, referring to a patch of per-pixel transfer-table ids;
//   * patches come from the rasteriser in text_raster.hpp (exact restatement of cv2.putText):
//     a character only ever lands on four sub-pixel pen phases (the pen advances in multiples of
//     0.75 px), so the interior atlas is 95 characters x 4 phases, built once at first use;
//     characters whose strokes come within kBorderMargin pixels of the image border are rasterised
//     against that border geometry (cv2 clips stroke segments at the image rectangle, which changes
//     their anti-aliasing) and cached by where the borders lie relative to the patch;
//   * per pixel the rasteriser yields an ordered sequence of "blend with alpha a" / "store colour"
//     events; a sequence is composed once into a 256-entry table per colour component (0 and 255,
//     image_io.py:101) and de-duplicated by content.
// No OpenCV at run time, no device code: plain C++ linked into libaiko_b200.so.
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "../../include/aiko_b200.h"
#include "overlay_marks.h"
#include "text_raster.hpp"

namespace aiko {
void set_last_error(const char* fmt, ...);
}

namespace {

using aiko_text::P2;
using aiko_text::XY_ONE;
using aiko_text::XY_SHIFT;

constexpr int kPatchW = 40, kPatchH = 48;      // patch window: origin = (floor(pen_x) - kPadX, baseline - kPadY)
constexpr int kPadX = 8, kPadY = 40;
constexpr int kBorderMargin = 6;
constexpr int kComponents[2] = {0, 255};
constexpr int kMaxPatches = 4096;
constexpr int kMaxLuts = 65535;
constexpr int kFar = 1 << 20;
constexpr double kFontScale = 0.75;            // image_io.py:103
constexpr int kThickness = 2;                  // image_io.py:104
constexpr int kTextHeight = 17;                // cvRound((12 + 9) * 0.75 + (2 + 1) / 2): cap + base line of the face

// ---- sink: per-pixel event sequences inside the patch window -----------------------------------
struct SeqSink {
  int W, H;                 // image size = clip rectangle of the rasteriser
  int wx, wy;               // window origin in image coordinates
  std::vector<std::string> seq;     // kPatchW * kPatchH
  bool escaped = false;     // something was drawn outside the window
  SeqSink(int w, int h, int ox, int oy) : W(w), H(h), wx(ox), wy(oy), seq(kPatchW * kPatchH) {}
  int width() const { return W; }
  int height() const { return H; }
  void blend(int x, int y, int a) {
    if (a == 0) return;                                  // ((c - d) * 0 + 127) >> 8 == 0: identity
    int px = x - wx, py = y - wy;
    if ((unsigned)px >= (unsigned)kPatchW || (unsigned)py >= (unsigned)kPatchH) { escaped = true; return; }
    seq[py * kPatchW + px].push_back((char)a);
  }
  void fill(int y, int x0, int x1) {
    int py = y - wy;
    for (int x = x0; x <= x1; x++) {
      int px = x - wx;
      if ((unsigned)px >= (unsigned)kPatchW || (unsigned)py >= (unsigned)kPatchH) { escaped = true; continue; }
      std::string& s = seq[py * kPatchW + px];
      s.assign(1, '\0');                                 // a store forgets everything before it
    }
  }
};

struct Box {
  int x0 = 1, y0 = 1, x1 = 0, y1 = 0;          // empty: x1 < x0
  bool empty() const { return x1 < x0; }
};

struct Atlas {
  std::vector<uint16_t> patch_ids;             // [n_patches][kPatchH][kPatchW]
  std::vector<Box> bbox;                       // per patch
  std::vector<uint8_t> luts;                   // [n_luts][256], row 0 = identity
  std::unordered_map<std::string, uint16_t> lut_by_content;
  std::unordered_map<std::string, std::pair<uint16_t, uint16_t>> lut_by_sequence;   // -> ids for components 0, 255
  std::map<std::tuple<int, int, int, int, int, int>, int> variants;                 // (char, phase, l, t, r, b) -> first patch
  int n_interior_patches = 0, n_interior_luts = 0;
  uint64_t version = 0;                        // grows with every append
  uint64_t epoch = 0;                          // bumped when border variants are dropped (indices reused)

  int n_patches() const { return (int)(patch_ids.size() / (kPatchW * kPatchH)); }
  int n_luts() const { return (int)(luts.size() / 256); }

  int lut_id(const uint8_t* table) {
    std::string key((const char*)table, 256);
    auto it = lut_by_content.find(key);
    if (it != lut_by_content.end()) return it->second;
    if (n_luts() >= kMaxLuts) return -1;
    uint16_t id = (uint16_t)n_luts();
    luts.insert(luts.end(), table, table + 256);
    lut_by_content.emplace(std::move(key), id);
    return id;
  }

  // ids of the two transfer tables (component 0, component 255) of an event sequence
  bool sequence_ids(const std::string& s, uint16_t out[2]) {
    auto it = lut_by_sequence.find(s);
    if (it != lut_by_sequence.end()) { out[0] = it->second.first; out[1] = it->second.second; return true; }
    for (int k = 0; k < 2; k++) {
      const int c = kComponents[k];
      uint8_t table[256];
      for (int d0 = 0; d0 < 256; d0++) {
        int d = d0;
        for (unsigned char a : s) {
          if (a == 0) { d = c; continue; }
          d += ((c - d) * a + 127) >> 8;
          d += ((c - d) * a + 127) >> 8;
        }
        table[d0] = (uint8_t)d;
      }
      int id = lut_id(table);
      if (id < 0) return false;
      out[k] = (uint16_t)id;
    }
    lut_by_sequence.emplace(s, std::make_pair(out[0], out[1]));
    return true;
  }

  // Appends the two patches (component 0, component 255) of a rasterised glyph; -> first index or -1.
  int add_glyph(const SeqSink& sink) {
    if (n_patches() + 2 > kMaxPatches) return -1;
    const int first = n_patches();
    patch_ids.resize(patch_ids.size() + 2 * kPatchW * kPatchH, 0);
    bbox.resize(bbox.size() + 2);
    for (int i = 0; i < kPatchW * kPatchH; i++) {
      if (sink.seq[i].empty()) continue;
      uint16_t ids[2];
      if (!sequence_ids(sink.seq[i], ids)) {
        patch_ids.resize((size_t)first * kPatchW * kPatchH);
        bbox.resize(first);
        return -1;
      }
      for (int k = 0; k < 2; k++) {
        if (!ids[k]) continue;
        patch_ids[(size_t)(first + k) * kPatchW * kPatchH + i] = ids[k];
        Box& b = bbox[first + k];
        const int x = i % kPatchW, y = i / kPatchW;
        if (b.empty()) b = Box{x, y, x, y};
        else { b.x0 = std::min(b.x0, x); b.y0 = std::min(b.y0, y); b.x1 = std::max(b.x1, x); b.y1 = std::max(b.y1, y); }
      }
    }
    version++;
    return first;
  }

  void drop_variants() {
    patch_ids.resize((size_t)n_interior_patches * kPatchW * kPatchH);
    bbox.resize(n_interior_patches);
    luts.resize((size_t)n_interior_luts * 256);
    for (auto it = lut_by_content.begin(); it != lut_by_content.end();)
      it = (it->second >= n_interior_luts) ? lut_by_content.erase(it) : std::next(it);
    for (auto it = lut_by_sequence.begin(); it != lut_by_sequence.end();)
      it = (it->second.first >= n_interior_luts || it->second.second >= n_interior_luts) ? lut_by_sequence.erase(it)
                                                                                         : std::next(it);
    variants.clear();
    version++;
    epoch++;
  }
};

Atlas* g_atlas = nullptr;
std::mutex g_mutex;

const int kHScale = aiko_text::cv_round(kFontScale * XY_ONE);

// Rasterise character c at sub-pixel phase q (quarter pixels) into a window whose origin sits at
// (ox, oy) of an image W x H: pen floor = ox + kPadX, base line = oy + kPadY.
void raster_glyph(SeqSink& sink, unsigned char c, int q) {
  const int64_t view_x = ((int64_t)(sink.wx + kPadX) << XY_SHIFT) + (int64_t)q * (XY_ONE / 4);
  aiko_text::draw_glyph(sink, c, view_x, sink.wy + kPadY, kHScale);
}

bool build_interior(Atlas& a) {
  a.patch_ids.reserve((size_t)kMaxPatches * kPatchW * kPatchH);
  a.luts.reserve((size_t)kMaxLuts * 256);
  a.bbox.reserve(kMaxPatches);
  uint8_t identity[256];
  for (int i = 0; i < 256; i++) identity[i] = (uint8_t)i;
  a.lut_id(identity);                                          // id 0
  a.patch_ids.resize((size_t)kPatchW * kPatchH, 0);            // patch 0: nothing
  a.bbox.resize(1);
  const int slack = 64;
  for (int ci = 0; ci < 95; ci++)
    for (int q = 0; q < 4; q++) {
      SeqSink sink(kPatchW + 2 * slack, kPatchH + 2 * slack, slack, slack);
      raster_glyph(sink, (unsigned char)(32 + ci), q);
      if (sink.escaped) { aiko::set_last_error("glyph atlas: character %d leaves its patch window", 32 + ci); return false; }
      if (a.add_glyph(sink) != 1 + (ci * 4 + q) * 2) { aiko::set_last_error("glyph atlas: capacity"); return false; }
    }
  a.n_interior_patches = a.n_patches();
  a.n_interior_luts = a.n_luts();
  return true;
}

Atlas* atlas() {          // call with g_mutex held
  if (!g_atlas) {
    Atlas* a = new Atlas();
    if (!build_interior(*a)) { delete a; return nullptr; }
    g_atlas = a;
  }
  return g_atlas;
}

inline int interior_patch(unsigned char c, int q, int k) { return 1 + (aiko_text::glyph_index(c) * 4 + q) * 2 + k; }

// Patch pair for a glyph whose window origin is (px, py) in a W x H image.  Interior glyphs use the
// atlas built at start-up; glyphs near a border get (and cache) a variant rasterised against it.
int glyph_patches(Atlas& a, unsigned char c, int q, int px, int py, int W, int H) {
  const int base = interior_patch(c, q, 0);
  Box u = a.bbox[base];
  const Box& b1 = a.bbox[base + 1];
  if (u.empty()) u = b1;
  else if (!b1.empty()) { u.x0 = std::min(u.x0, b1.x0); u.y0 = std::min(u.y0, b1.y0); u.x1 = std::max(u.x1, b1.x1); u.y1 = std::max(u.y1, b1.y1); }
  if (u.empty()) return base;                                   // blank
  const int m = kBorderMargin;
  if (px + u.x0 >= m && py + u.y0 >= m && px + u.x1 < W - m && py + u.y1 < H - m) return base;
  const int left = (px + u.x0 < m) ? px : -kFar, top = (py + u.y0 < m) ? py : -kFar;
  const int right = (px + u.x1 >= W - m) ? W - px : kFar, bottom = (py + u.y1 >= H - m) ? H - py : kFar;
  auto key = std::make_tuple((int)aiko_text::glyph_index(c), q, left, top, right, bottom);
  auto it = a.variants.find(key);
  if (it != a.variants.end()) return it->second;
  // canonical geometry: borders that are far stay far (slack), near ones at their true distance
  const int slack = kPatchW + 64;
  const int ox = (left > -kFar) ? left : slack, oy = (top > -kFar) ? top : slack;      // window origin in the image
  const int iw = ox + ((right < kFar) ? right : kPatchW + slack), ih = oy + ((bottom < kFar) ? bottom : kPatchH + slack);
  int first;
  if (iw <= 0 || ih <= 0) {
    first = 0;                                                   // nothing of the image under the window
  } else {
    SeqSink sink(iw, ih, ox, oy);
    raster_glyph(sink, c, q);
    first = a.add_glyph(sink);
    if (first < 0) return -1;
  }
  a.variants.emplace(key, first);
  return first;
}

void text_size(const char* text, int* w, int* h) {
  long units = 0;
  for (const unsigned char* p = (const unsigned char*)text; *p; p++) units += aiko_text::glyph_advance(*p);
  *w = aiko_text::cv_round(units * kFontScale + kThickness);
  *h = kTextHeight;
}

// Marks of an op (overlay_marks.h) -> their number; adds their cell count.
int op_marks(const aiko_b200_overlay_op& op, int W, int H, long long* cells) {
  aiko::OverlayMark m[aiko::kMaxMarksPerOp];
  const int n = aiko::overlay_op_marks(op.kind, op.x0, op.y0, op.x1, op.y1, W, H, m);
  for (int i = 0; i < n; i++) *cells += aiko::overlay_mark_cells(m[i]);
  return n;
}

}  // namespace

// What a planner needs to know about the 95 x 4 interior glyphs, for the device-side planner of
// aiko_b200_overlay_boxes_u8 (csrc/overlay_boxes.cu): advance per character in glyph units, and per
// (character, pen phase) the inclusive box x0, y0, x1, y1 of the pixels its two patches touch inside the patch
// window (x1 < x0: blank), plus the planner's constants.
extern "C" int aiko_b200_overlay_glyph_metrics_host(uint8_t* advance_out_host /* [95] */, int16_t* boxes_out_host /* [95*4][4] */,
                                                    int32_t* constants_out_host /* [8] */) {
  if (!advance_out_host || !boxes_out_host || !constants_out_host) {
    aiko::set_last_error("glyph_metrics: null argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  std::lock_guard<std::mutex> lock(g_mutex);
  Atlas* a = atlas();
  if (!a) return AIKO_B200_ERR_UNSUPPORTED;
  for (int ci = 0; ci < 95; ci++) {
    advance_out_host[ci] = (uint8_t)aiko_text::glyph_advance((unsigned char)(32 + ci));
    for (int q = 0; q < 4; q++) {
      const int base = interior_patch((unsigned char)(32 + ci), q, 0);
      Box u = a->bbox[base];
      const Box& b1 = a->bbox[base + 1];
      if (u.empty()) u = b1;
      else if (!b1.empty()) { u.x0 = std::min(u.x0, b1.x0); u.y0 = std::min(u.y0, b1.y0); u.x1 = std::max(u.x1, b1.x1); u.y1 = std::max(u.y1, b1.y1); }
      int16_t* o = boxes_out_host + (ci * 4 + q) * 4;
      o[0] = (int16_t)u.x0; o[1] = (int16_t)u.y0; o[2] = (int16_t)u.x1; o[3] = (int16_t)u.y1;
    }
  }
  const int32_t constants[8] = {kPatchW, kPatchH, kPadX, kPadY, kBorderMargin, kTextHeight, kThickness, 0};
  std::memcpy(constants_out_host, constants, sizeof(constants));
  return AIKO_B200_OK;
}

// The border variants rasterised so far, for the device-side planner: per variant a 64-bit key -- glyph slot, pen
// phase and where the image borders lie relative to the patch window, packed so that unsigned order is the order
// of the planner's own map (slot << 34 | phase << 32 | left << 24 | top << 16 | right << 8 | bottom, every border
// distance biased by 128, "far" = 0 for left / top and 255 for right / bottom) --, its first patch (0: nothing of
// the glyph lands on the image) and the inclusive box of the pixels its patches touch.  Sorted by key.
// -> number of variants (only the first `capacity` are written).
extern "C" int aiko_b200_overlay_variants_host(uint64_t* keys_out_host, int32_t* firsts_out_host, int16_t* boxes_out_host,
                                               int capacity, uint64_t* version_out_host, uint64_t* epoch_out_host) {
  std::lock_guard<std::mutex> lock(g_mutex);
  Atlas* a = atlas();
  if (!a) return AIKO_B200_ERR_UNSUPPORTED;
  if (version_out_host) *version_out_host = a->version;
  if (epoch_out_host) *epoch_out_host = a->epoch;
  int n = 0;
  for (const auto& kv : a->variants) {
    const int slot = std::get<0>(kv.first), q = std::get<1>(kv.first);
    const int l = std::get<2>(kv.first), t = std::get<3>(kv.first), r = std::get<4>(kv.first), b = std::get<5>(kv.first);
    const bool lf = l <= -kFar, tf = t <= -kFar, rf = r >= kFar, bf = b >= kFar;
    if ((!lf && (l < -127 || l > 126)) || (!tf && (t < -127 || t > 126)) || (!rf && (r < -127 || r > 126)) ||
        (!bf && (b < -127 || b > 126)))
      continue;                                            // (not representable in the key: the device asks the host)
    if (n < capacity && keys_out_host && firsts_out_host && boxes_out_host) {
      const uint64_t L = lf ? 0 : (uint64_t)(l + 128), T = tf ? 0 : (uint64_t)(t + 128);
      const uint64_t R = rf ? 255 : (uint64_t)(r + 128), B = bf ? 255 : (uint64_t)(b + 128);
      keys_out_host[n] = ((uint64_t)slot << 34) | ((uint64_t)q << 32) | (L << 24) | (T << 16) | (R << 8) | B;
      const int first = kv.second;
      firsts_out_host[n] = first;
      Box u;
      if (first > 0) {
        u = a->bbox[first];
        const Box& b1 = a->bbox[first + 1];
        if (u.empty()) u = b1;
        else if (!b1.empty()) { u.x0 = std::min(u.x0, b1.x0); u.y0 = std::min(u.y0, b1.y0); u.x1 = std::max(u.x1, b1.x1); u.y1 = std::max(u.y1, b1.y1); }
      }
      boxes_out_host[4 * n] = (int16_t)u.x0; boxes_out_host[4 * n + 1] = (int16_t)u.y0;
      boxes_out_host[4 * n + 2] = (int16_t)u.x1; boxes_out_host[4 * n + 3] = (int16_t)u.y1;
    }
    n++;
  }
  return n;
}

extern "C" int aiko_b200_overlay_text_size_host(const char* text_host, int* width_host, int* height_host) {
  if (!text_host || !width_host || !height_host) { aiko::set_last_error("text_size: null argument"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
  text_size(text_host, width_host, height_host);
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_overlay_atlas_host(aiko_b200_atlas_view* out) {
  if (!out) { aiko::set_last_error("atlas: null argument"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
  std::lock_guard<std::mutex> lock(g_mutex);
  Atlas* a = atlas();
  if (!a) return AIKO_B200_ERR_UNSUPPORTED;
  out->patch_ids_host = a->patch_ids.data();
  out->luts_host = a->luts.data();
  out->patch_w = kPatchW; out->patch_h = kPatchH;
  out->n_patches = a->n_patches(); out->n_luts = a->n_luts();
  out->max_patches = kMaxPatches; out->max_luts = kMaxLuts;
  out->version = a->version; out->epoch = a->epoch;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_overlay_cells_host(const aiko_b200_overlay_op* ops_host, int n_ops, int image_width,
                                            int image_height, int* n_marks_host, long long* n_cells_host) {
  if ((!ops_host && n_ops > 0) || n_ops < 0 || !n_marks_host || !n_cells_host) {
    aiko::set_last_error("overlay_cells: bad argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  int marks = 0;
  long long cells = 0;
  for (int i = 0; i < n_ops; i++) marks += op_marks(ops_host[i], image_width, image_height, &cells);
  *n_marks_host = marks;
  *n_cells_host = cells;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_overlay_marks_host(const aiko_b200_overlay_op* op_host, int image_width, int image_height,
                                            int16_t* marks_out_host) {
  if (!op_host || !marks_out_host) { aiko::set_last_error("overlay_marks: null argument"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
  aiko::OverlayMark m[aiko::kMaxMarksPerOp];
  const int n = aiko::overlay_op_marks(op_host->kind, op_host->x0, op_host->y0, op_host->x1, op_host->y1, image_width,
                                       image_height, m);
  for (int i = 0; i < n; i++) {
    marks_out_host[4 * i] = m[i].y0; marks_out_host[4 * i + 1] = m[i].y1;
    marks_out_host[4 * i + 2] = m[i].g0; marks_out_host[4 * i + 3] = m[i].g1;
  }
  return n;
}

namespace {

struct PlanArgs {
  int n_lists;
  const int32_t* object_offsets;
  const int32_t* rects;
  const char* labels;
  int W, H, layout;
  aiko_b200_overlay_op* ops;
  int capacity;
  int32_t* op_offsets;
  aiko_b200_overlay_plan_info* info;
};

// One pass over the call.  -> 0 done, 1 the atlas ran out of room (caller drops the border variants
// and runs the pass again).
int plan_pass(Atlas& a, const PlanArgs& g) {
  // colour (0, 255, 255) is drawn on the BGR copy of the frame (image_io.py:101,117): stored-channel
  // components are (255, 255, 0) for frames kept RGB (layout 0) and (0, 255, 255) for BGR (layout 1)
  const int comp[3] = {g.layout == 0 ? 1 : 0, 1, g.layout == 0 ? 0 : 1};         // index into kComponents
  const uint32_t rect_colour = g.layout == 0 ? 0x00ffffu : 0xffff00u;            // byte c = stored channel c
  const int W = g.W, H = g.H;
  const char* label = g.labels;
  int n_ops = 0, max_ops = 0, max_marks = 0;
  long long max_cells = 0;
  auto emit = [&](const aiko_b200_overlay_op& op, int* marks, long long* cells) {
    if (n_ops < g.capacity) g.ops[n_ops] = op;
    *marks += op_marks(op, W, H, cells);
    n_ops++;
  };
  for (int li = 0; li < g.n_lists; li++) {
    g.op_offsets[li] = n_ops;
    int marks = 0;
    long long cells = 0;
    const int list_begin = n_ops;
    for (int oi = g.object_offsets[li]; oi < g.object_offsets[li + 1]; oi++) {
      const int x = g.rects[4 * oi], y = g.rects[4 * oi + 1], w = g.rects[4 * oi + 2], h = g.rects[4 * oi + 3];
      const char* text = label ? label : "";
      if (label) label += std::strlen(label) + 1;
      if (*text) {
        int tw, th;                                     // image_io.py:140-150
        text_size(text, &tw, &th);
        int tx = x, ty;
        if (y > th * 1.5) ty = (int)(y - th / 2.0 - 2);
        else ty = y + h + th + 2;
        if (x + tw > W) { tx = x - tw; ty = y + th; }
        long units = 0;
        for (const unsigned char* p = (const unsigned char*)text; *p; p++) {
          const long pen_q = 3 * units;                 // pen position in quarter pixels
          units += aiko_text::glyph_advance(*p);
          const int q = (int)(pen_q & 3);
          const long long pxl = (long long)tx + (pen_q >> 2) - kPadX;
          const long long pyl = (long long)ty - kPadY;
          if (pxl > W || pxl + kPatchW < 0 || pyl > H || pyl + kPatchH < 0) continue;   // window off the image
          const int px = (int)pxl, py = (int)pyl;
          const int first = glyph_patches(a, *p, q, px, py, W, H);
          if (first < 0) return 1;
          if (first == 0) continue;
          Box u = a.bbox[first];
          const Box& b1 = a.bbox[first + 1];
          if (u.empty()) u = b1;
          else if (!b1.empty()) { u.x0 = std::min(u.x0, b1.x0); u.y0 = std::min(u.y0, b1.y0); u.x1 = std::max(u.x1, b1.x1); u.y1 = std::max(u.y1, b1.y1); }
          if (u.empty()) continue;
          const int cx0 = std::max(px + u.x0, 0), cy0 = std::max(py + u.y0, 0);
          const int cx1 = std::min(px + u.x1, W - 1), cy1 = std::min(py + u.y1, H - 1);
          if (cx1 < cx0 || cy1 < cy0) continue;
          aiko_b200_overlay_op op;
          op.kind = AIKO_B200_OP_GLYPH;
          op.x0 = cx0; op.y0 = cy0; op.x1 = cx1; op.y1 = cy1;
          op.p0 = (uint32_t)(first + comp[0]) | ((uint32_t)(first + comp[1]) << 16);
          op.p1 = (uint32_t)(first + comp[2]);
          op.p2 = (uint32_t)(cx0 - px) | ((uint32_t)(cy0 - py) << 16);
          emit(op, &marks, &cells);
        }
      }
      const long long x1l = (long long)x + w, y1l = (long long)y + h;       // image_io.py:154-155
      long long xa = std::min<long long>(x, x1l), xb = std::max<long long>(x, x1l);
      long long ya = std::min<long long>(y, y1l), yb = std::max<long long>(y, y1l);
      if (xb + 1 < 0 || yb + 1 < 0 || xa - 1 >= W || ya - 1 >= H) continue;
      const long long lim = 1 << 24;             // keeps +-2 arithmetic on the device inside int32
      xa = std::max(-lim, std::min(lim, xa)); xb = std::max(-lim, std::min(lim, xb));
      ya = std::max(-lim, std::min(lim, ya)); yb = std::max(-lim, std::min(lim, yb));
      aiko_b200_overlay_op op;
      op.kind = AIKO_B200_OP_RECT;
      op.x0 = (int)xa; op.y0 = (int)ya; op.x1 = (int)xb; op.y1 = (int)yb;
      op.p0 = rect_colour; op.p1 = 0; op.p2 = 0;
      emit(op, &marks, &cells);
    }
    max_ops = std::max(max_ops, n_ops - list_begin);
    max_marks = std::max(max_marks, marks);
    max_cells = std::max(max_cells, cells);
  }
  g.op_offsets[g.n_lists] = n_ops;
  g.info->n_ops = n_ops;
  g.info->max_ops_per_list = max_ops;
  g.info->max_marks_per_list = max_marks;
  g.info->max_cells_per_list = max_cells > 0x7fffffff ? 0x7fffffff : (int32_t)max_cells;
  g.info->atlas_version = a.version;
  g.info->atlas_epoch = a.epoch;
  return 0;
}

}  // namespace

extern "C" int aiko_b200_overlay_plan_host(int n_lists, const int32_t* object_offsets_host, const int32_t* rects_host,
                                           const char* labels_host, int image_width, int image_height, int layout,
                                           aiko_b200_overlay_op* ops_out_host, int ops_capacity,
                                           int32_t* op_offsets_out_host, aiko_b200_overlay_plan_info* info_host) {
  if (n_lists < 0 || !object_offsets_host || !op_offsets_out_host || !info_host || (ops_capacity > 0 && !ops_out_host) ||
      ops_capacity < 0 || image_width <= 0 || image_height <= 0 || (layout != 0 && layout != 1) ||
      (n_lists > 0 && object_offsets_host[n_lists] > 0 && !rects_host)) {
    aiko::set_last_error("overlay_plan: bad argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  std::lock_guard<std::mutex> lock(g_mutex);
  Atlas* a = atlas();
  if (!a) return AIKO_B200_ERR_UNSUPPORTED;
  const PlanArgs g{n_lists, object_offsets_host, rects_host, labels_host, image_width, image_height, layout,
                   ops_out_host, ops_capacity, op_offsets_out_host, info_host};
  if (plan_pass(*a, g) != 0) {
    a->drop_variants();               // atlas full: forget the cached border variants and start over
    if (plan_pass(*a, g) != 0) {
      aiko::set_last_error("overlay_plan: glyph atlas capacity exceeded by one call");
      return AIKO_B200_ERR_UNSUPPORTED;
    }
  }
  if (info_host->n_ops > ops_capacity) {
    aiko::set_last_error("overlay_plan: %d ops do not fit the output buffer (%d)", info_host->n_ops, ops_capacity);
    return AIKO_B200_ERR_WORKSPACE;
  }
  return AIKO_B200_OK;
}

// Plans straight into a caller-owned page-locked buffer and fills the aiko_b200_overlay that the
// drawing call takes: ops | op_offsets | list_of_frame laid out in buffer_host, device pointers set
// to the same offsets inside buffer_dev, upload_from_host / upload_bytes set so that the drawing call
// ships the bytes itself.  One call per new overlay on the per-frame path.
extern "C" int aiko_b200_overlay_plan_struct_host(int n_lists, const int32_t* object_offsets_host,
                                                  const int32_t* rects_host, const char* labels_host, int image_width,
                                                  int image_height, int layout, int n_frames,
                                                  const int32_t* list_of_frame_host, void* buffer_host,
                                                  size_t buffer_bytes, void* buffer_dev, aiko_b200_overlay* overlay_host,
                                                  aiko_b200_overlay_plan_info* info_host) {
  if (!buffer_host || !buffer_dev || !overlay_host || !info_host || n_frames < 0 || n_lists < 0 ||
      (reinterpret_cast<uintptr_t>(buffer_host) & 15) || (reinterpret_cast<uintptr_t>(buffer_dev) & 15)) {
    aiko::set_last_error("overlay_plan_struct: bad argument (buffers must be 16-byte aligned)");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  const size_t tail = sizeof(int32_t) * ((size_t)n_lists + 1 + (size_t)n_frames);
  if (buffer_bytes < tail + sizeof(aiko_b200_overlay_op)) {
    aiko::set_last_error("overlay_plan_struct: buffer of %zu bytes is too small", buffer_bytes);
    return AIKO_B200_ERR_WORKSPACE;
  }
  // the op area takes what the offsets and the frame map leave, rounded down to whole ops
  const int capacity = (int)std::min<size_t>((buffer_bytes - tail) / sizeof(aiko_b200_overlay_op), 1u << 24);
  std::vector<int32_t> offsets((size_t)n_lists + 1);
  aiko_b200_overlay_op* ops = static_cast<aiko_b200_overlay_op*>(buffer_host);
  const int rc = aiko_b200_overlay_plan_host(n_lists, object_offsets_host, rects_host, labels_host, image_width,
                                             image_height, layout, ops, capacity, offsets.data(), info_host);
  if (rc != AIKO_B200_OK) return rc;
  const size_t ops_bytes = sizeof(aiko_b200_overlay_op) * (size_t)std::max(info_host->n_ops, 1);
  int32_t* tailp = reinterpret_cast<int32_t*>(static_cast<char*>(buffer_host) + ops_bytes);
  std::memcpy(tailp, offsets.data(), sizeof(int32_t) * offsets.size());
  int32_t* frames = tailp + offsets.size();
  for (int f = 0; f < n_frames; f++) {
    const int li = list_of_frame_host ? list_of_frame_host[f] : 0;
    if (li < 0 || li >= n_lists) {
      aiko::set_last_error("overlay_plan_struct: frame %d names list %d of %d", f, li, n_lists);
      return AIKO_B200_ERR_INVALID_ARGUMENT;
    }
    frames[f] = li;
  }
  char* dev = static_cast<char*>(buffer_dev);
  overlay_host->ops = reinterpret_cast<const aiko_b200_overlay_op*>(dev);
  overlay_host->op_offsets = reinterpret_cast<const int32_t*>(dev + ops_bytes);
  overlay_host->list_of_frame = overlay_host->op_offsets + offsets.size();
  overlay_host->n_lists = n_lists;
  overlay_host->n_frames = n_frames;
  overlay_host->max_ops_per_list = info_host->max_ops_per_list;
  overlay_host->max_marks_per_list = info_host->max_marks_per_list;
  overlay_host->max_cells_per_list = info_host->max_cells_per_list;
  overlay_host->frame_width = image_width;
  overlay_host->frame_height = image_height;
  overlay_host->upload_from_host = buffer_host;
  overlay_host->upload_bytes = ops_bytes + tail;
  return AIKO_B200_OK;
}

// Plan + draw in one crossing of the ABI (the per-frame call when the overlay is new every time).
extern "C" int aiko_b200_resize_linear_objects_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                                  int64_t src_row_stride, int64_t src_frame_stride, uint8_t* dst, int dst_h,
                                                  int dst_w, int64_t dst_row_stride, int64_t dst_frame_stride, int swap_rb,
                                                  int n_lists, const int32_t* object_offsets_host,
                                                  const int32_t* rects_host, const char* labels_host, int layout,
                                                  const int32_t* list_of_frame_host, void* buffer_host, size_t buffer_bytes,
                                                  void* buffer_dev, aiko_b200_overlay* overlay_host,
                                                  aiko_b200_overlay_plan_info* info_host, uint64_t atlas_version,
                                                  uint64_t atlas_epoch, void* stream) {
  const int rc = aiko_b200_overlay_plan_struct_host(n_lists, object_offsets_host, rects_host, labels_host, dst_w, dst_h,
                                                    layout, n_frames, list_of_frame_host, buffer_host, buffer_bytes,
                                                    buffer_dev, overlay_host, info_host);
  if (rc != AIKO_B200_OK) return rc;
  if (info_host->atlas_version != atlas_version || info_host->atlas_epoch != atlas_epoch) return AIKO_B200_ATLAS_GREW;
  return aiko_b200_resize_linear_u8(src, n_frames, src_h, src_w, channels, src_row_stride, src_frame_stride, dst, dst_h,
                                    dst_w, dst_row_stride, dst_frame_stride, swap_rb,
                                    info_host->n_ops > 0 ? overlay_host : nullptr, stream);
}
