// Marks of a draw-list op: the axis-aligned blocks of 4-pixel groups (pixels 4g .. 4g+3 of a row)
// that the op can touch, clipped to the frame.  A glyph has one (its bounding box); a
// cv2.rectangle(thickness 2) outline has up to four (top band, bottom band, two side bands), so the
// untouched interior of a box costs nothing.  The same function runs on the host (planner: cell
// counts for the launch grid) and on the device (overlay_cells_kernel: one thread per cell).
#pragma once
#include "../../include/aiko_b200.h"

#if defined(__CUDACC__)
#define AIKO_HD __host__ __device__ __forceinline__
#else
#define AIKO_HD inline
#endif

namespace aiko {

struct OverlayMark {
  short y0, y1;      // rows, inclusive
  short g0, g1;      // groups, inclusive
};

constexpr int kMaxMarksPerOp = 4;
constexpr int kMarkLimit = 32766;      // frames up to 32766 rows / groups fit the 16-bit mark fields

// The four mark slots of an op: slot k is in use when bit k of `valid` is set.  (Fixed slots and
// compile-time indices keep the marks in registers on the device.)
struct OverlayMarks4 {
  OverlayMark m[kMaxMarksPerOp];
  unsigned valid;
};

AIKO_HD void overlay_set_mark(OverlayMarks4& r, int slot, int ya, int yb, int xa, int xb, int W, int H) {
  ya = ya < 0 ? 0 : ya;
  yb = yb > H - 1 ? H - 1 : yb;
  xa = xa < 0 ? 0 : xa;
  xb = xb > W - 1 ? W - 1 : xb;
  if (ya > yb || xa > xb) return;
  r.m[slot].y0 = (short)ya; r.m[slot].y1 = (short)yb;
  r.m[slot].g0 = (short)(xa >> 2); r.m[slot].g1 = (short)(xb >> 2);
  r.valid |= 1u << slot;
}

AIKO_HD OverlayMarks4 overlay_op_marks4(int kind, int x0, int y0, int x1, int y1, int W, int H) {
  OverlayMarks4 r;
  r.valid = 0;
  if (kind == AIKO_B200_OP_GLYPH) {
    overlay_set_mark(r, 0, y0, y1, x0, x1, W, H);
    return r;
  }
  // outline: [x0-1 .. x1+1] x [y0-1 .. y1+1] minus the interior [x0+2 .. x1-2] x [y0+2 .. y1-2]
  if (x1 - 2 < x0 + 2 || y1 - 2 < y0 + 2) {
    overlay_set_mark(r, 0, y0 - 1, y1 + 1, x0 - 1, x1 + 1, W, H);
    return r;
  }
  overlay_set_mark(r, 0, y0 - 1, y0 + 1, x0 - 1, x1 + 1, W, H);
  overlay_set_mark(r, 1, y1 - 1, y1 + 1, x0 - 1, x1 + 1, W, H);
  const int lg1 = (x0 + 1 > W - 1 ? W - 1 : x0 + 1) >> 2;          // last group of the left band
  const int rg0 = (x1 - 1 < 0 ? 0 : x1 - 1) >> 2;                  // first group of the right band
  if (rg0 <= lg1) {
    overlay_set_mark(r, 2, y0 + 2, y1 - 2, x0 - 1, x1 + 1, W, H);  // the side bands share a group
  } else {
    overlay_set_mark(r, 2, y0 + 2, y1 - 2, x0 - 1, x0 + 1, W, H);
    overlay_set_mark(r, 3, y0 + 2, y1 - 2, x1 - 1, x1 + 1, W, H);
  }
  return r;
}

// -> number of marks written to out[0 .. 4)
AIKO_HD int overlay_op_marks(int kind, int x0, int y0, int x1, int y1, int W, int H, OverlayMark* out) {
  const OverlayMarks4 r = overlay_op_marks4(kind, x0, y0, x1, y1, W, H);
  int n = 0;
  for (int k = 0; k < kMaxMarksPerOp; k++)
    if (r.valid >> k & 1) out[n++] = r.m[k];
  return n;
}

AIKO_HD int overlay_mark_cells(const OverlayMark& m) { return (m.y1 - m.y0 + 1) * (m.g1 - m.g0 + 1); }

}  // namespace aiko
