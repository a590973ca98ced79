// Device-side replay of ImageOverlay's drawing on one image row held in shared
// memory (reference: R/elements/media/image_io.py:127-155).  Used both by the
// fused resize kernel and by the stand-alone overlay kernel.
#pragma once
#include "common.cuh"

namespace aiko {


struct OverlayDev {
  const aiko_b200_overlay_op* ops;
  const int32_t* op_offsets;
  const int32_t* list_of_frame;   // or null: frame f uses list f
  const uint16_t* patch_ids;
  const uint8_t* luts;
  int patch_w, patch_h;
  int max_ops;     // host hints (include/aiko_b200.h): largest per-list op / mark / cell counts, 0 = unknown
  int max_marks, max_cells;
  int frame_w, frame_h;           // frame size max_marks / max_cells were computed for
};

// Checks an overlay given through the C ABI against the call's frame count and copies it.
// (and ships the planned buffer first when it is still in host memory).
int overlay_from_abi(const aiko_b200_overlay* overlay, int n_frames, cudaStream_t stream, OverlayDev* out);

__device__ __forceinline__ int overlay_list(const OverlayDev& ov, int frame) {
  return ov.list_of_frame ? __ldg(ov.list_of_frame + frame) : frame;
}

// cv2.rectangle(.., thickness=2) coverage, closed form.
__device__ __forceinline__ bool rect_covers(int x, int y, int xa, int ya, int xb, int yb) {
  if (x < xa - 1 || x > xb + 1 || y < ya - 1 || y > yb + 1) return false;
  if (x >= xa + 2 && x <= xb - 2 && y >= ya + 2 && y <= yb - 2) return false;
  if ((x == xa - 1 || x == xb + 1) && (y == ya - 1 || y == yb + 1)) return false;
  return true;
}

__device__ __forceinline__ void put_rgb(uint8_t* row, int x, uint32_t colour) {
  row[3 * x + 0] = colour & 0xff;
  row[3 * x + 1] = (colour >> 8) & 0xff;
  row[3 * x + 2] = (colour >> 16) & 0xff;
}

// ---------------------------------------------------------------------------------------
// Ownership: pixels 4g .. 4g+3 ("group" g) belong to worker thread g % nthreads -- the same
// mapping the resize kernels use when they produce the row -- so a thread only ever
// read-modify-writes its own pixels and the replay needs no barrier: draw order is kept per
// pixel by program order.

// Columns of row y that op can touch: two intervals [a0,a1] and [b0,b1] (equal for glyphs and
// for the top / bottom bands of a rectangle; the two 3-pixel side bands otherwise).
struct OpSpan {
  int a0, a1, b0, b1;
  bool row_hit;
};
__device__ __forceinline__ OpSpan op_span(const aiko_b200_overlay_op* op, int y) {
  OpSpan s;
  const int4 lo4 = reinterpret_cast<const int4*>(op)[0];   // kind, x0, y0, x1
  const int kind = lo4.x, x0 = lo4.y, y0 = lo4.z, x1 = lo4.w, y1 = op->y1;
  if (kind == AIKO_B200_OP_RECT) {
    s.row_hit = (y >= y0 - 1) && (y <= y1 + 1);
    const bool band = (y <= y0 + 1) || (y >= y1 - 1);
    s.a0 = x0 - 1;
    s.b1 = x1 + 1;
    s.a1 = band ? x1 + 1 : x0 + 1;
    s.b0 = band ? x0 - 1 : x1 - 1;
  } else {
    s.row_hit = (y >= y0) && (y <= y1);
    s.a0 = s.b0 = x0;
    s.a1 = s.b1 = x1;
  }
  return s;
}
__device__ __forceinline__ bool span_meets(const OpSpan& s, int lo, int hi) {
  return (s.a0 <= hi && s.a1 >= lo) || (s.b0 <= hi && s.b1 >= lo);
}

// One lane applies one op to its pixels [mlo, mhi] of row y.
__device__ __forceinline__ void apply_op_to_pixels(const aiko_b200_overlay_op* op, const OverlayDev& ov, int y,
                                                   int mlo, int mhi, uint8_t* row) {
  const int4 lo4 = reinterpret_cast<const int4*>(op)[0];   // kind, x0, y0, x1
  const int4 hi4 = reinterpret_cast<const int4*>(op)[1];   // y1, p0, p1, p2
  const int x0 = lo4.y, y0 = lo4.z, x1 = lo4.w, y1 = hi4.x;
  if (lo4.x == AIKO_B200_OP_RECT) {
    const uint32_t colour = (uint32_t)hi4.y;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int x = mlo + k;
      if (x <= mhi && rect_covers(x, y, x0, y0, x1, y1)) put_rgb(row, x, colour);
    }
  } else {
    const int xs = max(mlo, x0), xe = min(mhi, x1);
    const int psize = ov.patch_w * ov.patch_h;
    const uint32_t p0 = (uint32_t)hi4.y, p1 = (uint32_t)hi4.z, p2 = (uint32_t)hi4.w;
    const int px = x0 - int(p2 & 0xffff), py = y0 - int(p2 >> 16);
    const uint16_t* q = ov.patch_ids + (y - py) * ov.patch_w - px;
    const uint16_t* q0 = q + (size_t)(p0 & 0xffff) * psize;
    const uint16_t* q1 = q + (size_t)(p0 >> 16) * psize;
    const uint16_t* q2 = q + (size_t)(p1 & 0xffff) * psize;
    uint16_t id[4][3];
    uint8_t val[4][3];
#pragma unroll
    for (int k = 0; k < 4; k++) {          // all id look-ups first ...
      const int x = min(xs + k, xe);
      id[k][0] = __ldg(q0 + x); id[k][1] = __ldg(q1 + x); id[k][2] = __ldg(q2 + x);
    }
#pragma unroll
    for (int k = 0; k < 4; k++) {          // ... then all transfer-table look-ups
      const uint8_t* d = row + 3 * min(xs + k, xe);
#pragma unroll
      for (int c = 0; c < 3; c++) val[k][c] = id[k][c] ? __ldg(ov.luts + (size_t)id[k][c] * 256 + d[c]) : d[c];
    }
#pragma unroll
    for (int k = 0; k < 4; k++)
      if (xs + k <= xe) {
        uint8_t* d = row + 3 * (xs + k);
        d[0] = val[k][0]; d[1] = val[k][1]; d[2] = val[k][2];
      }
  }
}

// The warp applies the ops flagged in `hits` (bit b = ops32[b], in order) to the pixels its
// lanes own.  Every lane walks the list for the next op that touches ITS pixels, then the lanes
// apply their (different) ops together, so the two dependent atlas look-ups of a glyph are
// paid once per round rather than once per op.
__device__ __forceinline__ void overlay_apply_hits(const aiko_b200_overlay_op* ops32, uint32_t hits,
                                                   const OverlayDev& ov, int y, int mlo, int mhi, uint8_t* row) {
  uint32_t rem = hits;   // every lane keeps its own cursor into the hit list
  while (true) {
    const aiko_b200_overlay_op* op = nullptr;
    while (rem) {
      const int b = __ffs(rem) - 1;
      rem &= rem - 1;
      const aiko_b200_overlay_op* cand = ops32 + b;
      if (span_meets(op_span(cand, y), mlo, mhi)) { op = cand; break; }
    }
    if (!__any_sync(0xffffffffu, op != nullptr)) break;
    if (op != nullptr) apply_op_to_pixels(op, ov, y, mlo, mhi, row);
  }
}

// Self-contained replay for a CTA of `nthreads` worker threads: applies, in order, every op of
// ops[0 .. nops) that touches image row `y` to `row` (w pixels x 3 bytes, shared memory).
// Culling is warp-local (32 ops per ballot).  Callers order the row's producers before, and
// its consumers after, with __syncthreads().
__device__ __forceinline__ void overlay_apply_row(const aiko_b200_overlay_op* ops, int nops, const OverlayDev& ov,
                                                  int y, int w, uint8_t* row, int nthreads) {
  if (nops == 0) return;
  const int lane = threadIdx.x & 31;
  const int groups = (w + 3) >> 2;
  for (int gbase = (threadIdx.x & ~31); gbase < groups; gbase += nthreads) {   // warp-uniform
    const int wlo = gbase * 4, whi = min(wlo + 127, w - 1);                    // the warp's pixel span
    const int mlo = (gbase + lane) * 4, mhi = min(mlo + 3, w - 1);             // this lane's pixels
    for (int base = 0; base < nops; base += 32) {
      const int i = base + lane;
      bool hit = false;
      if (i < nops) {
        const OpSpan s = op_span(ops + i, y);
        hit = s.row_hit && span_meets(s, wlo, whi);
      }
      const uint32_t hits = __ballot_sync(0xffffffffu, hit);
      if (hits) overlay_apply_hits(ops + base, hits, ov, y, mlo, mhi, row);
    }
  }
}

// Host: apply the draw list in place to frames [n, h, w, 3] (overlay.cu, tile-parallel kernel).
int launch_overlay_tiles(uint8_t* frames, int n, int h, int w, long long row_stride, long long frame_stride,
                         const OverlayDev& ov, cudaStream_t stream);

}  // namespace aiko
