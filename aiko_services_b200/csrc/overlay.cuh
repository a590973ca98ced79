// Device-side replay of ImageOverlay's drawing on one image row held in shared
// memory (reference: R/elements/media/image_io.py:127-155).  Used both by the
// fused resize kernel and by the stand-alone overlay kernel.
#pragma once
#include "common.cuh"

namespace aiko {

struct OverlayDev {
  const aiko_b200_overlay_op* ops;
  const int32_t* op_offsets;
  const uint16_t* patch_ids;
  const uint8_t* luts;
  int patch_w, patch_h;
};

// cv2.rectangle(.., thickness=2) coverage, closed form.
__device__ __forceinline__ bool rect_covers(int x, int y, int xa, int ya, int xb, int yb) {
  if (x < xa - 1 || x > xb + 1 || y < ya - 1 || y > yb + 1) return false;
  if (x >= xa + 2 && x <= xb - 2 && y >= ya + 2 && y <= yb - 2) return false;
  if ((x == xa - 1 || x == xb + 1) && (y == ya - 1 || y == yb + 1)) return false;
  return true;
}

// Applies, in order, every op of frame `frame` that touches image row `y` to
// `row` (w pixels x 3 bytes, shared memory).  Must be called by all threads of
// the CTA with CTA-uniform arguments; `mask` is kThreads/32 words of shared
// scratch.  Ends with the row consistent for all threads (__syncthreads).
__device__ __forceinline__ void overlay_apply_row(const OverlayDev& ov, int frame, int y, int w, uint8_t* row,
                                                  uint32_t* mask) {
  const int ob = ov.op_offsets[frame], oe = ov.op_offsets[frame + 1];
  for (int base = ob; base < oe; base += kThreads) {
    // parallel cull by row range, order preserved through the bit position
    const int i = base + threadIdx.x;
    bool hit = false;
    if (i < oe) {
      const aiko_b200_overlay_op* op = ov.ops + i;
      const int kind = op->kind;
      const int ylo = op->y0 - (kind == AIKO_B200_OP_RECT ? 1 : 0);
      const int yhi = op->y1 + (kind == AIKO_B200_OP_RECT ? 1 : 0);
      hit = (y >= ylo) && (y <= yhi);
    }
    const uint32_t m = __ballot_sync(0xffffffffu, hit);
    if ((threadIdx.x & 31) == 0) mask[threadIdx.x >> 5] = m;
    __syncthreads();
#pragma unroll 1
    for (int wd = 0; wd < kThreads / 32; wd++) {
      uint32_t bits = mask[wd];
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const aiko_b200_overlay_op op = ov.ops[base + wd * 32 + b];
        if (op.kind == AIKO_B200_OP_RECT) {
          const int xs = max(op.x0 - 1, 0), xe = min(op.x1 + 1, w - 1);
          const bool full = (y <= op.y0 + 1) || (y >= op.y1 - 1);  // top / bottom band: whole span
          if (full || op.x1 - op.x0 < 6) {
            for (int x = xs + threadIdx.x; x <= xe; x += kThreads)
              if (rect_covers(x, y, op.x0, op.y0, op.x1, op.y1)) {
                row[3 * x + 0] = op.p0 & 0xff;
                row[3 * x + 1] = (op.p0 >> 8) & 0xff;
                row[3 * x + 2] = (op.p0 >> 16) & 0xff;
              }
          } else if (threadIdx.x < 6) {  // side bands: 3 pixels at each side
            const int x = (threadIdx.x < 3) ? op.x0 - 1 + threadIdx.x : op.x1 - 1 + (threadIdx.x - 3);
            if (x >= 0 && x < w) {
              row[3 * x + 0] = op.p0 & 0xff;
              row[3 * x + 1] = (op.p0 >> 8) & 0xff;
              row[3 * x + 2] = (op.p0 >> 16) & 0xff;
            }
          }
        } else {
          const int px = op.x0 - int(op.p2 & 0xffff), py = op.y0 - int(op.p2 >> 16);
          const int xs = max(op.x0, 0), xe = min(op.x1, w - 1);
          const int prow = (y - py) * ov.patch_w - px;
          const int span = xe - xs + 1;
          for (int t = threadIdx.x; t < span * 3; t += kThreads) {
            const int x = xs + t / 3, c = t - (t / 3) * 3;
            const uint32_t pid = (c == 0) ? (op.p0 & 0xffff) : (c == 1) ? (op.p0 >> 16) : (op.p1 & 0xffff);
            const uint16_t id = ov.patch_ids[(size_t)pid * (ov.patch_w * ov.patch_h) + prow + x];
            if (id) row[3 * x + c] = ov.luts[(size_t)id * 256 + row[3 * x + c]];
          }
        }
        __syncthreads();
      }
    }
    __syncthreads();  // mask is rewritten by the next batch of ops
  }
}

}  // namespace aiko
