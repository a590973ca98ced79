// Workspace layout and plan shared by the two bicubic kernels (resize_bicubic.cu: scalar rows,
// resize_bicubic_mma.cu: tensor-core tiles).  Everything here is a pure function of the four
// sizes, so aiko_b200_resize_bicubic_workspace_bytes / _prepare / _aa_u8 agree without talking.
#pragma once
#include "common.cuh"

namespace aiko {

constexpr int kPrecisionBits = 32 - 8 - 2;   // Pillow's PRECISION_BITS

struct AxisTable {
  int ksize;  // taps per output index (row stride of k)
  int off_bounds;
  int off_k;
};

// Tables of the tensor-core path.  Along x a tile is 16 output pixels and its weights form the A
// operand of mma.m16n8k32 (16 outputs x K source pixels, K = 32 * kx, window start `xa` a multiple
// of 4); along y a group is 8 output rows and its weights form the B operand (K = 32 * ky
// intermediate rows x 8 outputs, window start `ya`).  A 22-bit weight is stored as three byte
// planes (w & 255, (w >> 8) & 255, w >> 16 signed) in the register order of the instruction.
struct MmaPlan {
  int ok;            // sizes are inside what the tensor-core kernel is built for
  int kx, ky;        // K steps of 32 per tile / group
  int ntx, ngy;      // tiles along x, groups along y
  int off_xa;        // [ntx] window start (source pixel) per tile
  int off_ax;        // [ntx][3 planes][kx][32 lanes] uint4
  int off_ya;        // [ngy] window start (intermediate row) per group
  int off_yend;      // [ngy] one past the last intermediate row the group needs
  int off_by;        // [ngy][3 planes][ky][32 lanes] uint2
};

struct BicubicLayout {
  AxisTable x, y;
  MmaPlan mma;
  size_t ints;
};

constexpr int kMmaMaxKx = 4, kMmaMaxKy = 2;

__host__ __device__ inline int bicubic_ksize(int in_size, int out_size) {
  double scale = (double)in_size / out_size;
  if (scale < 1.0) scale = 1.0;
  const double support = 2.0 * scale;
  int c = (int)support;
  if ((double)c < support) c++;
  return c * 2 + 1;
}

// Upper bound of the source span of `outputs` consecutive output indices, plus 3 for rounding the
// window start down to a multiple of 4: (outputs - 1) * scale + 2 * support + 1, rounded up.
inline int bicubic_window_bound(int in_size, int out_size, int outputs) {
  const double scale = (double)in_size / out_size;
  const double support = 2.0 * (scale < 1.0 ? 1.0 : scale);
  return (int)((outputs - 1) * scale + 2.0 * support) + 5;
}

inline BicubicLayout bicubic_layout(int sh, int sw, int dh, int dw, bool with_mma = true) {
  BicubicLayout l;
  l.x.ksize = (bicubic_ksize(sw, dw) + 3) & ~3;  // rows of k padded to 16 bytes
  l.y.ksize = (bicubic_ksize(sh, dh) + 3) & ~3;
  size_t o = 4;   // ints 0..3: header (int 1: error flag raised by the table builders)
  l.x.off_bounds = (int)o; o += 2 * (size_t)dw;
  o = (o + 3) & ~(size_t)3;
  l.x.off_k = (int)o; o += (size_t)dw * l.x.ksize;
  l.y.off_bounds = (int)o; o += 2 * (size_t)dh;
  o = (o + 3) & ~(size_t)3;
  l.y.off_k = (int)o; o += (size_t)dh * l.y.ksize;
  MmaPlan& m = l.mma;
  m.kx = (bicubic_window_bound(sw, dw, 16) + 31) / 32;
  m.ky = (bicubic_window_bound(sh, dh, 8) + 31) / 32;
  m.ntx = (dw + 15) / 16;
  m.ngy = (dh + 7) / 8;
  m.ok = (with_mma && m.kx <= kMmaMaxKx && m.ky <= kMmaMaxKy) ? 1 : 0;
  m.off_xa = m.off_ax = m.off_ya = m.off_yend = m.off_by = 0;
  if (m.ok) {
    o = (o + 3) & ~(size_t)3;
    m.off_xa = (int)o; o += (size_t)m.ntx;
    o = (o + 3) & ~(size_t)3;
    m.off_ax = (int)o; o += (size_t)m.ntx * 3 * m.kx * 128;
    m.off_ya = (int)o; o += (size_t)m.ngy;
    m.off_yend = (int)o; o += (size_t)m.ngy;
    o = (o + 3) & ~(size_t)3;
    m.off_by = (int)o; o += (size_t)m.ngy * 3 * m.ky * 64;
  }
  l.ints = o;
  return l;
}

// The layout a workspace of `bytes` holds: with the tensor-core tables when they fit, else the
// coefficient tables alone (prepare and resize are given the same size, so they agree).
inline BicubicLayout bicubic_layout_for(int sh, int sw, int dh, int dw, size_t bytes) {
  const BicubicLayout full = bicubic_layout(sh, sw, dh, dw, true);
  return full.ints * sizeof(int) <= bytes ? full : bicubic_layout(sh, sw, dh, dw, false);
}

// resize_bicubic_mma.cu
int bicubic_mma_prepare(int* ws, const BicubicLayout& l, int sh, int sw, int dh, int dw, cudaStream_t stream);
// *done = false when the call is outside what the tensor-core kernel handles (the caller then
// runs the scalar kernel); otherwise the launch status.
int bicubic_mma_launch(const uint8_t* src, int n_frames, int sh, int sw, int channels, uint8_t* dst, int dh, int dw,
                       const int* ws, const BicubicLayout& l, cudaStream_t stream, bool* done);

}  // namespace aiko
