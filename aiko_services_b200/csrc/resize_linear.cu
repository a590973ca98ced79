// Batched bilinear resize, bit-exact with cv2.resize(image, (W, H)) [INTER_LINEAR, 8-bit]
// as called by the reference's ImageResize (R/elements/media/image_io.py:216), optionally
// fused with ImageOverlay's drawing (image_io.py:127-155) and an RGB<->BGR swap
// (R/elements/media/video_io.py:299).
//
// Work item = one output row of one frame; each CTA owns a contiguous run of items.
//
// resize3_staged (RGB and gray frames, horizontal scale <= 4): the two source rows an output
// row needs are pulled into shared memory by the TMA engine (1-D cp.async.bulk, mbarrier
// completion) through a ring of stages, so every 32-byte sector of a needed source row crosses
// HBM exactly once and the SM's load/store units only see shared memory.  A thread produces 4
// consecutive output pixels: taps are fetched as aligned words, the horizontal pass is one dp2a
// per channel (two 11-bit weights x two 8-bit taps), the vertical pass one mad.hi per source row.
// Rows that do not start on 16-byte boundaries are copied from the boundary below them and the
// tap offsets are shifted per row (register-resident row path only, as is gray).
//
// resize_rows_generic (RGBA, scale > 4 where most sectors of a source row carry no weight, and
// the shapes the register-resident path does not take): same row structure, taps read straight
// from global memory.
#include <atomic>
#include <math.h>
#include <vector>

#include "overlay.cuh"

namespace aiko {

struct ResizeParams {
  const uint8_t* src;
  uint8_t* dst;
  int n, sh, sw, dh, dw;
  long long src_row_stride, src_frame_stride, dst_row_stride, dst_frame_stride;
  double scale_x, scale_y;
  int swap_rb, has_overlay;
  OverlayDev ov;
  int stages, seg_stride, row_bytes;
  int items, items_per_cta;
  int wait_ns[4];              // producer wait: suspend hint, back-off; consumer wait: suspend hint, back-off
  int fast_rows;               // no draw list in the row loop, whole 4-pixel groups, 4-byte aligned output rows
  int evict_first;             // source rows are read once: keep them from displacing other data in L2
  int keep_output;             // a second pass over the output follows: store it with an L2 evict-last hint
  int unaligned;               // source rows do not start on 16-byte boundaries (register-resident row path only)
  int ch;                      // channels of the staged kernel: 3, or 1 (gray: register-resident row path only)
  int rows_per_slot;           // output rows (pairs of source rows) per ring slot: > 1 for short rows, register-resident path only
};

// cv2's coefficient derivation (float32 phase from a double position, 11-bit weights).
// Every step is an explicitly rounded IEEE operation: no FMA contraction.
__device__ __forceinline__ void linear_coef(int d, double scale, int* s_out, float* f_out) {
  const double pos = __dadd_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), -0.5);
  float f = __double2float_rn(pos);
  const int s = __float2int_rd(f);
  f = __fsub_rn(f, (float)s);
  *s_out = s;
  *f_out = f;
}
__device__ __forceinline__ int coef_q11(float f) { return __float2int_rn(__fmul_rn(f, 2048.f)); }

__device__ __forceinline__ void horizontal_coef(int dx, double scale, int sw, int* s_out, uint32_t* packed) {
  int s;
  float f;
  linear_coef(dx, scale, &s, &f);
  if (s < 0) { s = 0; f = 0.f; }
  if (s >= sw - 1) { s = sw - 1; f = 0.f; }
  const int a0 = coef_q11(__fsub_rn(1.f, f)), a1 = coef_q11(f);
  *s_out = s;
  *packed = (uint32_t)(a0 & 0xffff) | ((uint32_t)a1 << 16);
}

struct RowCoef {
  int y0, y1, b0, b1;
};
__device__ __forceinline__ RowCoef vertical_coef(int dy, double scale, int sh) {
  int sy;
  float fy;
  linear_coef(dy, scale, &sy, &fy);
  RowCoef r;
  r.b0 = coef_q11(__fsub_rn(1.f, fy));
  r.b1 = coef_q11(fy);
  r.y0 = min(max(sy, 0), sh - 1);
  r.y1 = min(max(sy + 1, 0), sh - 1);
  return r;
}

// ((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2 with b pre-shifted by 16.
// The sum never exceeds 1020, so no saturation is needed.
__device__ __forceinline__ uint32_t vblend_hi(uint32_t r0, uint32_t r1, uint32_t b0s, uint32_t b1s) {
  return (__umulhi(b0s, r0 >> 4) + __umulhi(b1s, r1 >> 4) + 2u) >> 2;
}

// ---------------------------------------------------------------------------------------
// Row scheduling shared by the producer warp and the consumer warps of resize3_staged: CTA c
// takes chunks c, c + G, c + 2G, ... of `chunk` consecutive output rows, so rows being read at
// the same time are neighbours in memory and label rows (slow) are spread over many CTAs.
constexpr int kRecHeader = 8;   // ints: n, dy, b0, b1, first op, op count, output row pointer (2)

// Row scheduling of resize3_staged: CTA c takes chunks c, c + G, c + 2G, ... of `chunk`
// consecutive output rows, so rows being read at the same time are neighbours in memory and
// label rows (slow) are spread over many CTAs.
struct RowIter {
  int item, left, n, dy;
  __device__ __forceinline__ void start(int first_item, int items, int chunk, int dh) {
    item = first_item;
    left = min(chunk, items - first_item);
    n = first_item / dh;
    dy = first_item - n * dh;
  }
  __device__ __forceinline__ bool valid(int items) const { return item < items; }
  __device__ __forceinline__ void next(int items, int chunk, int stride_items, int dh) {
    if (--left > 0) {
      ++item;
      if (++dy == dh) { dy = 0; ++n; }
    } else {
      item += stride_items - chunk + 1;
      if (item < items) start(item, items, chunk, dh);
    }
  }
};

// One RGB output pixel from the 2x2 taps around it.  lo/hi hold the 6 source bytes of a row
// (t0c0 t0c1 t0c2 t1c0 | t1c1 t1c2 . .), coef = a0 | a1 << 16.
__device__ __forceinline__ void pixel_general(uint32_t ulo, uint32_t uhi, uint32_t vlo, uint32_t vhi, uint32_t coef,
                                              uint32_t b0s, uint32_t b1s, uint32_t* c0, uint32_t* c1, uint32_t* c2) {
  const uint32_t up01 = __byte_perm(ulo, uhi, 0x4130), up2 = __byte_perm(ulo, uhi, 0x0052);
  const uint32_t vp01 = __byte_perm(vlo, vhi, 0x4130), vp2 = __byte_perm(vlo, vhi, 0x0052);
  *c0 = vblend_hi(__dp2a_lo(coef, up01, 0u), __dp2a_lo(coef, vp01, 0u), b0s, b1s);
  *c1 = vblend_hi(__dp2a_hi(coef, up01, 0u), __dp2a_hi(coef, vp01, 0u), b0s, b1s);
  *c2 = vblend_hi(__dp2a_lo(coef, up2, 0u), __dp2a_lo(coef, vp2, 0u), b0s, b1s);
}
// Same arithmetic when every horizontal weight pair is (2048, 0) -- an odd integer scale such
// as 1920 -> 640: R = 2048 * tap, so (b * (R >> 4)) >> 16 == (b * tap) >> 9, computed as
// umulhi(b << 15, tap << 8) (b <= 2048 keeps b << 15 inside 32 bits).
__device__ __forceinline__ uint32_t vblend_decim(uint32_t t0, uint32_t t1, uint32_t b0d, uint32_t b1d) {
  return (__umulhi(b0d, t0) + __umulhi(b1d, t1) + 2u) >> 2;
}

// Four consecutive output pixels (12 bytes, 3 words) from the two staged source rows.
// DECIM: horizontal pass degenerates to picking one source pixel (all weights 2048, 0).
template <bool DECIM>
__device__ __forceinline__ void group_pixels(const uint8_t* r0, const uint8_t* r1, const int (&wo)[4], const int (&sh)[4],
                                             const int (&wo1)[4], const int (&sh1)[4], const uint32_t (&cs)[4],
                                             uint32_t b0s, uint32_t b1s, uint32_t sel_rgb, uint32_t (&o)[3]) {
  uint32_t px[4];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint32_t* w0 = reinterpret_cast<const uint32_t*>(r0 + wo[k]);
    const uint32_t* w1 = reinterpret_cast<const uint32_t*>(r1 + wo1[k]);
    uint32_t c0, c1, c2;
    if (DECIM) {
      const uint32_t u = __funnelshift_r(w0[0], w0[1], sh[k]), v = __funnelshift_r(w1[0], w1[1], sh1[k]);
      c0 = vblend_decim(__byte_perm(u, 0, 0x4404), __byte_perm(v, 0, 0x4404), b0s, b1s);   // tap << 8
      c1 = vblend_decim(__byte_perm(u, 0, 0x4414), __byte_perm(v, 0, 0x4414), b0s, b1s);
      c2 = vblend_decim(__byte_perm(u, 0, 0x4424), __byte_perm(v, 0, 0x4424), b0s, b1s);
    } else {
      const uint32_t u0 = w0[0], u1 = w0[1], u2 = w0[2];
      const uint32_t v0 = w1[0], v1 = w1[1], v2 = w1[2];
      pixel_general(__funnelshift_r(u0, u1, sh[k]), __funnelshift_r(u1, u2, sh[k]), __funnelshift_r(v0, v1, sh1[k]),
                    __funnelshift_r(v1, v2, sh1[k]), cs[k], b0s, b1s, &c0, &c1, &c2);
    }
    // c0 c1 c2 -> 24-bit pixel in stored channel order (sel_rgb swaps R and B when asked)
    px[k] = __byte_perm(__byte_perm(c0, c1, 0x0040), c2, sel_rgb);
  }
  o[0] = __byte_perm(px[0], px[1], 0x4210);
  o[1] = __byte_perm(px[1], px[2], 0x5421);
  o[2] = __byte_perm(px[2], px[3], 0x6542);
}

template <bool DECIM>
__device__ __forceinline__ void consume_row(const uint8_t* r0, const uint8_t* r1, const int* xw, const int* xs,
                                            const uint32_t* xcoef, uint8_t* orow, int g_first, int g_step, int groups,
                                            uint32_t b0, uint32_t b1, uint32_t sel_rgb) {
  const uint32_t b0s = DECIM ? (b0 << 15) : (b0 << 16), b1s = DECIM ? (b1 << 15) : (b1 << 16);
  for (int g = g_first; g < groups; g += g_step) {
    const int4 w4 = reinterpret_cast<const int4*>(xw)[g];
    const int4 s4 = reinterpret_cast<const int4*>(xs)[g];
    const int wo[4] = {w4.x, w4.y, w4.z, w4.w};
    const int sh[4] = {s4.x, s4.y, s4.z, s4.w};
    uint32_t cs[4] = {0, 0, 0, 0};
    if (!DECIM) {
      const uint4 c4 = reinterpret_cast<const uint4*>(xcoef)[g];
      cs[0] = c4.x; cs[1] = c4.y; cs[2] = c4.z; cs[3] = c4.w;
    }
    uint32_t o[3];
    group_pixels<DECIM>(r0, r1, wo, sh, wo, sh, cs, b0s, b1s, sel_rgb, o);
    uint32_t* d = reinterpret_cast<uint32_t*>(orow) + 3 * g;
    d[0] = o[0]; d[1] = o[1]; d[2] = o[2];
  }
}

// The common case without a draw list in the row loop: every thread owns one group of 4 output
// pixels for the whole kernel, so its tap offsets and weights stay in registers and its 12 result
// bytes go from registers straight to global memory (the destination pointer of the row comes
// from the producer's record).  No per-row table loads, no output row in shared memory.
// UNALIGNED: source rows that do not start on a 16-byte boundary were copied from the boundary
// below, so each staged row starts `delta` (0..15, from the record) bytes into its buffer and the
// word / shift split of a tap offset is redone per row.
template <bool DECIM, bool UNALIGNED, int G>
__device__ __forceinline__ void consume_rows_fast(const ResizeParams& p, uint64_t* full, uint64_t* empty, const int* recs,
                                                  const uint8_t* stage, int my_items, const int* xw, const int* xs,
                                                  const uint32_t* xcoef, int g0, int gstep, int groups, uint32_t sel_rgb,
                                                  int lane) {
  // G groups per thread (rows wider than 1024 output pixels): group g0 + i * gstep, so the lanes of a
  // warp still write consecutive 12-byte pieces
  int wo[G][4], sh[G][4];
  uint32_t cs[G][4];
  bool on[G];
#pragma unroll
  for (int i = 0; i < G; i++) {
    const int g = g0 + i * gstep;
    on[i] = g < groups;
    const int4 w4 = on[i] ? reinterpret_cast<const int4*>(xw)[g] : make_int4(0, 0, 0, 0);
    const int4 s4 = on[i] ? reinterpret_cast<const int4*>(xs)[g] : make_int4(0, 0, 0, 0);
    wo[i][0] = w4.x; wo[i][1] = w4.y; wo[i][2] = w4.z; wo[i][3] = w4.w;
    sh[i][0] = s4.x; sh[i][1] = s4.y; sh[i][2] = s4.z; sh[i][3] = s4.w;
    const uint4 c4 = (!DECIM && on[i]) ? reinterpret_cast<const uint4*>(xcoef)[g] : make_uint4(0, 0, 0, 0);
    cs[i][0] = c4.x; cs[i][1] = c4.y; cs[i][2] = c4.z; cs[i][3] = c4.w;
  }
  int slot = 0, round = 0;
  uint64_t keep;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
  const int R = p.rows_per_slot;
  int sub = 0;                                     // row inside the slot
  for (int it = 0; it < my_items; ++it) {
    if (sub == 0) mbar_wait(&full[slot], (uint32_t)(round & 1), (uint32_t)p.wait_ns[2], (uint32_t)p.wait_ns[3]);
    const int* rec = recs + (slot * R + sub) * kRecHeader;
    const int2 bb = *reinterpret_cast<const int2*>(rec + 2);                       // b0, b1
    uint32_t* drow = *reinterpret_cast<uint32_t* const*>(rec + 6);                 // first byte of the output row
    const uint8_t* r0 = stage + (size_t)(slot * R + sub) * 2 * p.seg_stride;
    const uint32_t b0s = DECIM ? ((uint32_t)bb.x << 15) : ((uint32_t)bb.x << 16);
    const uint32_t b1s = DECIM ? ((uint32_t)bb.y << 15) : ((uint32_t)bb.y << 16);
    const int2 delta = UNALIGNED ? *reinterpret_cast<const int2*>(rec + 4) : make_int2(0, 0);
#pragma unroll
    for (int i = 0; i < G; i++) {
      if (!on[i]) continue;
      uint32_t o[3];
      if (UNALIGNED) {
        int wa[4], sa[4], wb[4], sb[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
          const int exact = wo[i][k] + (sh[i][k] >> 3);      // byte offset of the tap in the source row
          wa[k] = (exact + delta.x) & ~3; sa[k] = ((exact + delta.x) & 3) * 8;
          wb[k] = (exact + delta.y) & ~3; sb[k] = ((exact + delta.y) & 3) * 8;
        }
        group_pixels<DECIM>(r0, r0 + p.seg_stride, wa, sa, wb, sb, cs[i], b0s, b1s, sel_rgb, o);
      } else {
        group_pixels<DECIM>(r0, r0 + p.seg_stride, wo[i], sh[i], wo[i], sh[i], cs[i], b0s, b1s, sel_rgb, o);
      }
      uint32_t* d = drow + 3 * (g0 + i * gstep);
      if (p.keep_output) {      // the overlay pass reads these pixels next: ask L2 to hold on to them
        asm volatile("st.global.L1::no_allocate.L2::cache_hint.b32 [%0], %1, %2;" ::"l"(d), "r"(o[0]), "l"(keep) : "memory");
        asm volatile("st.global.L1::no_allocate.L2::cache_hint.b32 [%0], %1, %2;" ::"l"(d + 1), "r"(o[1]), "l"(keep) : "memory");
        asm volatile("st.global.L1::no_allocate.L2::cache_hint.b32 [%0], %1, %2;" ::"l"(d + 2), "r"(o[2]), "l"(keep) : "memory");
      } else {
        asm volatile("st.global.L1::no_allocate.b32 [%0], %1;" ::"l"(d), "r"(o[0]) : "memory");
        asm volatile("st.global.L1::no_allocate.b32 [%0], %1;" ::"l"(d + 1), "r"(o[1]) : "memory");
        asm volatile("st.global.L1::no_allocate.b32 [%0], %1;" ::"l"(d + 2), "r"(o[2]) : "memory");
      }
    }
    if (++sub == R || it + 1 == my_items) {
      sub = 0;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);    // the staged rows of this slot are free again
      if (++slot == p.stages) { slot = 0; ++round; }
    }
  }
}

template <int G>
__device__ __forceinline__ void consume_rows_fast_g(const ResizeParams& p, bool decim, uint64_t* full, uint64_t* empty,
                                                    const int* recs, const uint8_t* stage, int my_items, const int* xw,
                                                    const int* xs, const uint32_t* xcoef, int g0, int gstep, int groups,
                                                    uint32_t sel_rgb, int lane) {
  if (p.unaligned) {
    if (decim) consume_rows_fast<true, true, G>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g0, gstep, groups, sel_rgb, lane);
    else consume_rows_fast<false, true, G>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g0, gstep, groups, sel_rgb, lane);
  } else {
    if (decim) consume_rows_fast<true, false, G>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g0, gstep, groups, sel_rgb, lane);
    else consume_rows_fast<false, false, G>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g0, gstep, groups, sel_rgb, lane);
  }
}

// Gray frames on the same ring: a thread owns 4 consecutive output pixels = one output word; a tap
// pair is two adjacent bytes (one dp2a), taken from two aligned words with a funnel shift.
// DECIM (every horizontal weight pair is (2048, 0), e.g. 1920 -> 640): the tap is ONE byte per row, read as a
// byte, and the vertical blend is (b * tap) >> 9 as in vblend_decim.
template <bool DECIM>
__device__ __forceinline__ void consume_rows_gray(const ResizeParams& p, uint64_t* full, uint64_t* empty, const int* recs,
                                                  const uint8_t* stage, int my_items, const int* xw, const int* xs,
                                                  const uint32_t* xcoef, int g, bool on, int lane) {
  int exact[4] = {0, 0, 0, 0};
  uint32_t cs[4] = {0, 0, 0, 0};
  if (on) {
    const int4 w4 = reinterpret_cast<const int4*>(xw)[g];
    const int4 s4 = reinterpret_cast<const int4*>(xs)[g];
    const uint4 c4 = reinterpret_cast<const uint4*>(xcoef)[g];
    exact[0] = w4.x + (s4.x >> 3); exact[1] = w4.y + (s4.y >> 3); exact[2] = w4.z + (s4.z >> 3); exact[3] = w4.w + (s4.w >> 3);
    cs[0] = c4.x; cs[1] = c4.y; cs[2] = c4.z; cs[3] = c4.w;
  }
  int slot = 0, round = 0;
  const int R = p.rows_per_slot;
  int sub = 0;                                     // row inside the slot
  for (int it = 0; it < my_items; ++it) {
    if (sub == 0) mbar_wait(&full[slot], (uint32_t)(round & 1), (uint32_t)p.wait_ns[2], (uint32_t)p.wait_ns[3]);
    const int* rec = recs + (slot * R + sub) * kRecHeader;
    const int2 bb = *reinterpret_cast<const int2*>(rec + 2);                       // b0, b1
    const int2 delta = p.unaligned ? *reinterpret_cast<const int2*>(rec + 4) : make_int2(0, 0);
    uint32_t* drow = *reinterpret_cast<uint32_t* const*>(rec + 6);                 // first byte of the output row
    const uint8_t* r0 = stage + (size_t)(slot * R + sub) * 2 * p.seg_stride;
    const uint8_t* r1 = r0 + p.seg_stride;
    if (on) {
      const uint32_t b0s = (uint32_t)bb.x << (DECIM ? 15 : 16), b1s = (uint32_t)bb.y << (DECIM ? 15 : 16);
      uint32_t out = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int e0 = exact[k] + delta.x, e1 = exact[k] + delta.y;
        if (DECIM) {
          out |= vblend_decim((uint32_t)r0[e0] << 8, (uint32_t)r1[e1] << 8, b0s, b1s) << (8 * k);
          continue;
        }
        const uint32_t* w0 = reinterpret_cast<const uint32_t*>(r0 + (e0 & ~3));
        const uint32_t* w1 = reinterpret_cast<const uint32_t*>(r1 + (e1 & ~3));
        const uint32_t u = __funnelshift_r(w0[0], w0[1], (e0 & 3) * 8), v = __funnelshift_r(w1[0], w1[1], (e1 & 3) * 8);
        out |= vblend_hi(__dp2a_lo(cs[k], u, 0u), __dp2a_lo(cs[k], v, 0u), b0s, b1s) << (8 * k);
      }
      asm volatile("st.global.L1::no_allocate.b32 [%0], %1;" ::"l"(drow + g), "r"(out) : "memory");
    }
    if (++sub == R || it + 1 == my_items) {
      sub = 0;
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[slot]);
      if (++slot == p.stages) { slot = 0; ++round; }
    }
  }
}

// Warp roles: warps [0, ncons) consume rows; warp ncons is the TMA producer.  Per ring slot the
// producer brings in the two source rows; `full` completes when the bytes have landed, `empty`
// when every consumer warp is done with the slot.
__global__ void __launch_bounds__(352) resize3_staged(const ResizeParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int groups = (p.dw + 3) >> 2;  // 4 output pixels per group
  const int dwp = groups * 4;
  int* xw = reinterpret_cast<int*>(smem);                               // word-aligned byte offset of tap 0
  int* xs = reinterpret_cast<int*>(smem + (size_t)dwp * 4);             // 8 * (offset & 3): funnel shift
  uint32_t* xcoef = reinterpret_cast<uint32_t*>(smem + (size_t)dwp * 8);  // a0 | a1 << 16
  size_t off = (size_t)dwp * 12;
  uint8_t* orow = smem + off;                      // one output row; each warp owns 384-byte segments
  off += ((size_t)groups * 12 + 15) & ~(size_t)15;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + off);
  uint64_t* empty = full + 8;
  off += 128;
  int* recs = reinterpret_cast<int*>(smem + off);  // per slot: n, dy, b0, b1, first op, op count, -, -
  off += (size_t)p.stages * p.rows_per_slot * kRecHeader * 4;
  uint8_t* stage = smem + off;

  const int ncons = (int)blockDim.x / 32 - 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  bool not_decim = false;
  for (int dx = threadIdx.x; dx < dwp; dx += blockDim.x) {
    int s;
    uint32_t c;
    horizontal_coef(min(dx, p.dw - 1), p.scale_x, p.sw, &s, &c);
    xw[dx] = (s * p.ch) & ~3;
    xs[dx] = ((s * p.ch) & 3) * 8;
    xcoef[dx] = c;
    not_decim |= (c != 2048u);
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], ncons);
    }
    fence_mbar_init();
  }
  const bool decim = !__syncthreads_or(not_decim);

  const int chunk = p.items_per_cta;               // rows per chunk
  const int stride_items = (int)gridDim.x * chunk;
  // rows this CTA handles (chunks blockIdx, blockIdx + G, ...; the globally last chunk may be short)
  const int total_chunks = (p.items + chunk - 1) / chunk;
  const int my_chunks = ((int)blockIdx.x < total_chunks) ? (total_chunks - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  int my_items = my_chunks * chunk;
  if (my_chunks > 0 && ((int)blockIdx.x + (my_chunks - 1) * (int)gridDim.x) == total_chunks - 1)
    my_items -= total_chunks * chunk - p.items;

  if (warp == ncons) {
    // ---- producer: one lane feeds the TMA ring ---------------------------------------------
    if (lane != 0 || my_items == 0) return;
    RowIter ri, rn;                                // current row, next row (to prefetch its list offsets)
    ri.start(blockIdx.x * chunk, p.items, chunk, p.dh);
    rn = ri;
    int ob = 0, oe = 0, ob_next = 0, oe_next = 0;
    if (p.has_overlay) { const int li = overlay_list(p.ov, ri.n); ob = __ldg(p.ov.op_offsets + li); oe = __ldg(p.ov.op_offsets + li + 1); }
    int slot = 0, round = 0, sub = 0;               // a slot takes rows_per_slot consecutive rows (1 unless rows are short)
    uint32_t slot_bytes = 0;
    const int R = p.rows_per_slot;
    const uint64_t policy = l2_policy_evict_first();
    for (int it = 0; it < my_items; ++it) {
      if (sub == 0 && round > 0)
        mbar_wait(&empty[slot], (uint32_t)((round - 1) & 1), (uint32_t)p.wait_ns[0], (uint32_t)p.wait_ns[1]);
      const RowCoef rc = vertical_coef(ri.dy, p.scale_y, p.sh);
      uint8_t* s0 = stage + (size_t)(slot * R + sub) * 2 * p.seg_stride;
      const uint8_t* f = p.src + (long long)ri.n * p.src_frame_stride;
      // a row is copied from the 16-byte boundary below its first byte to the one above its last
      const uint8_t* a0 = f + (long long)rc.y0 * p.src_row_stride;
      const uint8_t* a1 = f + (long long)rc.y1 * p.src_row_stride;
      const uint32_t d0 = (uint32_t)(reinterpret_cast<uintptr_t>(a0) & 15), d1 = (uint32_t)(reinterpret_cast<uintptr_t>(a1) & 15);
      const uint32_t n0 = (d0 + (uint32_t)p.row_bytes + 15u) & ~15u, n1 = (d1 + (uint32_t)p.row_bytes + 15u) & ~15u;
      if (p.evict_first) {      // each source row is read once (vertical scale >= 2)
        bulk_g2s_hint(s0, a0 - d0, n0, &full[slot], policy);
        bulk_g2s_hint(s0 + p.seg_stride, a1 - d1, n1, &full[slot], policy);
      } else {
        bulk_g2s(s0, a0 - d0, n0, &full[slot]);
        bulk_g2s(s0 + p.seg_stride, a1 - d1, n1, &full[slot]);
      }
      const int nops = oe - ob;
      if (p.has_overlay && it + 1 < my_items) {    // next row's offsets: in flight while this row is consumed
        rn.next(p.items, chunk, stride_items, p.dh);
        const int li = overlay_list(p.ov, rn.n);
        ob_next = __ldg(p.ov.op_offsets + li);
        oe_next = __ldg(p.ov.op_offsets + li + 1);
      }
      int* rec = recs + (slot * R + sub) * kRecHeader;
      *reinterpret_cast<int4*>(rec) = make_int4(ri.n, ri.dy, rc.b0, rc.b1);
      rec[4] = p.unaligned ? (int)d0 : ob;          // (unaligned rows never carry a draw list in the row loop)
      rec[5] = p.unaligned ? (int)d1 : nops;
      *reinterpret_cast<uint8_t**>(rec + 6) = p.dst + (long long)ri.n * p.dst_frame_stride + (long long)ri.dy * p.dst_row_stride;
      slot_bytes += n0 + n1;
      ri.next(p.items, chunk, stride_items, p.dh);
      ob = ob_next; oe = oe_next;
      if (++sub == R || it + 1 == my_items) {       // the slot is complete: one arrival for all its bytes
        mbar_expect_tx(&full[slot], slot_bytes);    // release: records visible
        slot_bytes = 0;
        sub = 0;
        if (++slot == p.stages) { slot = 0; ++round; }
      }
    }
    return;
  }

  // ---- consumers: each warp produces, overlays and stores its own 128-pixel segments --------
  const int row_bytes_out = p.dw * 3;
  const uint32_t sel_rgb = p.swap_rb ? 0x7014u : 0x7410u;   // (c0 c1 . .),(c2) -> c0 c1 c2 0 | c2 c1 c0 0
  if (p.fast_rows) {
    const int g = warp * 32 + lane;
    const bool on = g < groups;
    if (p.ch == 1) {
      if (decim) consume_rows_gray<true>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g, on, lane);
      else consume_rows_gray<false>(p, full, empty, recs, stage, my_items, xw, xs, xcoef, g, on, lane);
    } else {
      const int gstep = ncons * 32, per_thread = (groups + gstep - 1) / gstep;
      if (per_thread <= 1) consume_rows_fast_g<1>(p, decim, full, empty, recs, stage, my_items, xw, xs, xcoef, g, gstep, groups, sel_rgb, lane);
      else if (per_thread == 2) consume_rows_fast_g<2>(p, decim, full, empty, recs, stage, my_items, xw, xs, xcoef, g, gstep, groups, sel_rgb, lane);
      else consume_rows_fast_g<4>(p, decim, full, empty, recs, stage, my_items, xw, xs, xcoef, g, gstep, groups, sel_rgb, lane);
    }
    return;
  }
  int slot = 0, round = 0;
  for (int it = 0; it < my_items; ++it) {
    mbar_wait(&full[slot], (uint32_t)(round & 1), (uint32_t)p.wait_ns[2], (uint32_t)p.wait_ns[3]);
    const int* rec = recs + slot * kRecHeader;
    const int4 hdr = *reinterpret_cast<const int4*>(rec);   // n, dy, b0, b1
    const uint8_t* r0 = stage + (size_t)slot * 2 * p.seg_stride;
    const uint8_t* r1 = r0 + p.seg_stride;
    if (decim)
      consume_row<true>(r0, r1, xw, xs, xcoef, orow, warp * 32 + lane, ncons * 32, groups, (uint32_t)hdr.z,
                        (uint32_t)hdr.w, sel_rgb);
    else
      consume_row<false>(r0, r1, xw, xs, xcoef, orow, warp * 32 + lane, ncons * 32, groups, (uint32_t)hdr.z,
                         (uint32_t)hdr.w, sel_rgb);
    __syncwarp();
    if (p.has_overlay) {
      const int ob = rec[4], nops = rec[5];
      if (nops > 0) overlay_apply_row(p.ov.ops + ob, nops, p.ov, hdr.y, p.dw, orow, ncons * 32);
      __syncwarp();
    }
    uint8_t* drow = p.dst + (long long)hdr.x * p.dst_frame_stride + (long long)hdr.y * p.dst_row_stride;
    for (int gbase = warp * 32; gbase < groups; gbase += ncons * 32) {
      const int b0 = gbase * 12, b1 = min(b0 + 384, row_bytes_out);   // this warp's bytes of the row
      if ((reinterpret_cast<uintptr_t>(drow) & 15) == 0) {
        const int nvec = (b1 - b0) >> 4;
        if (lane < nvec)
          stg_stream(reinterpret_cast<uint4*>(drow + b0) + lane, reinterpret_cast<const uint4*>(orow + b0)[lane]);
        for (int i = b0 + (nvec << 4) + lane; i < b1; i += 32) drow[i] = orow[i];
      } else {
        for (int i = b0 + lane; i < b1; i += 32) drow[i] = orow[i];
      }
    }
    __syncwarp();                                  // the segment is rewritten by the next row
    if (lane == 0) mbar_arrive(&empty[slot]);      // rows and records of this slot are free again
    if (++slot == p.stages) { slot = 0; ++round; }
  }
}

// ---------------------------------------------------------------------------------------
// kGenericRows output rows per round: the byte gathers of all of them are in flight before the
// first result is needed, so a round pays one memory latency instead of one per row.
template <int C>
struct GenericRows {
  // gray: 4 rows (16 byte loads in flight per thread).  Colour: 1 -- with more, the source rows the
  // resident CTAs of an SM gather from no longer fit its L1 (measured 81 / 90 / 85 us for 1 / 2 / 4 rows)
  static constexpr int value = C == 1 ? 4 : 1;
};

template <int C>
__global__ void __launch_bounds__(kThreads) resize_rows_generic(const ResizeParams p) {
  constexpr int kGenericRows = GenericRows<C>::value;
  extern __shared__ __align__(128) uint8_t smem[];
  int2* xtab = reinterpret_cast<int2*>(smem);
  const size_t off = ((size_t)p.dw * 8 + 15) & ~(size_t)15;
  const int row_pitch = (p.dw * C + 15) & ~15;
  uint8_t* orows = smem + off;
  __shared__ int4 s_coef[kGenericRows];   // y0, y1, b0 << 16, b1 << 16
  __shared__ int2 s_where[kGenericRows];  // frame, output row

  for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
    int s;
    uint32_t c;
    horizontal_coef(dx, p.scale_x, p.sw, &s, &c);
    xtab[dx] = make_int2(s * C, (int)c);
  }

  const int last_off = (p.sw - 1) * C;
  const int rounds = (p.items + kGenericRows - 1) / kGenericRows;
  for (int round = blockIdx.x; round < rounds; round += gridDim.x) {
    const int i0 = round * kGenericRows;
    const int nrows = min(kGenericRows, p.items - i0);
    if (threadIdx.x < nrows) {
      const int item = i0 + threadIdx.x;
      const int n = item / p.dh, dy = item - n * p.dh;
      const RowCoef rc = vertical_coef(dy, p.scale_y, p.sh);
      s_coef[threadIdx.x] = make_int4(rc.y0, rc.y1, rc.b0 << 16, rc.b1 << 16);
      s_where[threadIdx.x] = make_int2(n, dy);
    }
    __syncthreads();   // also orders xtab (first round) and the previous round's stores
    for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
      const int2 e = xtab[dx];
      const uint32_t a0 = (uint32_t)e.y & 0xffff, a1 = (uint32_t)e.y >> 16;
      const int o0 = e.x, o1 = min(e.x + C, last_off);
      uint32_t t[kGenericRows][C][4];
#pragma unroll
      for (int j = 0; j < kGenericRows; j++) {
        if (j < nrows) {
          const int4 rc = s_coef[j];
          const uint8_t* f = p.src + (long long)s_where[j].x * p.src_frame_stride;
          const uint8_t* r0 = f + (long long)rc.x * p.src_row_stride;
          const uint8_t* r1 = f + (long long)rc.y * p.src_row_stride;
#pragma unroll
          for (int c = 0; c < C; c++) {
            t[j][c][0] = __ldg(r0 + o0 + c);
            t[j][c][1] = __ldg(r0 + o1 + c);
            t[j][c][2] = __ldg(r1 + o0 + c);
            t[j][c][3] = __ldg(r1 + o1 + c);
          }
        }
      }
#pragma unroll
      for (int j = 0; j < kGenericRows; j++) {
        if (j < nrows) {
          const int4 rc = s_coef[j];
          uint8_t* orow = orows + j * row_pitch;
#pragma unroll
          for (int c = 0; c < C; c++) {
            const uint32_t v = vblend_hi(t[j][c][0] * a0 + t[j][c][1] * a1, t[j][c][2] * a0 + t[j][c][3] * a1,
                                         (uint32_t)rc.z, (uint32_t)rc.w);
            const int cc = (p.swap_rb && C >= 3 && c < 3) ? 2 - c : c;
            orow[dx * C + cc] = (uint8_t)v;
          }
        }
      }
    }
    __syncthreads();
    if (C == 3 && p.has_overlay) {
      for (int j = 0; j < nrows; j++) {
        const int n = s_where[j].x;
        const int li = overlay_list(p.ov, n);
        const int ob = p.ov.op_offsets[li];
        overlay_apply_row(p.ov.ops + ob, p.ov.op_offsets[li + 1] - ob, p.ov, s_where[j].y, p.dw, orows + j * row_pitch,
                          kThreads);
      }
      __syncthreads();
    }
    for (int j = 0; j < nrows; j++)
      store_row(p.dst + (long long)s_where[j].x * p.dst_frame_stride + (long long)s_where[j].y * p.dst_row_stride,
                orows + j * row_pitch, p.dw * C);
    __syncthreads();  // rows and their records are rewritten by the next round
  }
}

static int env_int(const char* name, int fallback) {
  const char* v = getenv(name);
  return v ? atoi(v) : fallback;
}

static int launch_staged3(ResizeParams p, cudaStream_t stream, bool* done) {
  *done = false;
  const int groups = (p.dw + 3) / 4;
  // consumer warps: one 4-pixel group per thread up to 10 warps (1280 output pixels), beyond that the
  // fewest groups per thread, spread evenly (1920 pixels: 2 groups x 8 warps)
  const int per_thread = (groups + 32 * 10 - 1) / (32 * 10);
  int ncons = (groups + 32 * per_thread - 1) / (32 * per_thread);
  ncons = env_int("AIKO_B200_RESIZE_WARPS", ncons);
  if (ncons < 1) ncons = 1;
  if (ncons > 10) ncons = 10;
  const int threads = (ncons + 1) * 32;            // + the TMA producer warp
  p.row_bytes = p.sw * p.ch;
  p.fast_rows = (!p.has_overlay && (p.dw & 3) == 0 && groups <= ncons * 32 * (p.ch == 1 ? 1 : 4) && !env_int("AIKO_B200_RESIZE_SLOW_ROWS", 0) &&
                 ((reinterpret_cast<uintptr_t>(p.dst) | (uintptr_t)p.dst_row_stride | (uintptr_t)p.dst_frame_stride) & 3) == 0)
                    ? 1 : 0;
  if ((p.unaligned || p.ch == 1) && !p.fast_rows) return AIKO_B200_OK;   // not staged: the generic kernel takes it
  p.seg_stride = ((p.row_bytes + 15) & ~15) + 32;  // room for the 16-byte rounding at both ends; tap reads may run a few bytes past the row end
  const size_t fixed = (size_t)groups * 48 + (((size_t)groups * 12 + 15) & ~(size_t)15) + 128;
  // Short rows (gray, small frames): several output rows share a ring slot, so the hand-off between
  // producer and consumers (one barrier round trip, ~1 us) is paid once per ~12 KB and not per row.
  p.rows_per_slot = 1;
  if (p.fast_rows) {
    int r = 12288 / (2 * p.seg_stride);
    r = env_int("AIKO_B200_RESIZE_ROWS_PER_SLOT", r);
    p.rows_per_slot = r < 1 ? 1 : (r > 8 ? 8 : r);
  }
  auto need = [&](int st) {
    return fixed + (size_t)st * ((size_t)p.rows_per_slot * kRecHeader * 4 +
                                 (size_t)p.rows_per_slot * 2 * (size_t)p.seg_stride);
  };
  // Measured on 32 x 1080p (us per launch; CTAs per SM x ring stages): 6x2 37.8, 5x2 35.5, 4x3 37.5, 5x3 39.5,
  // 4x4 39.3, 3x2 38.6, 3x3 35.1, 2x4 35.1, 2x6 35.1, 2x8 35.9, 1x8 55.4.  The consumers wait for the ring
  // half of the time, so few CTAs with deeper rings lose nothing, and large shared-memory footprints per SM
  // (> ~200 KB) were consistently slower: three stages, at most three CTAs per SM.
  // (The single-kernel overlay variant keeps the settings it was tuned with: every occupancy slot, two stages.)
  int stages = env_int("AIKO_B200_RESIZE_STAGES", p.has_overlay ? 2 : 3);
  if (stages > 8) stages = 8;
  const size_t max_smem = (size_t)device_max_smem_optin();
  while (stages > 2 && need(stages) > max_smem) stages--;
  if (stages < 2 || need(stages) > max_smem) return AIKO_B200_OK;  // not staged
  p.stages = stages;
  const size_t smem = need(stages);
  // (attribute + occupancy of the last launch shape are remembered: a stream repeats one shape)
  static thread_local struct { int dev = -1, threads = 0, per_sm = 0; size_t smem = 0; } seen;
  int device = 0, per_sm = 0;
  AIKO_CUDA(cudaGetDevice(&device));
  static std::atomic<int> smem_limit[64];          // per device: the limit only ever grows (it is process-wide)
  if (device >= 0 && device < 64 && smem_limit[device].load() < (int)smem) {
    AIKO_CUDA(cudaFuncSetAttribute(resize3_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int cur = smem_limit[device].load();
    while (cur < (int)smem && !smem_limit[device].compare_exchange_weak(cur, (int)smem)) {}
  }
  if (seen.dev == device && seen.threads == threads && seen.smem == smem) {
    per_sm = seen.per_sm;
  } else {
    AIKO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resize3_staged, threads, smem));
    seen.dev = device; seen.threads = threads; seen.smem = smem; seen.per_sm = per_sm;
  }
  if (per_sm < 1) per_sm = 1;
  // Source rows in page-locked host memory (the TMA engine reads them over PCIe): the link, not the
  // SMs, sets the pace, and one CTA per SM keeps far more bytes in flight than it needs.  Leaving
  // the rest of each SM free lets the kernels of other streams (the previous batch's overlay) run
  // beside this one instead of queueing behind its persistent CTAs for milliseconds.
  cudaPointerAttributes attr{};
  if (cudaPointerGetAttributes(&attr, p.src) == cudaSuccess && attr.type == cudaMemoryTypeHost) per_sm = 1;
  else (void)cudaGetLastError();
  // (short rows, < 4 KB: measured better with up to 6 CTAs -- 640x480 RGB 92 -> 82 us, gray 1080p 28.9 -> 26.9 us)
  { const int cap = env_int("AIKO_B200_RESIZE_CTAS_PER_SM", p.has_overlay ? 0 : (p.row_bytes < 4096 ? 6 : 3));
    if (cap > 0 && cap < per_sm) per_sm = cap; }
  int chunk = env_int("AIKO_B200_RESIZE_CHUNK", 1);
  if (chunk < 1) chunk = 1;
  if (p.rows_per_slot > 1) chunk = p.rows_per_slot;   // a slot is one chunk of consecutive rows
  long long grid = (long long)sm_count() * per_sm;
  const long long chunks = ((long long)p.items + chunk - 1) / chunk;
  if (grid > chunks) grid = chunks;
  p.items_per_cta = chunk;
  p.evict_first = (p.sh >= 2 * p.dh && env_int("AIKO_B200_RESIZE_EVICT_FIRST", 1)) ? 1 : 0;
  p.keep_output = env_int("AIKO_B200_RESIZE_KEEP_OUTPUT", p.keep_output);
  p.wait_ns[0] = env_int("AIKO_B200_RESIZE_PW", 1000); p.wait_ns[1] = env_int("AIKO_B200_RESIZE_PB", 100);
  p.wait_ns[2] = env_int("AIKO_B200_RESIZE_CW", 500); p.wait_ns[3] = env_int("AIKO_B200_RESIZE_CB", 0);
  resize3_staged<<<(unsigned)grid, threads, smem, stream>>>(p);
  *done = true;
  return cuda_status(cudaGetLastError(), "resize3_staged launch");
}

template <int C>
static int launch_generic(ResizeParams p, cudaStream_t stream) {
  constexpr int kGenericRows = GenericRows<C>::value;
  const size_t smem = (((size_t)p.dw * 8 + 15) & ~(size_t)15) + kGenericRows * (((size_t)p.dw * C + 15) & ~(size_t)15);
  const int max_smem = device_max_smem_optin();
  if (smem > (size_t)max_smem) {
    set_last_error("resize_linear: destination width %d needs %zu bytes of shared memory (max %d)", p.dw, smem,
                   max_smem);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  AIKO_CUDA(cudaFuncSetAttribute(resize_rows_generic<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  AIKO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, resize_rows_generic<C>, kThreads, smem));
  if (per_sm < 1) per_sm = 1;
  const long long rounds = ((long long)p.items + kGenericRows - 1) / kGenericRows;
  long long grid = (long long)sm_count() * per_sm;   // resident CTAs stride over the rounds
  if (grid > rounds) grid = rounds;
  resize_rows_generic<C><<<(unsigned)grid, kThreads, smem, stream>>>(p);
  return cuda_status(cudaGetLastError(), "resize_rows_generic launch");
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_resize_linear_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                          int64_t src_row_stride, int64_t src_frame_stride, uint8_t* dst,
                                          int dst_h, int dst_w, int64_t dst_row_stride, int64_t dst_frame_stride,
                                          int swap_rb, const aiko_b200_overlay* overlay, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0, "resize_linear: n_frames %d < 0", n_frames);
  if (n_frames == 0) return AIKO_B200_OK;  // an empty `images` list is legal (image_io.py:213-214)
  AIKO_CHECK_ARG(src && dst, "resize_linear: null frame pointer");
  AIKO_CHECK_ARG(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "resize_linear: bad size %dx%d -> %dx%d", src_w,
                 src_h, dst_w, dst_h);
  AIKO_CHECK_ARG(channels == 1 || channels == 3 || channels == 4, "resize_linear: channels %d not in {1,3,4}",
                 channels);
  AIKO_CHECK_ARG(src_row_stride >= (int64_t)src_w * channels && dst_row_stride >= (int64_t)dst_w * channels,
                 "resize_linear: row stride smaller than a row");
  AIKO_CHECK_ARG(!overlay || channels == 3, "resize_linear: fused overlay needs 3-channel frames");
  AIKO_CHECK_ARG(!swap_rb || channels >= 3, "resize_linear: swap_rb needs colour frames");
  AIKO_CHECK_ARG((int64_t)n_frames * dst_h < (int64_t)1 << 31, "resize_linear: more than 2^31 output rows");
  ResizeParams p{};
  p.src = src; p.dst = dst;
  p.n = n_frames; p.sh = src_h; p.sw = src_w; p.dh = dst_h; p.dw = dst_w;
  p.src_row_stride = src_row_stride; p.src_frame_stride = src_frame_stride;
  p.dst_row_stride = dst_row_stride; p.dst_frame_stride = dst_frame_stride;
  // cv::resize: inv_scale = (double)dsize/ssize; scale = 1./inv_scale
  p.scale_x = 1.0 / ((double)dst_w / (double)src_w);
  p.scale_y = 1.0 / ((double)dst_h / (double)src_h);
  p.swap_rb = swap_rb ? 1 : 0;
  p.has_overlay = overlay ? 1 : 0;
  if (overlay) {
    const int rc = overlay_from_abi(overlay, n_frames, static_cast<cudaStream_t>(stream), &p.ov);
    if (rc != AIKO_B200_OK) return rc;
  }
  p.items = n_frames * dst_h;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)src_row_stride | (uintptr_t)src_frame_stride |
                         (uintptr_t)((int64_t)src_w * channels)) & 15) == 0;
  if ((channels == 3 || channels == 1) && p.scale_x <= 4.0) {   // (gray, and rows off 16-byte boundaries: register-resident path only)
    // Default: resize (+ swap) with the row pipeline, then the draw list in a second, tile-parallel
    // kernel while the output frames are still in L2 -- measured faster than replaying the list
    // inside the row loop (AIKO_B200_OVERLAY_FUSED=1 selects that single-kernel variant).
    const bool two_pass = p.has_overlay && !env_int("AIKO_B200_OVERLAY_FUSED", 0);
    ResizeParams q = p;
    if (two_pass) { q.has_overlay = 0; q.keep_output = 1; }   // the overlay pass re-reads the pixels it touches
    q.unaligned = aligned ? 0 : 1;
    q.ch = channels;
    bool done = false;
    const int rc = launch_staged3(q, st, &done);
    if (rc != AIKO_B200_OK) return rc;
    if (done) return two_pass ? launch_overlay_tiles(dst, n_frames, dst_h, dst_w, dst_row_stride, dst_frame_stride, p.ov, st)
                              : AIKO_B200_OK;
  }
  switch (channels) {
    case 1: return launch_generic<1>(p, st);
    case 3: return launch_generic<3>(p, st);
    default: return launch_generic<4>(p, st);
  }
}

// Which source rows does aiko_b200_resize_linear_u8 read for src_h -> dst_h?  (Same arithmetic as
// vertical_coef above, on the host.)  needed_host[y] = 1 for every row that carries weight.
extern "C" int aiko_b200_resize_linear_rows_host(int src_h, int dst_h, uint8_t* needed_host) {
  AIKO_CHECK_ARG(src_h > 0 && dst_h > 0 && needed_host, "resize_linear_rows: bad argument");
  const double scale = 1.0 / ((double)dst_h / (double)src_h);
  for (int y = 0; y < src_h; y++) needed_host[y] = 0;
  for (int dy = 0; dy < dst_h; dy++) {
    volatile double pos = ((double)dy + 0.5) * scale;      // (volatile: every step rounded to double, as on the device)
    pos = pos + -0.5;
    const float f = (float)pos;
    const int s = (int)floorf(f);
    const int y0 = s < 0 ? 0 : (s > src_h - 1 ? src_h - 1 : s);
    const int y1 = s + 1 < 0 ? 0 : (s + 1 > src_h - 1 ? src_h - 1 : s + 1);
    needed_host[y0] = 1;
    needed_host[y1] = 1;
  }
  return AIKO_B200_OK;
}

// Host -> device copy of the listed row runs of every frame of a batch (runs_host: first row, row count
// pairs, ascending).  The frames are contiguous, so the batch is one long list of rows: runs that touch across a
// frame boundary are merged, and consecutive runs of equal length at equal distance become ONE strided copy
// (1080 -> 480 rows: every ninth source row carries no weight, the whole 32-frame batch is a head run, one
// strided copy of 3839 runs of 8 rows, and a tail run -- the last two lone runs of equal length again as one strided
// copy of two runs).  Patterns that would need more than 8 copies are copied whole.  Rows that are not listed are left as they are.
extern "C" int aiko_b200_copy_rows_h2d(uint8_t* dst, const uint8_t* src_host, int n_frames, int64_t frame_bytes,
                                       int64_t row_bytes, const int32_t* runs_host, int n_runs, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && n_runs >= 0 && frame_bytes > 0 && row_bytes > 0, "copy_rows: bad size");
  if (n_frames == 0 || n_runs == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(dst && src_host && runs_host, "copy_rows: null pointer");
  AIKO_CHECK_ARG(frame_bytes % row_bytes == 0, "copy_rows: frames must be whole rows (contiguous frames)");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t rows = frame_bytes / row_bytes;
  struct Piece { int64_t first, count, stride, repeats; };
  std::vector<Piece> pieces;
  int64_t run_first = -1, run_count = 0;               // the run being assembled, in batch rows
  auto close_run = [&]() {
    if (run_count == 0) return;
    if (!pieces.empty()) {
      Piece& p = pieces.back();
      const int64_t last_first = p.first + (p.repeats - 1) * p.stride;
      if (p.count == run_count && (p.repeats == 1 || run_first - last_first == p.stride)) {
        if (p.repeats == 1) p.stride = run_first - p.first;
        p.repeats++;
        run_count = 0;
        return;
      }
    }
    pieces.push_back({run_first, run_count, 0, 1});
    run_count = 0;
  };
  for (int f = 0; f < n_frames; f++)
    for (int r = 0; r < n_runs; r++) {
      const int64_t first = runs_host[2 * r], count = runs_host[2 * r + 1];
      AIKO_CHECK_ARG(first >= 0 && count > 0 && first + count <= rows, "copy_rows: run %d outside the frame", r);
      const int64_t at = (int64_t)f * rows + first;
      if (run_count > 0 && at == run_first + run_count) { run_count += count; continue; }
      close_run();
      run_first = at;
      run_count = count;
    }
  close_run();
  // Two lone runs of equal length (typically the head and the tail run either side of the strided piece) are one
  // strided copy of two runs: every copy command costs the DMA engine a few microseconds, which counts when the
  // frames of 64 streams are uploaded one by one.
  for (size_t i = 0; i < pieces.size(); i++)
    for (size_t j = i + 1; j < pieces.size(); j++)
      if (pieces[i].repeats == 1 && pieces[j].repeats == 1 && pieces[i].count == pieces[j].count) {
        pieces[i].stride = pieces[j].first - pieces[i].first;
        pieces[i].repeats = 2;
        pieces.erase(pieces.begin() + (long)j);
        break;
      }
  if (pieces.size() > 8) {                              // too fragmented: one contiguous copy of everything
    AIKO_CUDA(cudaMemcpyAsync(dst, src_host, (size_t)(frame_bytes * n_frames), cudaMemcpyHostToDevice, st));
    return AIKO_B200_OK;
  }
  for (const Piece& p : pieces) {
    const size_t offset = (size_t)(p.first * row_bytes), width = (size_t)(p.count * row_bytes);
    if (p.repeats == 1)
      AIKO_CUDA(cudaMemcpyAsync(dst + offset, src_host + offset, width, cudaMemcpyHostToDevice, st));
    else
      AIKO_CUDA(cudaMemcpy2DAsync(dst + offset, (size_t)(p.stride * row_bytes), src_host + offset, (size_t)(p.stride * row_bytes),
                                  width, (size_t)p.repeats, cudaMemcpyHostToDevice, st));
  }
  return AIKO_B200_OK;
}
