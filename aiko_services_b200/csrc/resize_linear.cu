// Batched bilinear resize, bit-exact with cv2.resize(image, (W, H)) [INTER_LINEAR, 8-bit]
// as called by the reference's ImageResize (R/elements/media/image_io.py:216), optionally
// fused with ImageOverlay's drawing (image_io.py:127-155) and an RGB<->BGR swap
// (R/elements/media/video_io.py:299).
//
// Work item = one output row of one frame.  Persistent CTAs stride over the items.  In the
// staged variant the two source rows an output row needs are pulled into shared memory by
// the TMA engine (1-D cp.async.bulk, mbarrier completion) through a ring of `stages`
// slots, so every 32-byte sector of a needed source row crosses HBM exactly once and the
// SM's load/store units only see shared memory.  The gather variant (unaligned rows, or
// horizontal scale > 4 where most sectors of a row carry no weight) reads taps directly
// from global memory.
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
  long long items;
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

__device__ __forceinline__ int vblend(int r0, int r1, int b0, int b1) {
  const int v = (((b0 * (r0 >> 4)) >> 16) + ((b1 * (r1 >> 4)) >> 16) + 2) >> 2;
  return min(max(v, 0), 255);
}

template <int C, bool STAGED>
__global__ void __launch_bounds__(kThreads) resize_linear_rows(const ResizeParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  int2* xtab = reinterpret_cast<int2*>(smem);
  size_t off = ((size_t)p.dw * 8 + 15) & ~(size_t)15;
  uint8_t* orow = smem + off;
  off += ((size_t)p.dw * C + 15) & ~(size_t)15;
  uint32_t* mask = reinterpret_cast<uint32_t*>(smem + off);
  off += 32;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + off);
  off += 64;
  uint8_t* stage = smem + off;

  // horizontal taps and weights, once per CTA
  for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
    int s;
    float f;
    linear_coef(dx, p.scale_x, &s, &f);
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= p.sw - 1) { s = p.sw - 1; f = 0.f; }
    const int a0 = coef_q11(__fsub_rn(1.f, f)), a1 = coef_q11(f);
    xtab[dx] = make_int2(s * C, (a0 & 0xffff) | (a1 << 16));
  }
  if (STAGED && threadIdx.x == 0) {
    for (int s = 0; s < p.stages; s++) mbar_init(&bars[s], 1);
    fence_mbar_init();
  }
  __syncthreads();

  const long long G = gridDim.x;
  auto rows_of = [&](long long item, int* n, int* dy, int* y0, int* y1, int* b0, int* b1) {
    *n = (int)(item / p.dh);
    *dy = (int)(item - (long long)(*n) * p.dh);
    int sy;
    float fy;
    linear_coef(*dy, p.scale_y, &sy, &fy);
    *b0 = coef_q11(__fsub_rn(1.f, fy));
    *b1 = coef_q11(fy);
    *y0 = min(max(sy, 0), p.sh - 1);
    *y1 = min(max(sy + 1, 0), p.sh - 1);
  };
  auto issue = [&](long long item, int slot) {
    int n, dy, y0, y1, b0, b1;
    rows_of(item, &n, &dy, &y0, &y1, &b0, &b1);
    const uint8_t* f = p.src + (long long)n * p.src_frame_stride;
    uint8_t* s0 = stage + (size_t)slot * 2 * p.seg_stride;
    mbar_expect_tx(&bars[slot], 2u * (uint32_t)p.row_bytes);
    bulk_g2s(s0, f + (long long)y0 * p.src_row_stride, (uint32_t)p.row_bytes, &bars[slot]);
    bulk_g2s(s0 + p.seg_stride, f + (long long)y1 * p.src_row_stride, (uint32_t)p.row_bytes, &bars[slot]);
  };

  if (STAGED && threadIdx.x == 0) {
    for (int s = 0; s < p.stages - 1; s++) {
      const long long item = blockIdx.x + (long long)s * G;
      if (item < p.items) issue(item, s);
    }
  }

  int it = 0;
  for (long long item = blockIdx.x; item < p.items; item += G, ++it) {
    int n, dy, y0, y1, b0, b1;
    rows_of(item, &n, &dy, &y0, &y1, &b0, &b1);
    const uint8_t *r0, *r1;
    if (STAGED) {
      const int slot = it % p.stages;
      if (threadIdx.x == 0) {
        const long long nxt = item + (long long)(p.stages - 1) * G;
        if (nxt < p.items) issue(nxt, (it + p.stages - 1) % p.stages);
      }
      mbar_wait(&bars[slot], (uint32_t)((it / p.stages) & 1));
      r0 = stage + (size_t)slot * 2 * p.seg_stride;
      r1 = r0 + p.seg_stride;
    } else {
      const uint8_t* f = p.src + (long long)n * p.src_frame_stride;
      r0 = f + (long long)y0 * p.src_row_stride;
      r1 = f + (long long)y1 * p.src_row_stride;
    }

    const int last_off = (p.sw - 1) * C;
    for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
      const int2 e = xtab[dx];
      const int a0 = (int)(short)(e.y & 0xffff), a1 = e.y >> 16;
      if (STAGED && C == 3) {
        // 6 source bytes per row (two adjacent RGB taps) from three aligned words
        const int sh8 = (e.x & 3) * 8;
        const uint32_t* w0 = reinterpret_cast<const uint32_t*>(r0 + (e.x & ~3));
        const uint32_t* w1 = reinterpret_cast<const uint32_t*>(r1 + (e.x & ~3));
        const uint32_t u0 = w0[0], u1 = w0[1], u2 = w0[2];
        const uint32_t v0 = w1[0], v1 = w1[1], v2 = w1[2];
        const uint32_t ulo = __funnelshift_r(u0, u1, sh8), uhi = __funnelshift_r(u1, u2, sh8);
        const uint32_t vlo = __funnelshift_r(v0, v1, sh8), vhi = __funnelshift_r(v1, v2, sh8);
        // lo = t0c0 t0c1 t0c2 t1c0 ; hi = t1c1 t1c2 ..
        int o[3];
        o[0] = vblend((ulo & 0xff) * a0 + (ulo >> 24) * a1, (vlo & 0xff) * a0 + (vlo >> 24) * a1, b0, b1);
        o[1] = vblend(((ulo >> 8) & 0xff) * a0 + (uhi & 0xff) * a1, ((vlo >> 8) & 0xff) * a0 + (vhi & 0xff) * a1,
                      b0, b1);
        o[2] = vblend(((ulo >> 16) & 0xff) * a0 + ((uhi >> 8) & 0xff) * a1,
                      ((vlo >> 16) & 0xff) * a0 + ((vhi >> 8) & 0xff) * a1, b0, b1);
        uint8_t* q = orow + dx * 3;
        if (p.swap_rb) { q[0] = o[2]; q[1] = o[1]; q[2] = o[0]; }
        else { q[0] = o[0]; q[1] = o[1]; q[2] = o[2]; }
      } else {
        const int o0 = e.x, o1 = min(e.x + C, last_off);
#pragma unroll
        for (int c = 0; c < C; c++) {
          int t00, t01, t10, t11;
          if (STAGED) {
            t00 = r0[o0 + c]; t01 = r0[o1 + c]; t10 = r1[o0 + c]; t11 = r1[o1 + c];
          } else {
            t00 = __ldg(r0 + o0 + c); t01 = __ldg(r0 + o1 + c);
            t10 = __ldg(r1 + o0 + c); t11 = __ldg(r1 + o1 + c);
          }
          const int v = vblend(t00 * a0 + t01 * a1, t10 * a0 + t11 * a1, b0, b1);
          const int cc = (p.swap_rb && C >= 3 && c < 3) ? 2 - c : c;
          orow[dx * C + cc] = (uint8_t)v;
        }
      }
    }
    __syncthreads();
    if (C == 3 && p.has_overlay) overlay_apply_row(p.ov, n, dy, p.dw, orow, mask);
    store_row(p.dst + (long long)n * p.dst_frame_stride + (long long)dy * p.dst_row_stride, orow, p.dw * C);
    __syncthreads();  // orow and the consumed stage slot are free again
  }
}

template <int C>
static int launch_resize(const ResizeParams& p0, bool staged, cudaStream_t stream) {
  ResizeParams p = p0;
  size_t smem = (((size_t)p.dw * 8 + 15) & ~(size_t)15) + (((size_t)p.dw * C + 15) & ~(size_t)15) + 32 + 64;
  const int max_smem = device_max_smem_optin();
  if (staged) {
    p.row_bytes = p.sw * C;
    p.seg_stride = p.row_bytes + 16;  // tap reads may run a few bytes past the row end
    int stages = 4;
    while (stages > 2 && smem + (size_t)stages * 2 * p.seg_stride > (size_t)56 * 1024) stages--;
    p.stages = stages;
    if (smem + (size_t)stages * 2 * p.seg_stride > (size_t)max_smem) staged = false;
    else smem += (size_t)stages * 2 * p.seg_stride;
  }
  if (smem > (size_t)max_smem) {
    set_last_error("resize_linear: destination width %d needs %zu bytes of shared memory (max %d)", p.dw, smem,
                   max_smem);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  auto kern = staged ? resize_linear_rows<C, true> : resize_linear_rows<C, false>;
  AIKO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  AIKO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kThreads, smem));
  if (per_sm < 1) per_sm = 1;
  long long grid = (long long)sm_count() * per_sm;
  if (grid > p.items) grid = p.items;
  kern<<<(unsigned)grid, kThreads, smem, stream>>>(p);
  return cuda_status(cudaGetLastError(), "resize_linear launch");
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
    AIKO_CHECK_ARG(overlay->ops && overlay->op_offsets, "resize_linear: overlay without ops/op_offsets");
    p.ov = OverlayDev{overlay->ops, overlay->op_offsets, overlay->patch_ids, overlay->luts, overlay->patch_w,
                      overlay->patch_h};
  }
  p.items = (long long)n_frames * dst_h;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)src_row_stride | (uintptr_t)src_frame_stride |
                         (uintptr_t)((int64_t)src_w * channels)) & 15) == 0;
  const bool staged = aligned && p.scale_x <= 4.0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (channels) {
    case 1: return launch_resize<1>(p, staged, st);
    case 3: return launch_resize<3>(p, staged, st);
    default: return launch_resize<4>(p, staged, st);
  }
}
