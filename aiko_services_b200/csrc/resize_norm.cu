// BASELINE config 4 (extension; the reference has no such element -- its closest consumer hands
// resized ndarrays to a torch model, R/examples/yolo/yolo.py:62-64): bilinear resize with cv2's
// exact 8-bit arithmetic (R/elements/media/image_io.py:216), then the model's input transform
//   out[n][c][y][x] = ((float)u8 / 255 - mean[c]) / std[c]
// written as fp32 NCHW.  Oracle: numpy restatement (oracle/image_oracle.c: orc_resize_norm_f32).
//
// Large downscales (4K -> 224x224) touch 2x2 taps out of every ~17x10 source pixels, so taps
// are gathered straight from global memory (only the sectors that carry weight are read); one
// CTA per output row at a time, the three planes written with fully coalesced 4-byte stores.
#include "common.cuh"

namespace aiko {

__device__ __forceinline__ void linear_tap(int d, double scale, int src, bool clamp_phase, int* s0, int* s1, int* w0,
                                           int* w1) {
  const double pos = __dadd_rn(__dmul_rn(__dadd_rn((double)d, 0.5), scale), -0.5);
  float f = __double2float_rn(pos);
  int s = __float2int_rd(f);
  f = __fsub_rn(f, (float)s);
  if (clamp_phase) {                       // horizontal: cv2 zeroes the phase at the borders
    if (s < 0) { s = 0; f = 0.f; }
    if (s >= src - 1) { s = src - 1; f = 0.f; }
  }
  *w0 = __float2int_rn(__fmul_rn(__fsub_rn(1.f, f), 2048.f));
  *w1 = __float2int_rn(__fmul_rn(f, 2048.f));
  *s0 = min(max(s, 0), src - 1);
  *s1 = min(max(s + 1, 0), src - 1);
}

struct NormParams {
  const uint8_t* src;
  float* dst;
  int n, sh, sw, dh, dw;
  double scale_x, scale_y;
  float mean[3], stdv[3];
};

// A CTA strides over output rows (frame, dy).  The horizontal taps of every output column are
// worked out once per CTA into shared memory (they are the same for every row), the vertical tap
// once per row; per pixel what is left is 12 byte loads, the fixed-point blend and the three
// IEEE divisions of the transform.
__global__ void __launch_bounds__(kThreads) resize_norm_kernel(const NormParams p) {
  extern __shared__ int4 xtap[];             // x0 * 3, x1 * 3, a0, a1 per output column
  for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
    int x0, x1, a0, a1;
    linear_tap(dx, p.scale_x, p.sw, true, &x0, &x1, &a0, &a1);
    xtap[dx] = make_int4(x0 * 3, x1 * 3, a0, a1);
  }
  __syncthreads();
  const size_t plane = (size_t)p.dh * p.dw;
  const int rows = p.n * p.dh;
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const int n = r / p.dh, dy = r - n * p.dh;
    int y0, y1, b0, b1;
    linear_tap(dy, p.scale_y, p.sh, false, &y0, &y1, &b0, &b1);
    const uint8_t* f = p.src + (size_t)n * p.sh * p.sw * 3;
    const uint8_t* r0 = f + (size_t)y0 * p.sw * 3;
    const uint8_t* r1 = f + (size_t)y1 * p.sw * 3;
    const uint32_t b0s = (uint32_t)b0 << 16, b1s = (uint32_t)b1 << 16;
    float* orow = p.dst + (size_t)n * 3 * plane + (size_t)dy * p.dw;
    for (int dx = threadIdx.x; dx < p.dw; dx += kThreads) {
      const int4 t = xtap[dx];
      uint32_t u[3][4];
#pragma unroll
      for (int c = 0; c < 3; c++) {          // all twelve gathers first
        u[c][0] = __ldg(r0 + t.x + c); u[c][1] = __ldg(r0 + t.y + c);
        u[c][2] = __ldg(r1 + t.x + c); u[c][3] = __ldg(r1 + t.y + c);
      }
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const uint32_t h0 = u[c][0] * t.z + u[c][1] * t.w;
        const uint32_t h1 = u[c][2] * t.z + u[c][3] * t.w;
        const uint32_t v = (__umulhi(b0s, h0 >> 4) + __umulhi(b1s, h1 >> 4) + 2u) >> 2;
        orow[c * plane + dx] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.f), p.mean[c]), p.stdv[c]);
      }
    }
  }
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_resize_norm_f32_nchw(const uint8_t* src, int n_frames, int src_h, int src_w, float* dst,
                                              int dst_h, int dst_w, const float* mean_host, const float* std_host,
                                              void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0, "resize_norm: n_frames < 0");
  if (n_frames == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst && mean_host && std_host, "resize_norm: null pointer");
  AIKO_CHECK_ARG(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "resize_norm: bad size");
  NormParams p{};
  p.src = src; p.dst = dst; p.n = n_frames; p.sh = src_h; p.sw = src_w; p.dh = dst_h; p.dw = dst_w;
  p.scale_x = 1.0 / ((double)dst_w / (double)src_w);
  p.scale_y = 1.0 / ((double)dst_h / (double)src_h);
  for (int c = 0; c < 3; c++) {
    AIKO_CHECK_ARG(std_host[c] != 0.f, "resize_norm: std[%d] is zero", c);
    p.mean[c] = mean_host[c];
    p.stdv[c] = std_host[c];
  }
  AIKO_CHECK_ARG((long long)n_frames * dst_h < ((long long)1 << 31), "resize_norm: more than 2^31 output rows");
  const size_t smem = (size_t)dst_w * sizeof(int4);
  if (smem > (size_t)device_max_smem_optin()) {
    set_last_error("resize_norm: destination width %d needs %zu bytes of shared memory", dst_w, smem);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  AIKO_CUDA(cudaFuncSetAttribute(resize_norm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  long long blocks = (long long)n_frames * dst_h;
  const long long cap = (long long)sm_count() * 16;
  if (blocks > cap) blocks = cap;
  resize_norm_kernel<<<(unsigned)blocks, kThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return cuda_status(cudaGetLastError(), "resize_norm launch");
}
