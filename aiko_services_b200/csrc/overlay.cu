// Stand-alone ImageOverlay on device-resident frames (R/elements/media/image_io.py:107-163).
// One CTA per image row: the row is staged in shared memory as 3-channel pixels, the
// frame's draw list is replayed on it in order, and the row is written back.  Grayscale
// frames take the reference's GRAY->BGR, draw, BGR->GRAY round trip (image_io.py:115,158).
#include "overlay.cuh"

namespace aiko {

struct OverlayParams {
  const uint8_t* src;
  uint8_t* dst;
  int n, h, w, channels;
  long long row_stride, frame_stride;
  OverlayDev ov;
};

__global__ void __launch_bounds__(kThreads) overlay_rows_kernel(const OverlayParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* row = smem;
  const long long rows = (long long)p.n * p.h;
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const int n = (int)(r / p.h), y = (int)(r - (long long)n * p.h);
    const long long o = (long long)n * p.frame_stride + (long long)y * p.row_stride;
    const uint8_t* s = p.src + o;
    if (p.channels == 3) {
      const int rb = p.w * 3;
      if ((((uintptr_t)s | (uintptr_t)rb) & 15) == 0) {
        for (int i = threadIdx.x; i < (rb >> 4); i += kThreads)
          reinterpret_cast<uint4*>(row)[i] = ldg_stream(reinterpret_cast<const uint4*>(s) + i);
      } else {
        for (int i = threadIdx.x; i < rb; i += kThreads) row[i] = s[i];
      }
    } else {
      for (int x = threadIdx.x; x < p.w; x += kThreads) row[3 * x] = row[3 * x + 1] = row[3 * x + 2] = s[x];
    }
    __syncthreads();
    const int ob = p.ov.op_offsets[n];
    overlay_apply_row(p.ov.ops + ob, p.ov.op_offsets[n + 1] - ob, p.ov, y, p.w, row, kThreads);
    __syncthreads();
    if (p.channels == 3) {
      store_row(p.dst + o, row, p.w * 3);
    } else {
      // the reference draws on a BGR copy: memory channel 0 = B, 1 = G, 2 = R
      uint8_t* d = p.dst + o;
      for (int x = threadIdx.x; x < p.w; x += kThreads)
        d[x] = (uint8_t)((row[3 * x] * 3735u + row[3 * x + 1] * 19235u + row[3 * x + 2] * 9798u + 16384u) >> 15);
    }
    __syncthreads();
  }
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_overlay_u8(const uint8_t* src, uint8_t* dst, int n_frames, int h, int w, int channels,
                                    int64_t row_stride, int64_t frame_stride, const aiko_b200_overlay* overlay,
                                    void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && h >= 0 && w >= 0, "overlay: negative size");
  if (n_frames == 0 || h == 0 || w == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "overlay: null frame pointer");
  AIKO_CHECK_ARG(channels == 1 || channels == 3, "overlay: channels %d not in {1,3}", channels);
  AIKO_CHECK_ARG(overlay && overlay->ops && overlay->op_offsets, "overlay: missing draw list");
  AIKO_CHECK_ARG(row_stride >= (int64_t)w * channels, "overlay: row stride smaller than a row");
  OverlayParams p{};
  p.src = src; p.dst = dst; p.n = n_frames; p.h = h; p.w = w; p.channels = channels;
  p.row_stride = row_stride; p.frame_stride = frame_stride;
  p.ov = OverlayDev{overlay->ops, overlay->op_offsets, overlay->patch_ids, overlay->luts, overlay->patch_w,
                    overlay->patch_h, overlay->max_ops_per_frame, overlay->index,
                      overlay->index_width, overlay->index_segments, overlay->index_words,
                      overlay->index_frame_words};
  const size_t smem = ((size_t)w * 3 + 15) & ~(size_t)15;
  AIKO_CHECK_ARG(smem <= (size_t)device_max_smem_optin(), "overlay: row of %zu bytes exceeds shared memory", smem);
  AIKO_CUDA(cudaFuncSetAttribute(overlay_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long rows = (long long)n_frames * h;
  long long grid = (long long)sm_count() * 8;
  if (grid > rows) grid = rows;
  overlay_rows_kernel<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  return cuda_status(cudaGetLastError(), "overlay launch");
}
