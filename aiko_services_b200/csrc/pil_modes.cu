// What PIL.Image.resize does around its filter for modes other than RGB / L
// (R/elements/media/image_io.py:218 calls image.resize((W, H)) on whatever mode the file had):
//   RGBA, LA : the image is premultiplied by alpha before the antialiased bicubic filter and
//              un-premultiplied after it (Pillow: Image.resize converts to "RGBa" / "La" and back;
//              Convert.c rgbA2rgba, rgba2rgbA, lA2la, la2lA).  LA pixels travel as (L, L, L, A),
//              exactly as Pillow stores them.
//   P, 1     : nearest neighbour (Image.resize forces NEAREST for palette and bilevel images):
//              out[y][x] = in[ytab[y]][xtab[x]] with the source positions Pillow's ImagingScaleAffine
//              pretabulates in double precision -- the caller computes the two tables the same way.
// Elementwise / gather byte work, HBM-bound.
#include "common.cuh"

namespace aiko {

// Pillow's MULDIV255(a, b): (a * b + 128 + ((a * b + 128) >> 8)) >> 8
__device__ __forceinline__ uint32_t muldiv255(uint32_t a, uint32_t b) {
  const uint32_t t = a * b + 128u;
  return ((t >> 8) + t) >> 8;
}
__device__ __forceinline__ uint32_t unmul(uint32_t v, uint32_t alpha) {     // CLIP8((255 * v) / alpha)
  const uint32_t q = (255u * v) / alpha;
  return q > 255u ? 255u : q;
}

// src: [n_pixels][C], C in {2, 4} -> dst [n_pixels][4] premultiplied (L, L, L, A) or (R, G, B, A)
__global__ void __launch_bounds__(kThreads) premultiply_kernel(const uint8_t* __restrict__ src,
                                                               uint32_t* __restrict__ dst, long long n, int c) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint32_t r, g, b, a;
    if (c == 4) {
      const uint32_t v = reinterpret_cast<const uint32_t*>(src)[i];
      r = v & 0xff; g = (v >> 8) & 0xff; b = (v >> 16) & 0xff; a = v >> 24;
    } else {
      r = g = b = src[2 * i]; a = src[2 * i + 1];
    }
    dst[i] = muldiv255(r, a) | (muldiv255(g, a) << 8) | (muldiv255(b, a) << 16) | (a << 24);
  }
}

// src: [n_pixels][4] premultiplied -> dst [n_pixels][C], C in {2, 4}
__global__ void __launch_bounds__(kThreads) unpremultiply_kernel(const uint32_t* __restrict__ src,
                                                                 uint8_t* __restrict__ dst, long long n, int c) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const uint32_t v = src[i];
    uint32_t r = v & 0xff, g = (v >> 8) & 0xff, b = (v >> 16) & 0xff;
    const uint32_t a = v >> 24;
    if (a != 255u && a != 0u) { r = unmul(r, a); g = unmul(g, a); b = unmul(b, a); }
    if (c == 4) {
      reinterpret_cast<uint32_t*>(dst)[i] = r | (g << 8) | (b << 16) | (a << 24);
    } else {
      dst[2 * i] = (uint8_t)r; dst[2 * i + 1] = (uint8_t)a;
    }
  }
}

__global__ void __launch_bounds__(kThreads) gather_nearest_kernel(const uint8_t* __restrict__ src, int n, int sh, int sw,
                                                                  int c, uint8_t* __restrict__ dst, int dh, int dw,
                                                                  const int* __restrict__ xtab,
                                                                  const int* __restrict__ ytab) {
  const long long total = (long long)n * dh * dw;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int x = (int)(i % dw);
    const long long t = i / dw;
    const int y = (int)(t % dh), f = (int)(t / dh);
    const int sx = __ldg(xtab + x), sy = __ldg(ytab + y);
    const bool inside = sx >= 0 && sx < sw && sy >= 0 && sy < sh;       // outside: Pillow's fill (0)
    const uint8_t* s = src + (((long long)f * sh + (inside ? sy : 0)) * sw + (inside ? sx : 0)) * c;
    uint8_t* d = dst + i * c;
    for (int k = 0; k < c; k++) d[k] = inside ? s[k] : (uint8_t)0;
  }
}

static unsigned grid_for(long long items) {
  long long g = (items + kThreads - 1) / kThreads;
  const long long cap = (long long)sm_count() * 16;
  return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_premultiply_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int channels, int inverse,
                                        void* stream) {
  AIKO_CHECK_ARG(n_pixels >= 0, "premultiply: n_pixels < 0");
  if (n_pixels == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "premultiply: null pointer");
  AIKO_CHECK_ARG(channels == 2 || channels == 4, "premultiply: channels %d not in {2 (LA), 4 (RGBA)}", channels);
  const uintptr_t four = reinterpret_cast<uintptr_t>(inverse ? src : dst) | (channels == 4 ? reinterpret_cast<uintptr_t>(inverse ? dst : src) : 0);
  AIKO_CHECK_ARG((four & 3) == 0, "premultiply: 4-channel buffers must be 4-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (inverse)
    unpremultiply_kernel<<<grid_for(n_pixels), kThreads, 0, st>>>(reinterpret_cast<const uint32_t*>(src), dst, n_pixels, channels);
  else
    premultiply_kernel<<<grid_for(n_pixels), kThreads, 0, st>>>(src, reinterpret_cast<uint32_t*>(dst), n_pixels, channels);
  return cuda_status(cudaGetLastError(), "premultiply launch");
}

extern "C" int aiko_b200_gather_nearest_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                           uint8_t* dst, int dst_h, int dst_w, const int32_t* xtab, const int32_t* ytab,
                                           void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "gather_nearest: bad size");
  if (n_frames == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst && xtab && ytab, "gather_nearest: null pointer");
  AIKO_CHECK_ARG(channels >= 1 && channels <= 4, "gather_nearest: channels %d not in 1..4", channels);
  gather_nearest_kernel<<<grid_for((long long)n_frames * dst_h * dst_w), kThreads, 0, static_cast<cudaStream_t>(stream)>>>(
      src, n_frames, src_h, src_w, channels, dst, dst_h, dst_w, xtab, ytab);
  return cuda_status(cudaGetLastError(), "gather_nearest launch");
}
