// Colour conversion, flip and crop call sites of the reference's media elements
// (R/elements/media/video_io.py:129,299; webcam_io.py:122-129; image_io.py:115-117,158-160).
// All are single-pass HBM-bound byte shuffles: each thread owns 16 pixels = 48 bytes =
// three 128-bit vectors (lcm of the 3-byte pixel and the 16-byte vector), so every global
// access is a full uint8x16 load or store.
#include "common.cuh"

namespace aiko {

__device__ __forceinline__ uint32_t bp(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }

// 12 bytes (4 pixels) in words a,b,c:  a = p0c0 p0c1 p0c2 p1c0 ; b = p1c1 p1c2 p2c0 p2c1 ; c = p2c2 p3c0 p3c1 p3c2
__device__ __forceinline__ void swap12(uint32_t a, uint32_t b, uint32_t c, uint32_t* oa, uint32_t* ob, uint32_t* oc) {
  // out bytes: p0c2 p0c1 p0c0 p1c2 | p1c1 p1c0 p2c2 p2c1 | p2c0 p3c2 p3c1 p3c0
  *oa = bp(a, b, 0x5012);                // a2 a1 a0 b1
  const uint32_t t = bp(a, c, 0x4003);   // a3 a0(dummy) a0(dummy) c0 -> bytes: [a3, a0, a0, c0]
  *ob = bp(b, t, 0x3740);                // b0 t0(a3) t3(c0) b3   => p1c1 p1c0 p2c2 p2c1
  *oc = bp(c, b, 0x1236);                // b2 c3 c2 c1           => p2c0 p3c2 p3c1 p3c0
}

__global__ void __launch_bounds__(kThreads) swap_rb_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                           long long n_groups, long long n_pixels, int vec_ok) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups; g += stride) {
    if (vec_ok && (g + 1) * 16 <= n_pixels) {
      const uint4* s = reinterpret_cast<const uint4*>(src) + g * 3;
      uint4 v0 = ldg_stream(s), v1 = ldg_stream(s + 1), v2 = ldg_stream(s + 2);
      uint4 o0, o1, o2;
      swap12(v0.x, v0.y, v0.z, &o0.x, &o0.y, &o0.z);
      swap12(v0.w, v1.x, v1.y, &o0.w, &o1.x, &o1.y);
      swap12(v1.z, v1.w, v2.x, &o1.z, &o1.w, &o2.x);
      swap12(v2.y, v2.z, v2.w, &o2.y, &o2.z, &o2.w);
      uint4* d = reinterpret_cast<uint4*>(dst) + g * 3;
      stg_stream(d, o0); stg_stream(d + 1, o1); stg_stream(d + 2, o2);
    } else {
      const long long pe = min((g + 1) * 16, n_pixels);
      for (long long px = g * 16; px < pe; px++) {
        const uint8_t a = src[3 * px], b = src[3 * px + 1], c = src[3 * px + 2];
        dst[3 * px] = c; dst[3 * px + 1] = b; dst[3 * px + 2] = a;
      }
    }
  }
}

// cv2 BGR2GRAY, 15-bit fixed point
__device__ __forceinline__ uint32_t gray_of(uint32_t c0, uint32_t c1, uint32_t c2, int rgb_order) {
  const uint32_t b = rgb_order ? c2 : c0, r = rgb_order ? c0 : c2;
  return (b * 3735u + c1 * 19235u + r * 9798u + 16384u) >> 15;
}
__device__ __forceinline__ uint32_t gray4(uint32_t a, uint32_t b, uint32_t c, int rgb) {
  const uint32_t g0 = gray_of(a & 0xff, (a >> 8) & 0xff, (a >> 16) & 0xff, rgb);
  const uint32_t g1 = gray_of(a >> 24, b & 0xff, (b >> 8) & 0xff, rgb);
  const uint32_t g2 = gray_of((b >> 16) & 0xff, b >> 24, c & 0xff, rgb);
  const uint32_t g3 = gray_of((c >> 8) & 0xff, (c >> 16) & 0xff, c >> 24, rgb);
  return g0 | (g1 << 8) | (g2 << 16) | (g3 << 24);
}

__global__ void __launch_bounds__(kThreads) bgr2gray_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                            long long n_groups, long long n_pixels, int rgb_order,
                                                            int vec_ok) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups; g += stride) {
    if (vec_ok && (g + 1) * 16 <= n_pixels) {
      const uint4* s = reinterpret_cast<const uint4*>(src) + g * 3;
      const uint4 v0 = ldg_stream(s), v1 = ldg_stream(s + 1), v2 = ldg_stream(s + 2);
      uint4 o;
      o.x = gray4(v0.x, v0.y, v0.z, rgb_order);
      o.y = gray4(v0.w, v1.x, v1.y, rgb_order);
      o.z = gray4(v1.z, v1.w, v2.x, rgb_order);
      o.w = gray4(v2.y, v2.z, v2.w, rgb_order);
      stg_stream(reinterpret_cast<uint4*>(dst) + g, o);
    } else {
      const long long pe = min((g + 1) * 16, n_pixels);
      for (long long px = g * 16; px < pe; px++)
        dst[px] = (uint8_t)gray_of(src[3 * px], src[3 * px + 1], src[3 * px + 2], rgb_order);
    }
  }
}

__device__ __forceinline__ void rep4(uint32_t g, uint32_t* a, uint32_t* b, uint32_t* c) {
  *a = bp(g, g, 0x1000);  // g0 g0 g0 g1
  *b = bp(g, g, 0x2211);  // g1 g1 g2 g2
  *c = bp(g, g, 0x3332);  // g2 g3 g3 g3
}

__global__ void __launch_bounds__(kThreads) gray2bgr_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                            long long n_groups, long long n_pixels, int vec_ok) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups; g += stride) {
    if (vec_ok && (g + 1) * 16 <= n_pixels) {
      const uint4 v = ldg_stream(reinterpret_cast<const uint4*>(src) + g);
      uint4 o0, o1, o2;
      rep4(v.x, &o0.x, &o0.y, &o0.z);
      rep4(v.y, &o0.w, &o1.x, &o1.y);
      rep4(v.z, &o1.z, &o1.w, &o2.x);
      rep4(v.w, &o2.y, &o2.z, &o2.w);
      uint4* d = reinterpret_cast<uint4*>(dst) + g * 3;
      stg_stream(d, o0); stg_stream(d + 1, o1); stg_stream(d + 2, o2);
    } else {
      const long long pe = min((g + 1) * 16, n_pixels);
      for (long long px = g * 16; px < pe; px++) dst[3 * px] = dst[3 * px + 1] = dst[3 * px + 2] = src[px];
    }
  }
}

// ---------------------------------------------------------------------------------------------
// RGB / BGR <-> YUV (extension: north_star names it, the reference has no call site; the oracle is
// cv2.cvtColor(COLOR_RGB2YUV / BGR2YUV / YUV2RGB / YUV2BGR) on 8-bit images, whose 14-bit fixed-point
// arithmetic this restates -- checked against cv2 over all 2^24 colours in tests/test_oracle_golden.py):
//   Y = (4899 R + 9617 G + 1868 B + 8192) >> 14
//   U = ((B - Y) 8061 + (128 << 14) + 8192) >> 14      V = ((R - Y) 14369 + (128 << 14) + 8192) >> 14
//   R = Y + ((V - 128) 18678 + 8192 >> 14)   G = Y + ((V - 128)(-9519) + (U - 128)(-6472) + 8192 >> 14)
//   B = Y + ((U - 128) 33292 + 8192 >> 14)    each result saturated to 0..255.
// 16 pixels (three uint8x16 vectors) per thread in and out.
__device__ __forceinline__ int sat8(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }
__device__ __forceinline__ int descale14(int v) { return (v + (1 << 13)) >> 14; }

// p = c0 | c1 << 8 | c2 << 16 (memory order) -> converted pixel in the same packing
__device__ __forceinline__ uint32_t yuv_pixel(uint32_t p, int bgr_order, int inverse) {
  const int c0 = p & 0xff, c1 = (p >> 8) & 0xff, c2 = (p >> 16) & 0xff;
  if (!inverse) {
    const int r = bgr_order ? c2 : c0, g = c1, b = bgr_order ? c0 : c2;
    const int y = descale14(r * 4899 + g * 9617 + b * 1868);
    const int u = descale14((b - y) * 8061 + (128 << 14)), v = descale14((r - y) * 14369 + (128 << 14));
    return (uint32_t)sat8(y) | ((uint32_t)sat8(u) << 8) | ((uint32_t)sat8(v) << 16);
  }
  const int y = c0, u = c1 - 128, v = c2 - 128;
  const int r = sat8(y + descale14(v * 18678)), g = sat8(y + descale14(v * -9519 + u * -6472));
  const int b = sat8(y + descale14(u * 33292));
  return bgr_order ? ((uint32_t)b | ((uint32_t)g << 8) | ((uint32_t)r << 16))
                   : ((uint32_t)r | ((uint32_t)g << 8) | ((uint32_t)b << 16));
}

__global__ void __launch_bounds__(kThreads) yuv_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                       long long n_groups, long long n_pixels, int bgr_order,
                                                       int inverse, int vec_ok) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < n_groups; g += stride) {
    if (vec_ok && (g + 1) * 16 <= n_pixels) {
      const uint4* s = reinterpret_cast<const uint4*>(src) + g * 3;
      const uint4 v0 = ldg_stream(s), v1 = ldg_stream(s + 1), v2 = ldg_stream(s + 2);
      const uint32_t w[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
      uint32_t o[12];
#pragma unroll
      for (int q = 0; q < 4; q++) {        // 4 pixels = 3 words in, 3 words out
        const uint32_t a = w[3 * q], b = w[3 * q + 1], c = w[3 * q + 2];
        const uint32_t p0 = yuv_pixel(a & 0xffffffu, bgr_order, inverse);
        const uint32_t p1 = yuv_pixel(((a >> 24) | (b << 8)) & 0xffffffu, bgr_order, inverse);
        const uint32_t p2 = yuv_pixel(((b >> 16) | (c << 16)) & 0xffffffu, bgr_order, inverse);
        const uint32_t p3 = yuv_pixel(c >> 8, bgr_order, inverse);
        o[3 * q] = p0 | (p1 << 24);
        o[3 * q + 1] = (p1 >> 8) | (p2 << 16);
        o[3 * q + 2] = (p2 >> 16) | (p3 << 8);
      }
      uint4* d = reinterpret_cast<uint4*>(dst) + g * 3;
      stg_stream(d, make_uint4(o[0], o[1], o[2], o[3]));
      stg_stream(d + 1, make_uint4(o[4], o[5], o[6], o[7]));
      stg_stream(d + 2, make_uint4(o[8], o[9], o[10], o[11]));
    } else {
      const long long pe = min((g + 1) * 16, n_pixels);
      for (long long px = g * 16; px < pe; px++) {
        const uint32_t p = src[3 * px] | ((uint32_t)src[3 * px + 1] << 8) | ((uint32_t)src[3 * px + 2] << 16);
        const uint32_t r = yuv_pixel(p, bgr_order, inverse);
        dst[3 * px] = (uint8_t)r; dst[3 * px + 1] = (uint8_t)(r >> 8); dst[3 * px + 2] = (uint8_t)(r >> 16);
      }
    }
  }
}

// Horizontal flip of RGB rows whose width is a multiple of 16 pixels, entirely in registers:
// a thread loads 16 pixels (3 aligned vectors), reverses their order with byte permutes and
// stores them at the mirrored position of the (possibly vertically mirrored) row.
__device__ __forceinline__ void reverse16_rgb(const uint32_t (&w)[12], uint32_t (&o)[12]) {
  uint32_t px[16];   // pixel i in the low three bytes
#pragma unroll
  for (int i = 0; i < 16; i++) {
    const int b = 3 * i, word = b >> 2, sh = b & 3;
    const uint32_t hi = word + 1 < 12 ? w[word + 1] : 0u;
    px[i] = __byte_perm(w[word], hi, 0x3210 + 0x1111 * sh);
  }
#pragma unroll
  for (int j = 0; j < 4; j++) {   // output pixels 4j .. 4j + 3 = input pixels 15 - 4j .. 12 - 4j
    const uint32_t a = px[15 - 4 * j], b = px[14 - 4 * j], c = px[13 - 4 * j], d = px[12 - 4 * j];
    o[3 * j] = __byte_perm(a, b, 0x4210);
    o[3 * j + 1] = __byte_perm(b, c, 0x5421);
    o[3 * j + 2] = __byte_perm(c, d, 0x6542);
  }
}

__global__ void __launch_bounds__(kThreads) flip_rgb16_kernel(const uint8_t* __restrict__ src,
                                                              uint8_t* __restrict__ dst, long long n_groups,
                                                              int h, int groups_per_row, int flip_vertical) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long gi = (long long)blockIdx.x * blockDim.x + threadIdx.x; gi < n_groups; gi += stride) {
    const long long row = gi / groups_per_row;
    const int gs = (int)(gi - row * groups_per_row);
    long long drow = row;
    if (flip_vertical) {
      const long long n = row / h;
      drow = n * h + (h - 1 - (row - n * h));
    }
    const uint4* s = reinterpret_cast<const uint4*>(src) + gi * 3;
    const uint4 v0 = ldg_stream(s), v1 = ldg_stream(s + 1), v2 = ldg_stream(s + 2);
    const uint32_t w[12] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, v2.x, v2.y, v2.z, v2.w};
    uint32_t o[12];
    reverse16_rgb(w, o);
    uint4* d = reinterpret_cast<uint4*>(dst) + (drow * groups_per_row + (groups_per_row - 1 - gs)) * 3;
    stg_stream(d, make_uint4(o[0], o[1], o[2], o[3]));
    stg_stream(d + 1, make_uint4(o[4], o[5], o[6], o[7]));
    stg_stream(d + 2, make_uint4(o[8], o[9], o[10], o[11]));
  }
}

// Flip / crop: one CTA per output row; the source row (segment) goes through shared memory
// so both the global read and the global write are coalesced vectors.
__global__ void __launch_bounds__(kThreads) flip_rows_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                             long long n_rows, int h, int w, int ch, int flip_code,
                                                             int rev_offset) {
  extern __shared__ __align__(16) uint8_t srow[];
  const int rb = w * ch;
  for (long long r = blockIdx.x; r < n_rows; r += gridDim.x) {
    const long long n = r / h;
    const int y = (int)(r - n * h);
    const int sy = (flip_code <= 0) ? h - 1 - y : y;
    const uint8_t* s = src + (n * h + sy) * (long long)rb;
    uint8_t* d = dst + r * (long long)rb;
    if (flip_code == 0) {  // vertical only: straight row copy
      if ((((uintptr_t)s | (uintptr_t)d | (uintptr_t)rb) & 15) == 0) {
        for (int i = threadIdx.x; i < (rb >> 4); i += kThreads)
          stg_stream(reinterpret_cast<uint4*>(d) + i, ldg_stream(reinterpret_cast<const uint4*>(s) + i));
      } else {
        for (int i = threadIdx.x; i < rb; i += kThreads) d[i] = s[i];
      }
      continue;
    }
    // stage the source row, then emit pixels in reverse order
    const int head = (int)((uintptr_t)s & 15);
    const uint8_t* sa = s - head;  // aligned down; bytes before s are never used
    const int nv = (head + rb + 15) >> 4;
    for (int i = threadIdx.x; i < nv; i += kThreads) {
      // first / last vector may straddle the allocation edge: fall back to bytes there
      if (i == 0 || i == nv - 1) {
        for (int b = 0; b < 16; b++) {
          const int o = i * 16 + b - head;
          srow[i * 16 + b] = (o >= 0 && o < rb) ? s[o] : 0;
        }
      } else {
        reinterpret_cast<uint4*>(srow)[i] = ldg_stream(reinterpret_cast<const uint4*>(sa) + i);
      }
    }
    __syncthreads();
    const uint8_t* sp = srow + head;
    // reversed row, pixel by pixel, into a second shared buffer; then one coalesced copy out
    uint8_t* rrow = srow + rev_offset;
    if (ch == 3) {
      for (int px = threadIdx.x; px < w; px += kThreads) {
        const uint8_t* q = sp + 3 * (w - 1 - px);
        rrow[3 * px] = q[0]; rrow[3 * px + 1] = q[1]; rrow[3 * px + 2] = q[2];
      }
    } else if (ch == 1) {
      for (int px = threadIdx.x; px < w; px += kThreads) rrow[px] = sp[w - 1 - px];
    } else {
      for (int px = threadIdx.x; px < w; px += kThreads)
        for (int c = 0; c < ch; c++) rrow[px * ch + c] = sp[(w - 1 - px) * ch + c];
    }
    __syncthreads();
    store_row(d, rrow, rb);
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads) crop_rows_kernel(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst,
                                                             long long n_rows, int h, int w, int ch, int x0, int y0,
                                                             int cw, int chh) {
  const int rb = cw * ch;
  for (long long r = blockIdx.x; r < n_rows; r += gridDim.x) {
    const long long n = r / chh;
    const int y = (int)(r - n * chh);
    const uint8_t* s = src + ((n * h + y0 + y) * (long long)w + x0) * ch;
    uint8_t* d = dst + r * (long long)rb;
    if ((((uintptr_t)s | (uintptr_t)d | (uintptr_t)rb) & 15) == 0) {
      for (int i = threadIdx.x; i < (rb >> 4); i += kThreads)
        stg_stream(reinterpret_cast<uint4*>(d) + i, ldg_stream(reinterpret_cast<const uint4*>(s) + i));
    } else {
      for (int i = threadIdx.x; i < rb; i += kThreads) d[i] = s[i];
    }
  }
}

static unsigned stream_grid(long long work_items) {
  long long blocks = (work_items + kThreads - 1) / kThreads;
  // enough blocks that a launch of a few hundred MB runs without serialising many iterations per
  // thread (measured: 8 per SM -> 0.67 of the HBM roofline for swap_rb, 64 per SM -> 0.88)
  const long long cap = (long long)sm_count() * 64;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (unsigned)blocks;
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_swap_rb_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, void* stream) {
  AIKO_CHECK_ARG(n_pixels >= 0, "swap_rb: n_pixels < 0");
  if (n_pixels == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "swap_rb: null pointer");
  const long long groups = (n_pixels + 15) / 16;
  const int vec_ok = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  swap_rb_kernel<<<stream_grid(groups), kThreads, 0, (cudaStream_t)stream>>>(src, dst, groups, n_pixels, vec_ok);
  return cuda_status(cudaGetLastError(), "swap_rb launch");
}

extern "C" int aiko_b200_yuv_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int bgr_order, int inverse,
                                void* stream) {
  AIKO_CHECK_ARG(n_pixels >= 0, "yuv: n_pixels < 0");
  if (n_pixels == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "yuv: null pointer");
  const long long groups = (n_pixels + 15) / 16;
  const int vec_ok = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  yuv_kernel<<<stream_grid(groups), kThreads, 0, (cudaStream_t)stream>>>(src, dst, groups, n_pixels, bgr_order ? 1 : 0,
                                                                          inverse ? 1 : 0, vec_ok);
  return cuda_status(cudaGetLastError(), "yuv launch");
}

// Alpha blending (extension: named by BASELINE.json's north_star, no call site in the reference): an RGBA image
// composited onto every RGB frame of a batch the way PIL's frame.paste(image, (x, y), image) does it -- per byte
// t = dst * (255 - a) + src * a + 128; out = ((t >> 8) + t) >> 8 (Paste.c, BLEND8).  One thread per 4 pixels of
// a row of the covered region; the 1.2 MB of a frame-sized RGBA image stay L2-resident across the batch.
__global__ void alpha_paste_kernel(uint8_t* __restrict__ frames, long long frame_stride, int width,
                                   const uint8_t* __restrict__ rgba, int sticker_w, int sx0, int sy0, int dx0, int dy0,
                                   int region_w, int region_h) {
  const int groups = (region_w + 3) >> 2;
  const int item = (int)blockIdx.x * kThreads + (int)threadIdx.x;
  if (item >= groups * region_h) return;
  const int r = item / groups, x = (item - r * groups) * 4;
  const int npx = min(4, region_w - x);
  const uint8_t* s = rgba + ((size_t)(sy0 + r) * sticker_w + sx0 + x) * 4;
  uint8_t* d = frames + (size_t)blockIdx.y * frame_stride + ((size_t)(dy0 + r) * width + dx0 + x) * 3;
  uint32_t sp[4] = {0, 0, 0, 0};
  if (npx == 4 && (reinterpret_cast<uintptr_t>(s) & 15) == 0) {
    const uint4 q = __ldg(reinterpret_cast<const uint4*>(s));
    sp[0] = q.x; sp[1] = q.y; sp[2] = q.z; sp[3] = q.w;
  } else {
    for (int k = 0; k < npx; k++) sp[k] = __ldg(reinterpret_cast<const uint32_t*>(s) + k);
  }
  const bool words = npx == 4 && (reinterpret_cast<uintptr_t>(d) & 3) == 0;
  uint32_t dw[3] = {0, 0, 0};
  if (words) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(d);
    dw[0] = p[0]; dw[1] = p[1]; dw[2] = p[2];
  } else {
    for (int q = 0; q < npx * 3; q++) dw[q >> 2] |= (uint32_t)d[q] << (8 * (q & 3));
  }
  uint32_t out[3] = {0, 0, 0};
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const unsigned a = sp[k] >> 24;
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const int at = 3 * k + c;
      const unsigned old = (dw[at >> 2] >> (8 * (at & 3))) & 0xffu, src = (sp[k] >> (8 * c)) & 0xffu;
      const unsigned t = old * (255u - a) + src * a + 128u;
      out[at >> 2] |= ((((t >> 8) + t) >> 8) & 0xffu) << (8 * (at & 3));
    }
  }
  if (words) {
    uint32_t* p = reinterpret_cast<uint32_t*>(d);
    p[0] = out[0]; p[1] = out[1]; p[2] = out[2];
  } else {
    for (int q = 0; q < npx * 3; q++) d[q] = (uint8_t)(out[q >> 2] >> (8 * (q & 3)));
  }
}

// The same for rows that are 16-byte aligned in both images (a frame-sized or tile-aligned RGBA image): one thread
// per 16 pixels, three 16-byte words of the frame against four of the RGBA image.
__global__ void __launch_bounds__(kThreads) alpha_paste16_kernel(uint8_t* __restrict__ frames, long long frame_stride,
                                                                 int width, const uint8_t* __restrict__ rgba, int sticker_w,
                                                                 int sx0, int sy0, int dx0, int dy0, int region_w,
                                                                 int region_h) {
  const int groups = region_w >> 4;
  const int item = (int)blockIdx.x * kThreads + (int)threadIdx.x;
  if (item >= groups * region_h) return;
  const int r = item / groups, x = (item - r * groups) * 16;
  const uint4* s = reinterpret_cast<const uint4*>(rgba + ((size_t)(sy0 + r) * sticker_w + sx0 + x) * 4);
  uint4* d = reinterpret_cast<uint4*>(frames + (size_t)blockIdx.y * frame_stride + ((size_t)(dy0 + r) * width + dx0 + x) * 3);
  uint32_t sp[16], dw[12], out[12];
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const uint4 q = __ldg(s + k);
    sp[4 * k] = q.x; sp[4 * k + 1] = q.y; sp[4 * k + 2] = q.z; sp[4 * k + 3] = q.w;
  }
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const uint4 q = d[k];
    dw[4 * k] = q.x; dw[4 * k + 1] = q.y; dw[4 * k + 2] = q.z; dw[4 * k + 3] = q.w;
    out[4 * k] = out[4 * k + 1] = out[4 * k + 2] = out[4 * k + 3] = 0;
  }
#pragma unroll
  for (int k = 0; k < 16; k++) {
    const unsigned a = sp[k] >> 24;
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const int at = 3 * k + c;
      const unsigned old = (dw[at >> 2] >> (8 * (at & 3))) & 0xffu, src = (sp[k] >> (8 * c)) & 0xffu;
      const unsigned t = old * (255u - a) + src * a + 128u;
      out[at >> 2] |= ((((t >> 8) + t) >> 8) & 0xffu) << (8 * (at & 3));
    }
  }
#pragma unroll
  for (int k = 0; k < 3; k++) d[k] = make_uint4(out[4 * k], out[4 * k + 1], out[4 * k + 2], out[4 * k + 3]);
}

extern "C" int aiko_b200_alpha_paste_u8(uint8_t* frames, int n_frames, int height, int width, int64_t frame_stride,
                                        const uint8_t* rgba, int image_h, int image_w, int x, int y, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && height > 0 && width > 0 && image_h >= 0 && image_w >= 0, "alpha_paste: bad size");
  AIKO_CHECK_ARG(frame_stride >= (int64_t)height * width * 3, "alpha_paste: frame stride smaller than a frame");
  // the part of the image that lands on the frame (PIL clips the box to the frame)
  const long long x0 = x < 0 ? 0 : x, y0 = y < 0 ? 0 : y;
  const long long x1 = (long long)x + image_w < width ? (long long)x + image_w : width;
  const long long y1 = (long long)y + image_h < height ? (long long)y + image_h : height;
  if (n_frames == 0 || x1 <= x0 || y1 <= y0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(frames && rgba, "alpha_paste: null pointer");
  AIKO_CHECK_ARG((reinterpret_cast<uintptr_t>(rgba) & 3) == 0, "alpha_paste: the RGBA image must be 4-byte aligned");
  const int region_w = (int)(x1 - x0), region_h = (int)(y1 - y0);
  const bool aligned16 = region_w % 16 == 0 && ((size_t)width * 3) % 16 == 0 && (size_t)frame_stride % 16 == 0 &&
                         ((reinterpret_cast<uintptr_t>(frames) + ((size_t)y0 * width + (size_t)x0) * 3) & 15) == 0 &&
                         image_w % 4 == 0 &&
                         ((reinterpret_cast<uintptr_t>(rgba) + ((size_t)(y0 - y) * image_w + (size_t)(x0 - x)) * 4) & 15) == 0;
  if (aligned16) {
    const long long items16 = (long long)(region_w / 16) * region_h;
    alpha_paste16_kernel<<<dim3((unsigned)((items16 + kThreads - 1) / kThreads), (unsigned)n_frames), kThreads, 0,
                           (cudaStream_t)stream>>>(frames, frame_stride, width, rgba, image_w, (int)(x0 - x), (int)(y0 - y),
                                                   (int)x0, (int)y0, region_w, region_h);
    return cuda_status(cudaGetLastError(), "alpha_paste launch");
  }
  const long long items = (long long)((region_w + 3) / 4) * region_h;
  alpha_paste_kernel<<<dim3((unsigned)((items + kThreads - 1) / kThreads), (unsigned)n_frames), kThreads, 0,
                       (cudaStream_t)stream>>>(frames, frame_stride, width, rgba, image_w, (int)(x0 - x), (int)(y0 - y), (int)x0,
                                               (int)y0, region_w, region_h);
  return cuda_status(cudaGetLastError(), "alpha_paste launch");
}

extern "C" int aiko_b200_bgr2gray_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, int rgb_order,
                                     void* stream) {
  AIKO_CHECK_ARG(n_pixels >= 0, "bgr2gray: n_pixels < 0");
  if (n_pixels == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "bgr2gray: null pointer");
  const long long groups = (n_pixels + 15) / 16;
  const int vec_ok = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  bgr2gray_kernel<<<stream_grid(groups), kThreads, 0, (cudaStream_t)stream>>>(src, dst, groups, n_pixels,
                                                                               rgb_order ? 1 : 0, vec_ok);
  return cuda_status(cudaGetLastError(), "bgr2gray launch");
}

extern "C" int aiko_b200_gray2bgr_u8(const uint8_t* src, uint8_t* dst, int64_t n_pixels, void* stream) {
  AIKO_CHECK_ARG(n_pixels >= 0, "gray2bgr: n_pixels < 0");
  if (n_pixels == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "gray2bgr: null pointer");
  const long long groups = (n_pixels + 15) / 16;
  const int vec_ok = ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0;
  gray2bgr_kernel<<<stream_grid(groups), kThreads, 0, (cudaStream_t)stream>>>(src, dst, groups, n_pixels, vec_ok);
  return cuda_status(cudaGetLastError(), "gray2bgr launch");
}

extern "C" int aiko_b200_flip_u8(const uint8_t* src, uint8_t* dst, int n_frames, int h, int w, int channels,
                                 int flip_code, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && h >= 0 && w >= 0, "flip: negative size");
  if (n_frames == 0 || h == 0 || w == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst && src != dst, "flip: null or aliased pointers");
  AIKO_CHECK_ARG(channels >= 1 && channels <= 4, "flip: channels %d not in 1..4", channels);
  if (flip_code != 0 && channels == 3 && (w & 15) == 0 && ((((uintptr_t)src) | ((uintptr_t)dst)) & 15) == 0) {
    const long long groups = (long long)n_frames * h * (w / 16);
    flip_rgb16_kernel<<<stream_grid(groups), kThreads, 0, (cudaStream_t)stream>>>(src, dst, groups, h, w / 16,
                                                                                  flip_code < 0 ? 1 : 0);
    return cuda_status(cudaGetLastError(), "flip launch");
  }
  const int rev_offset = (int)((((size_t)w * channels + 48) + 15) & ~(size_t)15);   // staged row, then reversed row
  const size_t smem = (size_t)rev_offset + (((size_t)w * channels + 15) & ~(size_t)15);
  AIKO_CHECK_ARG(smem <= (size_t)device_max_smem_optin(), "flip: row of %d bytes exceeds shared memory", w * channels);
  AIKO_CUDA(cudaFuncSetAttribute(flip_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long rows = (long long)n_frames * h;
  long long grid = (long long)sm_count() * 32;
  if (grid > rows) grid = rows;
  flip_rows_kernel<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(src, dst, rows, h, w, channels,
                                                                            flip_code, rev_offset);
  return cuda_status(cudaGetLastError(), "flip launch");
}

extern "C" int aiko_b200_crop_u8(const uint8_t* src, int n_frames, int h, int w, int channels, uint8_t* dst, int x0,
                                 int y0, int crop_w, int crop_h, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0, "crop: n_frames < 0");
  AIKO_CHECK_ARG(channels >= 1 && channels <= 4, "crop: channels %d not in 1..4", channels);
  AIKO_CHECK_ARG(x0 >= 0 && y0 >= 0 && crop_w >= 0 && crop_h >= 0 && x0 + crop_w <= w && y0 + crop_h <= h,
                 "crop: window %d,%d %dx%d outside %dx%d", x0, y0, crop_w, crop_h, w, h);
  if (n_frames == 0 || crop_w == 0 || crop_h == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "crop: null pointer");
  const long long rows = (long long)n_frames * crop_h;
  long long grid = (long long)sm_count() * 32;
  if (grid > rows) grid = rows;
  crop_rows_kernel<<<(unsigned)grid, kThreads, 0, (cudaStream_t)stream>>>(src, dst, rows, h, w, channels, x0, y0,
                                                                          crop_w, crop_h);
  return cuda_status(cudaGetLastError(), "crop launch");
}
