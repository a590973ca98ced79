// Baseline JPEG encoding on the GPU, hand-written: frames LEAVE the device compressed
// and the files are byte-identical to what the reference writes.  The reference encodes inside
// ImageWriteFile's image.save(path) (R/elements/media/image_io.py:250), i.e. Pillow driving libjpeg-turbo with its
// defaults: quality 75, 4:2:0 for RGB images, standard Huffman tables, JFIF header.  This file restates that
// pipeline (oracle: oracle/jpeg_encode_oracle.c, itself pinned against the bytes Pillow writes):
//
//   kernel 0: jpeg_planes_kernel  -- RGB -> YCbCr (jccolor.c 16-bit fixed point) and h2v1 / h2v2 box downsampling
//             with libjpeg's alternating bias and replicated edges, into one plane per component.
//   kernel 1: jpeg_fdct_kernel    -- one thread per 8x8 block in scan (MCU) order: level shift, the accurate integer
//             forward DCT (jfdctint.c), quantisation (jcdctmgr.c: rounded division, done with an exact 40-bit
//             reciprocal); coefficients stored in zigzag order.  Blocks outside the component's real block grid
//             ("dummy" blocks that complete an MCU) are zero apart from the DC they inherit in kernel 2.
//   kernel 2: jpeg_entropy_kernel -- one CTA per image, every thread owns a contiguous run of blocks: pass 1 sums
//             the code lengths of its run (jchuff.c encode_one_block), a block scan turns them into bit offsets,
//             pass 2 writes the codes at its offset (whole words with plain stores, the two words it shares with
//             its neighbours with atomicOr); the last thread pads with 1 bits.
//   kernel 3: jpeg_finish_kernel  -- one CTA per image: header, the scan with a zero stuffed after every FF byte
//             (count pass, block scan, write pass), EOI, the file's length.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "jpeg_std_tables.inc"

namespace aiko {

namespace {

struct EncGeometry {
  int W, H, ncomp, hs, vs;
  int mcux, mcuy, bpm;
  int cw[3], ch[3];              // downsampled size of the component
  int bw[3], bh[3];              // its real blocks
  int ph[3], pv[3];              // blocks per MCU, horizontally / vertically
  int comp_of[6], bx_of[6], by_of[6];
  int first_q[3];                // first block of the component inside an MCU
  int n_blocks;                  // per image, scan order
  long long plane_offset[3];     // bytes from the start of an image's plane area; plane c is bh*8 rows of bw*8 bytes
  long long plane_per_image;
};

struct EncTables {               // by value as a kernel argument
  uint16_t q8[2][64];            // quantisation divisors (table value * 8), natural order
  unsigned long long recip[2][64];   // ceil(2^40 / divisor): n / d = (n * recip) >> 40 exactly for n < 2^18
  uint16_t dc_code[2][12];
  uint8_t dc_size[2][12];
  uint16_t ac_code[2][256];
  uint8_t ac_size[2][256];
};

struct EncHeader {               // by value as a kernel argument
  int length;
  uint8_t bytes[640];
};

// ---------------------------------------------------------------------------------------------
// kernel 1
#define AIKO_FDCT_DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))
#define AIKO_FDCT_1D(d0, d1, d2, d3, d4, d5, d6, d7, FIRST)                                            \
  {                                                                                                    \
    int tmp0 = (d0) + (d7), tmp7 = (d0) - (d7), tmp1 = (d1) + (d6), tmp6 = (d1) - (d6);                \
    int tmp2 = (d2) + (d5), tmp5 = (d2) - (d5), tmp3 = (d3) + (d4), tmp4 = (d3) - (d4);                \
    const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;      \
    constexpr int SH = (FIRST) ? 11 : 15;                                                              \
    if (FIRST) { (d0) = (tmp10 + tmp11) * 4; (d4) = (tmp10 - tmp11) * 4; }                             \
    else { (d0) = AIKO_FDCT_DESCALE(tmp10 + tmp11, 2); (d4) = AIKO_FDCT_DESCALE(tmp10 - tmp11, 2); }   \
    int z1 = (tmp12 + tmp13) * 4433;                                                                   \
    (d2) = AIKO_FDCT_DESCALE(z1 + tmp13 * 6270, SH);                                                   \
    (d6) = AIKO_FDCT_DESCALE(z1 + tmp12 * (-15137), SH);                                               \
    z1 = tmp4 + tmp7;                                                                                  \
    int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;                                          \
    const int z5 = (z3 + z4) * 9633;                                                                   \
    tmp4 *= 2446; tmp5 *= 16819; tmp6 *= 25172; tmp7 *= 12299;                                         \
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;                                              \
    z3 += z5; z4 += z5;                                                                                \
    (d7) = AIKO_FDCT_DESCALE(tmp4 + z1 + z3, SH);                                                      \
    (d5) = AIKO_FDCT_DESCALE(tmp5 + z2 + z4, SH);                                                      \
    (d3) = AIKO_FDCT_DESCALE(tmp6 + z2 + z3, SH);                                                      \
    (d1) = AIKO_FDCT_DESCALE(tmp7 + z1 + z4, SH);                                                      \
  }

// kernel 0: component planes.  One thread per chroma sample position: its Cb and Cr samples and the hs x vs luma
// samples it covers (gray frames: one thread per sample).  Planes hold the component's real blocks; samples
// right of / below the image repeat the edge as libjpeg does (expand_right_edge on the source pixels,
// expand_bottom_edge on the colour rows and on the downsampled rows).
__device__ __forceinline__ int ycc_sample(const uint8_t* __restrict__ px, int comp) {
  const int r = px[0], g = px[1], b = px[2];
  if (comp == 0) return (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
  if (comp == 1) return (-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16;
  return (32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16;
}

template <int HS, int VS>
__global__ void __launch_bounds__(256) jpeg_planes_kernel(const uint8_t* __restrict__ src, long long frame_stride,
                                                          const EncGeometry g, uint8_t* __restrict__ planes_all) {
  const uint8_t* frame = src + (size_t)blockIdx.y * frame_stride;
  uint8_t* planes = planes_all + (size_t)blockIdx.y * g.plane_per_image;
  const int cpw = g.bw[g.ncomp == 1 ? 0 : 1] * 8, cph = g.bh[g.ncomp == 1 ? 0 : 1] * 8;   // the grid: chroma plane (gray: luma)
  const int item = (int)blockIdx.x * 256 + (int)threadIdx.x;
  if (item >= cpw * cph) return;
  const int Y = item / cpw, X = item - Y * cpw;
  if (g.ncomp == 1) {
    planes[(size_t)Y * cpw + X] = frame[(size_t)min(Y, g.H - 1) * g.W + min(X, g.W - 1)];
    return;
  }
  const int lpw = g.bw[0] * 8, lph = g.bh[0] * 8;
#pragma unroll
  for (int dy = 0; dy < VS; dy++)
#pragma unroll
    for (int dx = 0; dx < HS; dx++) {
      const int lx = X * HS + dx, ly = Y * VS + dy;
      if (lx < lpw && ly < lph)
        planes[g.plane_offset[0] + (size_t)ly * lpw + lx] =
            (uint8_t)ycc_sample(frame + ((size_t)min(ly, g.H - 1) * g.W + min(lx, g.W - 1)) * 3, 0);
    }
  const int sy = min(Y, g.ch[1] - 1);                     // below the image: the last sample row again
  int cb = 0, cr = 0;
#pragma unroll
  for (int dy = 0; dy < VS; dy++)
#pragma unroll
    for (int dx = 0; dx < HS; dx++) {
      const uint8_t* px = frame + ((size_t)min(sy * VS + dy, g.H - 1) * g.W + min(X * HS + dx, g.W - 1)) * 3;
      cb += ycc_sample(px, 1);
      cr += ycc_sample(px, 2);
    }
  if (HS == 2 && VS == 1) { const int bias = X & 1; cb = (cb + bias) >> 1; cr = (cr + bias) >> 1; }          // h2v1: 0, 1, 0, 1
  else if (HS == 2) { const int bias = (X & 1) ? 2 : 1; cb = (cb + bias) >> 2; cr = (cr + bias) >> 2; }      // h2v2: 1, 2, 1, 2
  planes[g.plane_offset[1] + (size_t)Y * cpw + X] = (uint8_t)cb;
  planes[g.plane_offset[2] + (size_t)Y * cpw + X] = (uint8_t)cr;
}

constexpr int kFdctThreads = 128;

__global__ void __launch_bounds__(kFdctThreads) jpeg_fdct_kernel(const uint8_t* __restrict__ planes_all, const EncGeometry g,
                                                                 const EncTables t, int16_t* __restrict__ coef) {
  const int blk = (int)blockIdx.x * kFdctThreads + (int)threadIdx.x;
  if (blk >= g.n_blocks) return;
  const int m = blk / g.bpm, q = blk - m * g.bpm;
  const int my = m / g.mcux, mx = m - my * g.mcux;
  const int c = g.comp_of[q];
  const int gx = mx * g.ph[c] + g.bx_of[q], gy = my * g.pv[c] + g.by_of[q];
  uint4* out = reinterpret_cast<uint4*>(coef + ((size_t)blockIdx.y * g.n_blocks + blk) * 64);
  if (gx >= g.bw[c] || gy >= g.bh[c]) {                   // a dummy block: zero, its DC is resolved by the entropy kernel
#pragma unroll
    for (int k = 0; k < 8; k++) out[k] = make_uint4(0, 0, 0, 0);
    return;
  }
  const int pitch = g.bw[c] * 8;
  const uint8_t* plane = planes_all + (size_t)blockIdx.y * g.plane_per_image + g.plane_offset[c] + (size_t)gy * 8 * pitch + gx * 8;
  int x[64];
#pragma unroll
  for (int yy = 0; yy < 8; yy++) {
    const uint2 v = __ldg(reinterpret_cast<const uint2*>(plane + (size_t)yy * pitch));
#pragma unroll
    for (int xx = 0; xx < 4; xx++) {
      x[8 * yy + xx] = (int)((v.x >> (8 * xx)) & 0xffu) - 128;
      x[8 * yy + 4 + xx] = (int)((v.y >> (8 * xx)) & 0xffu) - 128;
    }
  }
#pragma unroll
  for (int r = 0; r < 8; r++)
    AIKO_FDCT_1D(x[8 * r], x[8 * r + 1], x[8 * r + 2], x[8 * r + 3], x[8 * r + 4], x[8 * r + 5], x[8 * r + 6], x[8 * r + 7], true)
#pragma unroll
  for (int col = 0; col < 8; col++)
    AIKO_FDCT_1D(x[col], x[8 + col], x[16 + col], x[24 + col], x[32 + col], x[40 + col], x[48 + col], x[56 + col], false)
  const uint16_t* q8 = t.q8[c == 0 ? 0 : 1];
  const unsigned long long* recip = t.recip[c == 0 ? 0 : 1];
  uint32_t packed[32];
#pragma unroll
  for (int k = 0; k < 64; k++) {
    constexpr uint8_t zz[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
    const int i = zz[k];
    const int v = x[i];
    const unsigned n = (unsigned)(v < 0 ? -v : v) + (q8[i] >> 1);
    const int mag = (int)(((unsigned long long)n * recip[i]) >> 40);       // n / q8[i]
    const int qv = v < 0 ? -mag : mag;
    if (k & 1) packed[k >> 1] |= (uint32_t)(qv & 0xffff) << 16;
    else packed[k >> 1] = (uint32_t)(qv & 0xffff);
  }
#pragma unroll
  for (int k = 0; k < 8; k++) out[k] = make_uint4(packed[4 * k], packed[4 * k + 1], packed[4 * k + 2], packed[4 * k + 3]);
}

// ---------------------------------------------------------------------------------------------
// kernel 2
constexpr int kEntropyThreads = 1024;

struct EntropyShared {
  uint16_t dc_code[2][12];
  uint8_t dc_size[2][12];
  uint16_t ac_code[2][256];
  uint8_t ac_size[2][256];
  unsigned long long warp_sums[kEntropyThreads / 32];
};

__device__ __forceinline__ bool is_dummy(const EncGeometry& g, int blk) {
  const int m = blk / g.bpm, q = blk - m * g.bpm;
  const int my = m / g.mcux, mx = m - my * g.mcux;
  const int c = g.comp_of[q];
  return mx * g.ph[c] + g.bx_of[q] >= g.bw[c] || my * g.pv[c] + g.by_of[q] >= g.bh[c];
}

// DC of a block as the Huffman coder sees it: a dummy block carries the DC of the block before it in its MCU
// (jccoefct.c compress_data), which may be a dummy block itself; an MCU's first block is always real.
__device__ __forceinline__ int dc_value(const EncGeometry& g, const int16_t* __restrict__ coef, int blk) {
  while (is_dummy(g, blk)) blk--;
  return coef[(size_t)blk * 64];
}

struct BitSink {
  uint32_t* words;               // the image's scan as big-endian 32-bit words, zero-filled
  long long word;                // index of the word being assembled
  unsigned long long acc;        // pending bits, left-aligned at bit 63
  int n;                         // pending bits (incl. the leading gap of a shared first word)
  bool first;                    // the next word stored is shared with the thread before
};

__device__ __forceinline__ void sink_put(BitSink& s, unsigned code, int size) {        // size <= 26
  s.acc |= (unsigned long long)code << (64 - s.n - size);
  s.n += size;
  if (s.n >= 32) {
    const uint32_t w = __byte_perm((uint32_t)(s.acc >> 32), 0, 0x0123);
    if (s.first) { atomicOr(s.words + s.word, w); s.first = false; }
    else s.words[s.word] = w;
    s.word++;
    s.acc <<= 32;
    s.n -= 32;
  }
}

// One block (coefficients in zigzag order): WRITE = false -> its code length in bits.
template <bool WRITE>
__device__ __forceinline__ int encode_block(const EntropyShared& S, const uint4* __restrict__ blk, int table, int dc_diff,
                                            BitSink& sink) {
  int bits = 0;
  {
    int temp = dc_diff, temp2 = dc_diff;
    if (temp < 0) { temp = -temp; temp2--; }
    const int nbits = 32 - __clz(temp);
    if (WRITE) {
      sink_put(sink, ((unsigned)S.dc_code[table][nbits] << nbits) | ((unsigned)temp2 & ((1u << nbits) - 1u)),
               S.dc_size[table][nbits] + nbits);
    } else {
      bits += S.dc_size[table][nbits] + nbits;
    }
  }
  int run = 0;
#pragma unroll 1
  for (int v = 0; v < 8; v++) {
    const uint4 quad = __ldg(blk + v);
    const uint32_t w[4] = {quad.x, quad.y, quad.z, quad.w};
#pragma unroll
    for (int j = 0; j < 8; j++) {
      if (v == 0 && j == 0) continue;                      // the DC
      int temp = (int)(short)(j & 1 ? w[j >> 1] >> 16 : w[j >> 1] & 0xffffu);
      if (temp == 0) { run++; continue; }
      while (run > 15) {
        if (WRITE) sink_put(sink, S.ac_code[table][0xf0], S.ac_size[table][0xf0]);
        else bits += S.ac_size[table][0xf0];
        run -= 16;
      }
      int temp2 = temp;
      if (temp < 0) { temp = -temp; temp2--; }
      const int nbits = 32 - __clz(temp);
      const int sym = (run << 4) + nbits;
      if (WRITE) {
        sink_put(sink, ((unsigned)S.ac_code[table][sym] << nbits) | ((unsigned)temp2 & ((1u << nbits) - 1u)),
                 S.ac_size[table][sym] + nbits);
      } else {
        bits += S.ac_size[table][sym] + nbits;
      }
      run = 0;
    }
  }
  if (run > 0) {
    if (WRITE) sink_put(sink, S.ac_code[table][0], S.ac_size[table][0]);
    else bits += S.ac_size[table][0];
  }
  return bits;
}

__global__ void __launch_bounds__(kEntropyThreads) jpeg_entropy_kernel(const int16_t* __restrict__ coef_all, const EncGeometry g,
                                                                       const EncTables t, uint32_t* __restrict__ scan_all,
                                                                       long long scan_words, long long* __restrict__ scan_bits) {
  __shared__ EntropyShared S;
  const int tid = (int)threadIdx.x;
  for (int i = tid; i < 2 * 12; i += kEntropyThreads) { (&S.dc_code[0][0])[i] = (&t.dc_code[0][0])[i]; (&S.dc_size[0][0])[i] = (&t.dc_size[0][0])[i]; }
  for (int i = tid; i < 2 * 256; i += kEntropyThreads) { (&S.ac_code[0][0])[i] = (&t.ac_code[0][0])[i]; (&S.ac_size[0][0])[i] = (&t.ac_size[0][0])[i]; }
  __syncthreads();
  const int16_t* coef = coef_all + (size_t)blockIdx.x * g.n_blocks * 64;
  const int per = (g.n_blocks + kEntropyThreads - 1) / kEntropyThreads;
  const int b0 = min(tid * per, g.n_blocks), b1 = min(b0 + per, g.n_blocks);
  BitSink sink;
  sink.words = scan_all + (size_t)blockIdx.x * scan_words;
  // the difference coded for block b: its DC minus the DC of the block of the same component before it in scan order
  auto dc_diff = [&](int b) {
    const int m = b / g.bpm, q = b - m * g.bpm;
    const int c = g.comp_of[q];
    int prev = -1;
    if (q > g.first_q[c]) prev = b - 1;
    else if (m > 0) prev = b - g.bpm + g.ph[c] * g.pv[c] - 1;
    return dc_value(g, coef, b) - (prev >= 0 ? dc_value(g, coef, prev) : 0);
  };
  unsigned long long mine = 0;
  for (int b = b0; b < b1; b++) {
    const int q = b % g.bpm;
    mine += (unsigned long long)encode_block<false>(S, reinterpret_cast<const uint4*>(coef + (size_t)b * 64),
                                                    g.comp_of[q] == 0 ? 0 : 1, dc_diff(b), sink);
  }
  // exclusive scan of the bit counts over the CTA
  const int lane = tid & 31, warp = tid >> 5;
  unsigned long long incl = mine;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned long long v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += v;
  }
  if (lane == 31) S.warp_sums[warp] = incl;
  __syncthreads();
  unsigned long long before = 0, total = 0;
  for (int w = 0; w < kEntropyThreads / 32; w++) {
    const unsigned long long v = S.warp_sums[w];
    before += w < warp ? v : 0;
    total += v;
  }
  const unsigned long long start = before + incl - mine;
  const unsigned long long padded = (total + 7) & ~7ull;
  if (tid == 0) scan_bits[blockIdx.x] = padded <= (unsigned long long)scan_words * 32 ? (long long)padded : -1;
  if (padded > (unsigned long long)scan_words * 32) return;            // (uniform) the scan does not fit: reported
  sink.word = (long long)(start >> 5);
  sink.n = (int)(start & 31);
  sink.acc = 0;
  sink.first = true;
  for (int b = b0; b < b1; b++) {
    const int q = b % g.bpm;
    encode_block<true>(S, reinterpret_cast<const uint4*>(coef + (size_t)b * 64), g.comp_of[q] == 0 ? 0 : 1, dc_diff(b), sink);
  }
  const bool last = b1 == g.n_blocks && b0 < b1 || (g.n_blocks == 0 && tid == 0);
  if (last && padded > total) sink_put(sink, (1u << (int)(padded - total)) - 1u, (int)(padded - total));   // jchuff.c flush_bits
  if (sink.n > 0 && mine + (last ? padded - total : 0) > 0) {
    const uint32_t w = __byte_perm((uint32_t)(sink.acc >> 32), 0, 0x0123);
    atomicOr(sink.words + sink.word, w);
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 3
constexpr int kFinishThreads = 1024;

__global__ void __launch_bounds__(kFinishThreads) jpeg_finish_kernel(const uint8_t* __restrict__ scan_all, long long scan_bytes_cap,
                                                                     const long long* __restrict__ scan_bits, const EncHeader hdr,
                                                                     uint8_t* __restrict__ out_all, long long out_stride,
                                                                     int32_t* __restrict__ sizes) {
  __shared__ int warp_sums[kFinishThreads / 32];
  const int tid = (int)threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long bits = scan_bits[blockIdx.x];
  uint8_t* out = out_all + (size_t)blockIdx.x * out_stride;
  if (bits < 0) {                                                       // (uniform) the scan did not fit its buffer
    if (tid == 0) sizes[blockIdx.x] = -1;
    return;
  }
  const uint8_t* scan = scan_all + (size_t)blockIdx.x * scan_bytes_cap;
  const int bytes = (int)(bits >> 3);
  const int vecs = (bytes + 15) >> 4;
  const int per_warp = (vecs + kFinishThreads / 32 - 1) / (kFinishThreads / 32);
  const int v_lo = min(warp * per_warp, vecs), v_hi = min(v_lo + per_warp, vecs);
  auto ff_mask = [&](int v, uint4* q) -> unsigned {                     // bit p: byte 16 v + p is FF
    *q = __ldg(reinterpret_cast<const uint4*>(scan) + v);
    const uint32_t w[4] = {q->x, q->y, q->z, q->w};
    unsigned m = 0;
#pragma unroll
    for (int p = 0; p < 16; p++)
      if (((w[p >> 2] >> (8 * (p & 3))) & 0xffu) == 0xffu && 16 * v + p < bytes) m |= 1u << p;
    return m;
  };
  uint4 vec;
  int n_ff = 0;
  for (int v = v_lo + lane; v < v_hi; v += 32) n_ff += __popc(ff_mask(v, &vec));
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) n_ff += __shfl_xor_sync(0xffffffffu, n_ff, d);
  if (lane == 0) warp_sums[warp] = n_ff;
  __syncthreads();
  int ff_before = 0, ff_total = 0;
  for (int w = 0; w < kFinishThreads / 32; w++) {
    const int v = warp_sums[w];
    ff_before += w < warp ? v : 0;
    ff_total += v;
  }
  const long long file_bytes = (long long)hdr.length + bytes + ff_total + 2;
  if (file_bytes > out_stride) {                                        // (uniform) the file does not fit its slot
    if (tid == 0) sizes[blockIdx.x] = -1;
    return;
  }
  for (int i = tid; i < hdr.length; i += kFinishThreads) out[i] = hdr.bytes[i];
  if (tid == 0) {
    out[file_bytes - 2] = 0xff;
    out[file_bytes - 1] = 0xd9;
    sizes[blockIdx.x] = (int32_t)file_bytes;
  }
  uint8_t* dst = out + hdr.length;
  int running = ff_before;
  for (int v0 = v_lo; v0 < v_hi; v0 += 32) {
    const int v = v0 + lane;
    unsigned m = 0;
    if (v < v_hi) m = ff_mask(v, &vec);
    const int c = __popc(m);
    int incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int u = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += u;
    }
    int at = 16 * v + running + incl - c;                               // where byte 16 v goes
    running += __shfl_sync(0xffffffffu, incl, 31);
    if (v < v_hi) {
      const uint32_t w[4] = {vec.x, vec.y, vec.z, vec.w};
#pragma unroll
      for (int p = 0; p < 16; p++) {
        if (16 * v + p < bytes) {
          dst[at++] = (uint8_t)(w[p >> 2] >> (8 * (p & 3)));
          if ((m >> p) & 1u) dst[at++] = 0;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host side
int make_enc_geometry(int w, int h, int channels, int hs, int vs, EncGeometry* g) {
  memset(g, 0, sizeof(*g));
  if (w <= 0 || h <= 0 || w > 65535 || h > 65535) { set_last_error("jpeg_encode: bad size %dx%d", w, h); return AIKO_B200_ERR_INVALID_ARGUMENT; }
  if (channels != 1 && channels != 3) { set_last_error("jpeg_encode: %d channels", channels); return AIKO_B200_ERR_UNSUPPORTED; }
  if (channels == 1) { hs = 1; vs = 1; }
  if (!((hs == 1 && vs == 1) || (hs == 2 && vs == 1) || (hs == 2 && vs == 2))) {
    set_last_error("jpeg_encode: luma sampling %dx%d", hs, vs);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  g->W = w; g->H = h; g->ncomp = channels == 1 ? 1 : 3; g->hs = hs; g->vs = vs;
  g->mcux = (w + 8 * hs - 1) / (8 * hs);
  g->mcuy = (h + 8 * vs - 1) / (8 * vs);
  int q = 0;
  for (int c = 0; c < g->ncomp; c++) {
    g->ph[c] = c == 0 ? hs : 1; g->pv[c] = c == 0 ? vs : 1;
    g->cw[c] = (w * g->ph[c] + hs - 1) / hs; g->ch[c] = (h * g->pv[c] + vs - 1) / vs;
    g->bw[c] = (g->cw[c] + 7) / 8; g->bh[c] = (g->ch[c] + 7) / 8;
    g->first_q[c] = q;
    for (int by = 0; by < g->pv[c]; by++)
      for (int bx = 0; bx < g->ph[c]; bx++, q++) { g->comp_of[q] = c; g->bx_of[q] = bx; g->by_of[q] = by; }
  }
  g->bpm = q;
  const long long blocks = (long long)g->mcux * g->mcuy * g->bpm;
  if (blocks > (1ll << 26)) { set_last_error("jpeg_encode: image too large"); return AIKO_B200_ERR_UNSUPPORTED; }
  g->n_blocks = (int)blocks;
  long long plane = 0;
  for (int c = 0; c < g->ncomp; c++) {
    g->plane_offset[c] = plane;
    plane += (long long)g->bw[c] * 8 * g->bh[c] * 8;
  }
  g->plane_per_image = (plane + 15) & ~15ll;
  return AIKO_B200_OK;
}

void scaled_quant(const unsigned char* base, int quality, uint8_t* out) {             // jcparam.c jpeg_set_quality
  quality = quality < 1 ? 1 : (quality > 100 ? 100 : quality);
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; i++) {
    long v = ((long)base[i] * scale + 50) / 100;
    out[i] = (uint8_t)(v < 1 ? 1 : (v > 255 ? 255 : v));
  }
}

void derive_codes(const unsigned char* bits, const unsigned char* vals, uint16_t* code, uint8_t* size) {   // jchuff.c
  unsigned c = 0;
  int k = 0;
  for (int len = 1; len <= 16; len++) {
    for (int i = 0; i < bits[len - 1]; i++, k++, c++) { code[vals[k]] = (uint16_t)c; size[vals[k]] = (uint8_t)len; }
    c <<= 1;
  }
}

void put_dht(uint8_t*& p, int id, const unsigned char* bits, const unsigned char* vals) {
  int total = 0;
  for (int i = 0; i < 16; i++) total += bits[i];
  *p++ = 0xff; *p++ = 0xc4; *p++ = (uint8_t)((19 + total) >> 8); *p++ = (uint8_t)((19 + total) & 255);
  *p++ = (uint8_t)id;
  memcpy(p, bits, 16); p += 16;
  memcpy(p, vals, (size_t)total); p += total;
}

// jcmarker.c: SOI, APP0 (JFIF 1.01, density 1:1), DQT per table, SOF0, DHT per table, SOS
void build_header(const EncGeometry& g, const uint8_t q[2][64], EncHeader* hdr) {
  static const uint8_t zz[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,  12, 19, 26, 33, 40, 48,
                                 41, 34, 27, 20, 13, 6,  7,  14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23,
                                 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
  uint8_t* p = hdr->bytes;
  static const uint8_t head[] = {0xff, 0xd8, 0xff, 0xe0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
  memcpy(p, head, sizeof(head)); p += sizeof(head);
  for (int t = 0; t < (g.ncomp == 1 ? 1 : 2); t++) {
    *p++ = 0xff; *p++ = 0xdb; *p++ = 0; *p++ = 67; *p++ = (uint8_t)t;
    for (int k = 0; k < 64; k++) *p++ = q[t][zz[k]];
  }
  *p++ = 0xff; *p++ = 0xc0; *p++ = 0; *p++ = (uint8_t)(8 + 3 * g.ncomp); *p++ = 8;
  *p++ = (uint8_t)(g.H >> 8); *p++ = (uint8_t)(g.H & 255); *p++ = (uint8_t)(g.W >> 8); *p++ = (uint8_t)(g.W & 255);
  *p++ = (uint8_t)g.ncomp;
  for (int c = 0; c < g.ncomp; c++) {
    *p++ = (uint8_t)(c + 1);
    *p++ = (uint8_t)(c == 0 ? (g.hs << 4) | g.vs : 0x11);
    *p++ = (uint8_t)(c == 0 ? 0 : 1);
  }
  put_dht(p, 0x00, kStdDcLuminanceBits, kStdDcLuminanceVals);
  put_dht(p, 0x10, kStdAcLuminanceBits, kStdAcLuminanceVals);
  if (g.ncomp == 3) {
    put_dht(p, 0x01, kStdDcChrominanceBits, kStdDcChrominanceVals);
    put_dht(p, 0x11, kStdAcChrominanceBits, kStdAcChrominanceVals);
  }
  *p++ = 0xff; *p++ = 0xda; *p++ = 0; *p++ = (uint8_t)(6 + 2 * g.ncomp); *p++ = (uint8_t)g.ncomp;
  for (int c = 0; c < g.ncomp; c++) { *p++ = (uint8_t)(c + 1); *p++ = (uint8_t)(c == 0 ? 0x00 : 0x11); }
  *p++ = 0; *p++ = 63; *p++ = 0;
  hdr->length = (int)(p - hdr->bytes);
}

struct EncLayout { size_t planes, coef, scan, bits, total; long long scan_words; };

size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

EncLayout enc_layout(int n, const EncGeometry& g, size_t scan_capacity) {
  EncLayout L{};
  L.scan_words = (long long)((scan_capacity + 15) / 16 * 4);            // a multiple of 16 bytes per image
  size_t at = 0;
  L.planes = at; at = align256(at + (size_t)n * g.plane_per_image);
  L.coef = at; at = align256(at + (size_t)n * g.n_blocks * 64 * sizeof(int16_t));
  L.scan = at; at = align256(at + (size_t)n * (size_t)L.scan_words * 4 + 64);
  L.bits = at; at = align256(at + (size_t)n * sizeof(long long));
  L.total = at;
  return L;
}

}  // namespace

}  // namespace aiko

using namespace aiko;

extern "C" size_t aiko_b200_jpeg_gpu_encode_workspace_bytes(int n_images, int width, int height, int channels, int hs, int vs,
                                                            size_t scan_capacity) {
  EncGeometry g;
  if (n_images <= 0 || make_enc_geometry(width, height, channels, hs, vs, &g) != AIKO_B200_OK) return 0;
  return enc_layout(n_images, g, scan_capacity).total;
}

extern "C" int aiko_b200_jpeg_gpu_encode_u8(const uint8_t* src, int n_images, int height, int width, int channels,
                                            int64_t frame_stride, int quality, int hs, int vs, void* workspace,
                                            size_t workspace_bytes, size_t scan_capacity, uint8_t* out, size_t out_stride,
                                            int32_t* sizes_out, void* stream) {
  AIKO_CHECK_ARG(n_images >= 0, "jpeg_encode: n_images < 0");
  if (n_images == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && workspace && out && sizes_out, "jpeg_encode: null pointer");
  AIKO_CHECK_ARG(frame_stride >= (int64_t)height * width * channels, "jpeg_encode: frame stride smaller than a frame");
  AIKO_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "jpeg_encode: workspace must be 256-byte aligned");
  EncGeometry g;
  const int rg = make_enc_geometry(width, height, channels, hs, vs, &g);
  if (rg != AIKO_B200_OK) return rg;
  const EncLayout L = enc_layout(n_images, g, scan_capacity);
  if (workspace_bytes < L.total) {
    set_last_error("jpeg_encode: workspace of %zu bytes, need %zu", workspace_bytes, L.total);
    return AIKO_B200_ERR_WORKSPACE;
  }
  uint8_t q[2][64];
  scaled_quant(kStdLuminanceQuant, quality, q[0]);
  scaled_quant(kStdChrominanceQuant, quality, q[1]);
  EncTables t;
  memset(&t, 0, sizeof(t));
  for (int k = 0; k < 2; k++)
    for (int i = 0; i < 64; i++) {
      t.q8[k][i] = (uint16_t)(q[k][i] * 8);
      t.recip[k][i] = ((1ull << 40) + t.q8[k][i] - 1) / t.q8[k][i];
    }
  {
    uint16_t code[256];
    uint8_t size[256];
    const unsigned char* dc_bits[2] = {kStdDcLuminanceBits, kStdDcChrominanceBits};
    const unsigned char* dc_vals[2] = {kStdDcLuminanceVals, kStdDcChrominanceVals};
    const unsigned char* ac_bits[2] = {kStdAcLuminanceBits, kStdAcChrominanceBits};
    const unsigned char* ac_vals[2] = {kStdAcLuminanceVals, kStdAcChrominanceVals};
    for (int k = 0; k < 2; k++) {
      memset(code, 0, sizeof(code)); memset(size, 0, sizeof(size));
      derive_codes(dc_bits[k], dc_vals[k], code, size);
      for (int i = 0; i < 12; i++) { t.dc_code[k][i] = code[i]; t.dc_size[k][i] = size[i]; }
      derive_codes(ac_bits[k], ac_vals[k], t.ac_code[k], t.ac_size[k]);
    }
  }
  EncHeader hdr;
  memset(&hdr, 0, sizeof(hdr));
  build_header(g, q, &hdr);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  int16_t* d_coef = reinterpret_cast<int16_t*>(ws + L.coef);
  uint32_t* d_scan = reinterpret_cast<uint32_t*>(ws + L.scan);
  long long* d_bits = reinterpret_cast<long long*>(ws + L.bits);
  AIKO_CUDA(cudaMemsetAsync(d_scan, 0, (size_t)n_images * (size_t)L.scan_words * 4, st));
  {
    uint8_t* d_planes = ws + L.planes;
    const int cc = g.ncomp == 1 ? 0 : 1;
    const long long items = (long long)g.bw[cc] * 8 * g.bh[cc] * 8;
    const dim3 grid((unsigned)((items + 255) / 256), (unsigned)n_images);
    if (g.hs == 1) jpeg_planes_kernel<1, 1><<<grid, 256, 0, st>>>(src, frame_stride, g, d_planes);
    else if (g.vs == 1) jpeg_planes_kernel<2, 1><<<grid, 256, 0, st>>>(src, frame_stride, g, d_planes);
    else jpeg_planes_kernel<2, 2><<<grid, 256, 0, st>>>(src, frame_stride, g, d_planes);
    AIKO_CUDA(cudaGetLastError());
    jpeg_fdct_kernel<<<dim3((unsigned)((g.n_blocks + kFdctThreads - 1) / kFdctThreads), (unsigned)n_images), kFdctThreads, 0, st>>>(
        d_planes, g, t, d_coef);
    AIKO_CUDA(cudaGetLastError());
  }
  jpeg_entropy_kernel<<<(unsigned)n_images, kEntropyThreads, 0, st>>>(d_coef, g, t, d_scan, L.scan_words, d_bits);
  AIKO_CUDA(cudaGetLastError());
  jpeg_finish_kernel<<<(unsigned)n_images, kFinishThreads, 0, st>>>(reinterpret_cast<const uint8_t*>(d_scan), L.scan_words * 4,
                                                                    d_bits, hdr, out, (long long)out_stride, sizes_out);
  return cuda_status(cudaGetLastError(), "jpeg_finish launch");
}
