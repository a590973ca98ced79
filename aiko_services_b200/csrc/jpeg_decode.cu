// Baseline JPEG decoding on the GPU, hand-written: frames are BORN device-resident
// and are bit-identical to what the reference's decoder produces.  The reference decodes inside
// PIL.Image.open(path) ... image.resize(...) (R/elements/media/image_io.py:184,218), i.e. with
// libjpeg-turbo's defaults: Huffman entropy decoding, the accurate integer inverse DCT ("islow"), "fancy"
// triangle-filter chroma upsampling, 16-bit fixed-point YCbCr -> RGB.  This file restates that pipeline
// (oracle: oracle/jpeg_oracle.c, itself pinned against Pillow):
//
//   host    : parse the markers of every image (quantisation and Huffman tables, frame geometry, restart
//             interval, where the entropy-coded scan starts); build 9-bit look-up tables for the Huffman
//             codes; ship tables + compressed bytes (a few hundred KB per 1080p frame instead of 6.2 MB
//             of pixels) with one copy.
//   kernel 1: jpeg_unstuff_kernel  -- the scan without stuffed zeros and markers ("clean" stream) + where every
//             restart segment starts in it (one CTA per image).
//   kernel 2: entropy decoding into quantised coefficients in HBM.  jpeg_sync_huffman_kernel: one CTA per image,
//             ONE THREAD PER SUBSEQUENCE of the stream; the decoder state at the subsequence boundaries is found
//             by iteration (Huffman streams re-synchronise), so files WITHOUT restart markers decode in
//             parallel too.  jpeg_huffman_kernel (images with more than 512 restart segments): one thread per
//             segment -- a restart marker resets the DC predictors and byte-aligns the stream.
//   kernel 3: jpeg_idct_kernel     -- dequantise + islow IDCT, one thread per 8x8 block, into planes.
//   kernel 4: jpeg_colour_kernel   -- fancy upsampling of the chroma planes + YCbCr -> RGB, interleaved
//             uint8 RGB frames: the layout np.asarray(PIL image) has.
//
// Subset: SOF0, 8 bit, one interleaved scan, gray or YCbCr with luma sampling 1x1 / 2x1 / 2x2 and chroma
// 1x1, with or without restart markers, all images of a call of one geometry.
// Everything else returns AIKO_B200_ERR_UNSUPPORTED and the caller keeps its other decoder.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace aiko {

constexpr int kLutBits = 9;
constexpr int kMaxSegments = 65535;
constexpr int kSyncMaxSegments = 512;   // images with more restart segments than this take jpeg_huffman_kernel

struct HuffDev {                 // one Huffman table, ready for the device
  uint16_t lut[1 << kLutBits];   // next 9 bits -> (length << 8) | symbol, 0 = longer than 9 bits
  int maxcode[18];               // per length; -1 = none
  int mincode[17];
  int valptr[17];
  uint8_t vals[256];
};

struct JpegTables {              // per distinct set of tables (most images of a stream share one)
  uint16_t q[4][64];             // natural order
  HuffDev dc[4], ac[4];
};

struct JpegImage {               // per image
  uint32_t scan_offset;          // of the entropy-coded data, inside the compressed area
  uint32_t scan_bytes;
  int32_t restart_interval;      // MCUs per segment
  int32_t n_segments;
  int32_t table_set;
  uint8_t td[3], ta[3], tq[3];
  uint8_t pad[3];
};

struct JpegGeometry {            // the same for every image of a call
  int W, H, ncomp, hmax, vmax;
  int mcux, mcuy;
  int h[3], v[3];                // sampling factors (1 for chroma, 1 or 2 for luma)
  int bw[3], bh[3];              // blocks per row / column, padded to whole MCUs
  int dw[3], dh[3];              // true downsampled size
  long long coef_offset[3];      // int16 elements from the start of an image's coefficient area
  long long coef_per_image;
  long long plane_offset[3];     // bytes from the start of an image's plane area
  long long plane_per_image;
};

struct ParsedImage {
  int W = 0, H = 0, ncomp = 0, restart = 0;
  int id[3], h[3], v[3], tq[3], td[3], ta[3];
  const uint8_t* scan = nullptr;
  size_t scan_bytes = 0;
  JpegTables tables;
};

static const uint8_t kZigzagHost[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                        12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                        35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                        58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};
__constant__ uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

static int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

static bool build_huff(HuffDev* h, const uint8_t* counts /* [1..16] at counts[0..15] */, const uint8_t* vals, int total) {
  memset(h, 0, sizeof(*h));
  memcpy(h->vals, vals, (size_t)total);
  int code = 0, k = 0;
  for (int len = 1; len <= 16; len++) {
    const int n = counts[len - 1];
    h->valptr[len] = k;
    h->mincode[len] = code;
    for (int i = 0; i < n; i++, code++, k++) {
      if (len <= kLutBits) {
        const int first = code << (kLutBits - len), span = 1 << (kLutBits - len);
        if (first + span > (1 << kLutBits)) return false;
        for (int j = 0; j < span; j++) h->lut[first + j] = (uint16_t)((len << 8) | vals[k]);
      }
    }
    h->maxcode[len] = n ? code - 1 : -1;
    if (code > (1 << len)) return false;
    code <<= 1;
  }
  h->maxcode[17] = 0x7fffffff;
  return true;
}

// -> 0, AIKO_B200_ERR_INVALID_ARGUMENT (malformed) or AIKO_B200_ERR_UNSUPPORTED (outside the subset)
static int parse_jpeg(const uint8_t* data, size_t len, ParsedImage* out, bool header_only) {
  if (len < 4 || data[0] != 0xff || data[1] != 0xd8) {
    set_last_error("jpeg: not a JPEG stream");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  memset(&out->tables, 0, sizeof(out->tables));
  bool have_sof = false, saw_jfif = false;
  size_t pos = 2;
  while (pos + 4 <= len) {
    if (data[pos] != 0xff) { set_last_error("jpeg: marker expected at byte %zu", pos); return AIKO_B200_ERR_INVALID_ARGUMENT; }
    const int m = data[pos + 1];
    if (m == 0xff) { pos++; continue; }
    const int seg = be16(data + pos + 2);
    const uint8_t* s = data + pos + 4;
    if (seg < 2 || pos + 2 + (size_t)seg > len) { set_last_error("jpeg: truncated segment"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
    const uint8_t* e = data + pos + 2 + seg;
    if (m == 0xdb) {
      while (s < e) {
        const int pq = s[0] >> 4, tq = s[0] & 15;
        if (tq > 3 || s + 1 + (pq ? 128 : 64) > e) { set_last_error("jpeg: bad DQT"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
        s++;
        for (int i = 0; i < 64; i++) {
          out->tables.q[tq][kZigzagHost[i]] = (uint16_t)(pq ? be16(s) : s[0]);
          s += pq ? 2 : 1;
        }
      }
    } else if (m == 0xc4) {
      while (s < e) {
        if (s + 17 > e) { set_last_error("jpeg: bad DHT"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
        const int tc = s[0] >> 4, th = s[0] & 15;
        int total = 0;
        for (int k = 1; k <= 16; k++) total += s[k];
        if (th > 3 || tc > 1 || total > 256 || s + 17 + total > e) { set_last_error("jpeg: bad DHT"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
        if (!build_huff(tc ? &out->tables.ac[th] : &out->tables.dc[th], s + 1, s + 17, total)) {
          set_last_error("jpeg: inconsistent Huffman table");
          return AIKO_B200_ERR_INVALID_ARGUMENT;
        }
        s += 17 + total;
      }
    } else if (m == 0xc0) {
      if (seg < 8 || seg < 8 + 3 * s[5]) { set_last_error("jpeg: bad SOF0"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
      if (s[0] != 8) { set_last_error("jpeg: %d-bit samples", s[0]); return AIKO_B200_ERR_UNSUPPORTED; }
      out->H = be16(s + 1); out->W = be16(s + 3); out->ncomp = s[5];
      if ((out->ncomp != 1 && out->ncomp != 3) || out->W <= 0 || out->H <= 0) {
        set_last_error("jpeg: %d components", out->ncomp);
        return AIKO_B200_ERR_UNSUPPORTED;
      }
      for (int i = 0; i < out->ncomp; i++) {
        out->id[i] = s[6 + 3 * i]; out->h[i] = s[7 + 3 * i] >> 4; out->v[i] = s[7 + 3 * i] & 15;
        out->tq[i] = s[8 + 3 * i] & 3;
      }
      have_sof = true;
      if (header_only) return AIKO_B200_OK;
    } else if (m >= 0xc1 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) {
      set_last_error("jpeg: not baseline (SOF%d)", m - 0xc0);
      return AIKO_B200_ERR_UNSUPPORTED;
    } else if (m == 0xee) {
      // APP14 "Adobe": its transform flag can declare the three components RGB (or CMYK / YCCK) instead of YCbCr
      if (seg >= 14 && memcmp(s, "Adobe", 5) == 0 && s[11] != 1) {
        set_last_error("jpeg: Adobe marker with colour transform %d", s[11]);
        return AIKO_B200_ERR_UNSUPPORTED;
      }
    } else if (m == 0xe0) {
      if (seg >= 7 && memcmp(s, "JFIF", 5) == 0) saw_jfif = true;
    } else if (m == 0xdd) {
      if (seg < 4) { set_last_error("jpeg: bad DRI"); return AIKO_B200_ERR_INVALID_ARGUMENT; }
      out->restart = be16(s);
    } else if (m == 0xda) {
      if (!have_sof || seg < 6 + 2 * out->ncomp || s[0] != out->ncomp) { set_last_error("jpeg: scan does not cover all components"); return AIKO_B200_ERR_UNSUPPORTED; }
      for (int i = 0; i < out->ncomp; i++) {
        if (s[1 + 2 * i] != out->id[i]) { set_last_error("jpeg: component order"); return AIKO_B200_ERR_UNSUPPORTED; }
        out->td[i] = (s[2 + 2 * i] >> 4) & 3; out->ta[i] = s[2 + 2 * i] & 3;
      }
      {                                       // the entropy kernels keep two DC and two AC tables in shared memory
        int dcs = 0, acs = 0;
        for (int i = 0; i < out->ncomp; i++) { dcs |= 1 << out->td[i]; acs |= 1 << out->ta[i]; }
        if (__builtin_popcount(dcs) > 2 || __builtin_popcount(acs) > 2) {
          set_last_error("jpeg: more than two Huffman tables of a kind in one scan");
          return AIKO_B200_ERR_UNSUPPORTED;
        }
      }
      // without a JFIF marker libjpeg takes component ids 'R', 'G', 'B' for RGB data (jdapimin.c default_decompress_parms)
      if (out->ncomp == 3 && !saw_jfif && out->id[0] == 'R' && out->id[1] == 'G' && out->id[2] == 'B') {
        set_last_error("jpeg: RGB components (no colour transform)");
        return AIKO_B200_ERR_UNSUPPORTED;
      }
      out->scan = e;
      out->scan_bytes = len - (size_t)(e - data);
      return AIKO_B200_OK;
    }
    pos += 2 + (size_t)seg;
  }
  set_last_error("jpeg: no scan found");
  return AIKO_B200_ERR_INVALID_ARGUMENT;
}

static int make_geometry(const ParsedImage& im, JpegGeometry* g) {
  memset(g, 0, sizeof(*g));
  g->W = im.W; g->H = im.H; g->ncomp = im.ncomp;
  for (int i = 0; i < 3; i++) { g->h[i] = g->v[i] = 1; }
  if (im.ncomp == 3) {
    if (im.h[1] != 1 || im.v[1] != 1 || im.h[2] != 1 || im.v[2] != 1) { set_last_error("jpeg: subsampled chroma layout"); return AIKO_B200_ERR_UNSUPPORTED; }
    g->h[0] = im.h[0]; g->v[0] = im.v[0];
  }
  g->hmax = g->h[0]; g->vmax = g->v[0];
  if (!((g->hmax == 1 && g->vmax == 1) || (g->hmax == 2 && g->vmax == 1) || (g->hmax == 2 && g->vmax == 2))) {
    set_last_error("jpeg: luma sampling %dx%d", g->hmax, g->vmax);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  g->mcux = (g->W + 8 * g->hmax - 1) / (8 * g->hmax);
  g->mcuy = (g->H + 8 * g->vmax - 1) / (8 * g->vmax);
  long long coef = 0, plane = 0;
  for (int i = 0; i < g->ncomp; i++) {
    g->bw[i] = g->mcux * g->h[i]; g->bh[i] = g->mcuy * g->v[i];
    g->dw[i] = (g->W * g->h[i] + g->hmax - 1) / g->hmax;
    g->dh[i] = (g->H * g->v[i] + g->vmax - 1) / g->vmax;
    g->coef_offset[i] = coef;
    coef += (long long)g->bw[i] * g->bh[i] * 64;
    g->plane_offset[i] = plane;
    plane += (long long)g->bw[i] * g->bh[i] * 64;
  }
  g->coef_per_image = coef;
  g->plane_per_image = (plane + 15) & ~15ll;
  return AIKO_B200_OK;
}

// ---------------------------------------------------------------------------------------------
// kernel 1: the scan as a clean bit stream.  Inside entropy-coded data a byte FF is always followed by a stuffed
// 00, so FF + non-zero can only be a marker (RSTn between restart segments, EOI at the end).  One CTA per image
// copies the scan without stuffed zeros and markers ("clean" bytes: segments end on byte boundaries, so they
// simply follow each other) and records where every restart segment begins in it.  A warp owns a contiguous run
// of 16-byte vectors (lanes interleaved, so loads coalesce); pass 1 counts, pass 2 writes.
constexpr int kSegThreads = 1024;

struct ByteClasses {
  unsigned data;         // bit p: byte 16 v + p is entropy-coded data
  unsigned rst_end;      // bit p: a restart marker ends with byte 16 v + p
};

__device__ __forceinline__ ByteClasses classify16(const uint8_t* __restrict__ data, int v, int bytes, uint4* vec) {
  const uint4 q = __ldg(reinterpret_cast<const uint4*>(data) + v);
  *vec = q;
  const uint32_t w[4] = {q.x, q.y, q.z, q.w};
  const int at = 16 * v;
  const unsigned valid = at + 16 <= bytes ? 0xffffu : (1u << (bytes - at)) - 1u;
  unsigned any_ff = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) any_ff |= (~w[k] - 0x01010101u) & w[k] & 0x80808080u;   // exact for "no FF byte at all"
  const unsigned before = at > 0 ? (unsigned)__ldg(data + at - 1) : 0u;
  ByteClasses c;
  if (any_ff == 0 && before != 0xffu) {
    c.data = valid;
    c.rst_end = 0;
    return c;
  }
  const unsigned after = at + 16 < bytes ? (unsigned)__ldg(data + at + 16) : 0xd9u;
  unsigned d = 0, r = 0, prev = before;
#pragma unroll
  for (int p = 0; p < 16; p++) {
    const unsigned b = (w[p >> 2] >> (8 * (p & 3))) & 0xffu;
    const unsigned n = p < 15 ? (w[(p + 1) >> 2] >> (8 * ((p + 1) & 3))) & 0xffu : after;
    const bool stuffed = b == 0u && prev == 0xffu;
    const bool marker = (b == 0xffu && n != 0u) || (prev == 0xffu && b != 0u);
    if (!stuffed && !marker) d |= 1u << p;
    if (prev == 0xffu && (b & 0xf8u) == 0xd0u) r |= 1u << p;
    prev = b;
  }
  c.data = d & valid;
  c.rst_end = r & valid;
  return c;
}

__global__ void __launch_bounds__(kSegThreads) jpeg_unstuff_kernel(const uint8_t* __restrict__ compressed,
                                                                   const JpegImage* __restrict__ images,
                                                                   uint8_t* __restrict__ clean_all,
                                                                   uint32_t* __restrict__ seg_offsets, int seg_pitch,
                                                                   int* __restrict__ status) {
  __shared__ int warp_data[kSegThreads / 32], warp_rst[kSegThreads / 32];
  const JpegImage im = images[blockIdx.x];
  const uint8_t* data = compressed + im.scan_offset;      // 16-byte aligned, padded to a multiple of 16
  uint8_t* clean = clean_all + im.scan_offset;
  const int bytes = (int)im.scan_bytes;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int vecs = (bytes + 15) >> 4;
  const int per_warp = (vecs + kSegThreads / 32 - 1) / (kSegThreads / 32);
  const int v_lo = min(warp * per_warp, vecs), v_hi = min(v_lo + per_warp, vecs);
  int n_data = 0, n_rst = 0;
  uint4 vec;
  for (int v = v_lo + lane; v < v_hi; v += 32) {
    const ByteClasses c = classify16(data, v, bytes, &vec);
    n_data += __popc(c.data);
    n_rst += __popc(c.rst_end);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    n_data += __shfl_xor_sync(0xffffffffu, n_data, d);
    n_rst += __shfl_xor_sync(0xffffffffu, n_rst, d);
  }
  if (lane == 0) { warp_data[warp] = n_data; warp_rst[warp] = n_rst; }
  __syncthreads();
  int data_before = 0, data_total = 0, rst_before = 0, rst_total = 0;
  for (int w = 0; w < kSegThreads / 32; w++) {
    const int a = warp_data[w], b = warp_rst[w];
    data_before += w < warp ? a : 0;
    data_total += a;
    rst_before += w < warp ? b : 0;
    rst_total += b;
  }
  uint32_t* out = seg_offsets + (size_t)blockIdx.x * seg_pitch;
  if (threadIdx.x == 0) {
    out[0] = 0;
    out[im.n_segments] = (uint32_t)data_total;            // end of the last segment
    if (rst_total != im.n_segments - 1) atomicExch(status, 1 + (int)blockIdx.x);   // markers do not match the restart interval
    for (int k = rst_total + 1; k < im.n_segments; k++) out[k] = (uint32_t)data_total;   // (corrupt) empty segments
  }
  int run_data = data_before, run_rst = rst_before;
  for (int v0 = v_lo; v0 < v_hi; v0 += 32) {
    const int v = v0 + lane;
    ByteClasses c;
    c.data = 0; c.rst_end = 0;
    if (v < v_hi) c = classify16(data, v, bytes, &vec);
    const int cd = __popc(c.data);
    int incl = cd;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    int at = run_data + incl - cd;
    run_data += __shfl_sync(0xffffffffu, incl, 31);
    const uint32_t w[4] = {vec.x, vec.y, vec.z, vec.w};
    {
      int o = at;                                          // (predicated, no divergence: most vectors are all data)
#pragma unroll
      for (int p = 0; p < 16; p++)
        if ((c.data >> p) & 1u) clean[o++] = (uint8_t)(w[p >> 2] >> (8 * (p & 3)));
    }
    if (__ballot_sync(0xffffffffu, c.rst_end != 0) != 0) {
      const int cr = __popc(c.rst_end);
      int incl_r = cr;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl_r, d);
        if (lane >= d) incl_r += t;
      }
      int k = run_rst + incl_r - cr;
      run_rst += __shfl_sync(0xffffffffu, incl_r, 31);
      unsigned m = c.rst_end;
      while (m) {
        const int p = __ffs(m) - 1;
        m &= m - 1;
        k++;
        if (k < im.n_segments) out[k] = (uint32_t)(at + __popc(c.data & ((1u << p) - 1u)));
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Reading the clean stream: 64-bit accumulator, refilled with aligned big-endian 32-bit words.
struct CleanReader {
  const uint32_t* words;         // the image's clean stream (16-byte aligned)
  int wi;                        // next word to load
  unsigned long long acc;        // head bit = bit 63
  int n;                         // valid bits in acc (after a refill: 33..64)
};

__device__ __forceinline__ uint32_t be32(uint32_t x) { return __byte_perm(x, 0, 0x0123); }

__device__ __forceinline__ void reader_init(CleanReader& r, const uint32_t* words, uint32_t bit) {
  r.words = words;
  const int w = (int)(bit >> 5), drop = (int)(bit & 31u);
  r.acc = (((unsigned long long)be32(__ldg(words + w)) << 32) | be32(__ldg(words + w + 1))) << drop;
  r.n = 64 - drop;
  r.wi = w + 2;
}

__device__ __forceinline__ void reader_refill(CleanReader& r) {
  if (r.n <= 32) {
    r.acc |= (unsigned long long)be32(__ldg(r.words + r.wi)) << (32 - r.n);
    r.wi++;
    r.n += 32;
  }
}

__device__ __forceinline__ uint32_t reader_position(const CleanReader& r) { return 32u * (uint32_t)r.wi - (uint32_t)r.n; }

// The Huffman code at the head of the stream (at least 16 valid bits): -> symbol, *len = its length; -1 = no such code.
__device__ __forceinline__ int peek_symbol(const CleanReader& r, const HuffDev& h, int* len) {
  const uint32_t hi = (uint32_t)(r.acc >> 32);
  const unsigned e = h.lut[hi >> (32 - kLutBits)];
  if (e) {
    *len = (int)(e >> 8);
    return (int)(e & 0xffu);
  }
  for (int l = kLutBits + 1; l <= 16; l++) {
    const int code = (int)(hi >> (32 - l));
    if (code <= h.maxcode[l]) {
      *len = l;
      return h.vals[(h.valptr[l] + code - h.mincode[l]) & 255];
    }
  }
  *len = 1;
  return -1;
}

// Consumes a code of `len` bits and the `size` magnitude bits that follow it (len + size <= 31, all within the
// top 32 bits of the accumulator).  -> the magnitude, sign-extended as jdhuff.c's HUFF_EXTEND does (0 for size 0).
__device__ __forceinline__ int take_code_and_value(CleanReader& r, int len, int size) {
  const uint32_t hi = (uint32_t)(r.acc >> 32);
  int v = 0;
  if (size) {
    v = (int)((hi << len) >> (32 - size));
    v = v < (1 << (size - 1)) ? v - (1 << size) + 1 : v;
  }
  r.acc <<= (len + size);
  r.n -= len + size;
  return v;
}

// ---------------------------------------------------------------------------------------------
// kernel 2: Huffman decoding (jdhuff.c) of images with many restart segments, one thread per segment.
constexpr int kHuffThreads = 64;

__global__ void __launch_bounds__(kHuffThreads) jpeg_huffman_kernel(const uint8_t* __restrict__ clean_all,
                                                                    const JpegImage* __restrict__ images,
                                                                    const JpegTables* __restrict__ tables,
                                                                    const uint32_t* __restrict__ seg_offsets,
                                                                    int seg_pitch, const JpegGeometry g,
                                                                    int16_t* __restrict__ coef, int* __restrict__ status) {
  __shared__ HuffDev s_dc[2], s_ac[2];                    // the tables the image's components use (luma, chroma)
  __shared__ int s_slot_dc[4], s_slot_ac[4];
  __shared__ uint8_t s_zigzag[64];                        // (divergent indices: shared memory, not the constant cache)
  if (threadIdx.x < 64) s_zigzag[threadIdx.x] = kZigzag[threadIdx.x];
  const JpegImage im = images[blockIdx.y];
  if (im.n_segments <= kSyncMaxSegments) return;          // (uniform) jpeg_sync_huffman_kernel's image
  const JpegTables& t = tables[im.table_set];
  // components use at most two DC and two AC tables in this subset (luma's and chroma's)
  if (threadIdx.x < 4) { s_slot_dc[threadIdx.x] = -1; s_slot_ac[threadIdx.x] = -1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    int nd = 0, na = 0;
    for (int c = 0; c < g.ncomp; c++) {
      if (s_slot_dc[im.td[c]] < 0 && nd < 2) s_slot_dc[im.td[c]] = nd++;
      if (s_slot_ac[im.ta[c]] < 0 && na < 2) s_slot_ac[im.ta[c]] = na++;
    }
  }
  __syncthreads();
  for (int id = 0; id < 4; id++) {
    if (s_slot_dc[id] >= 0) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&t.dc[id]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&s_dc[s_slot_dc[id]]);
      for (int i = threadIdx.x; i < (int)(sizeof(HuffDev) / 4); i += kHuffThreads) dst[i] = __ldg(src + i);
    }
    if (s_slot_ac[id] >= 0) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&t.ac[id]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&s_ac[s_slot_ac[id]]);
      for (int i = threadIdx.x; i < (int)(sizeof(HuffDev) / 4); i += kHuffThreads) dst[i] = __ldg(src + i);
    }
  }
  __syncthreads();
  const int seg = blockIdx.x * kHuffThreads + threadIdx.x;
  if (seg >= im.n_segments) return;
  bool table_ok = true;
  for (int c = 0; c < g.ncomp; c++) table_ok &= s_slot_dc[im.td[c]] >= 0 && s_slot_ac[im.ta[c]] >= 0;
  if (!table_ok) {                                        // three different tables: not produced by any common encoder
    atomicExch(status, 1000 + (int)blockIdx.y);
    return;
  }
  const uint32_t* offs = seg_offsets + (size_t)blockIdx.y * seg_pitch;
  CleanReader r;
  reader_init(r, reinterpret_cast<const uint32_t*>(clean_all + im.scan_offset), 8u * offs[seg]);
  const uint32_t end_bit = 8u * offs[seg + 1];
  int16_t* image_coef = coef + (size_t)blockIdx.y * g.coef_per_image;
  const int total_mcus = g.mcux * g.mcuy;
  const int first = seg * im.restart_interval, last = min(first + im.restart_interval, total_mcus);
  int pred[3] = {0, 0, 0};
  int bad = 0;
  for (int m = first; m < last && !bad; m++) {
    const int my = m / g.mcux, mx = m - my * g.mcux;
#pragma unroll 1
    for (int c = 0; c < g.ncomp; c++) {
      const HuffDev& hd = s_dc[s_slot_dc[im.td[c]]];
      const HuffDev& ha = s_ac[s_slot_ac[im.ta[c]]];
      const int nblocks = g.h[c] * g.v[c];
#pragma unroll 1
      for (int q = 0; q < nblocks; q++) {
        const int by = q / g.h[c], bx = q - by * g.h[c];
        int16_t* blk = image_coef + g.coef_offset[c] + ((size_t)(my * g.v[c] + by) * g.bw[c] + (mx * g.h[c] + bx)) * 64;
        reader_refill(r);
        if (reader_position(r) > end_bit) { bad = 4000; break; }          // (corrupt) ran past the segment's data
        int len;
        const int s = peek_symbol(r, hd, &len);
        if (s < 0 || s > 15) { bad = 2000; break; }
        pred[c] += take_code_and_value(r, len, s);
        if (pred[c]) blk[0] = (int16_t)pred[c];
        for (int k = 1; k < 64; k++) {
          reader_refill(r);
          const int rs = peek_symbol(r, ha, &len);
          if (rs < 0) { bad = 2000; break; }
          const int run = rs >> 4, size = rs & 15;
          const int v = take_code_and_value(r, len, size);
          if (size) {
            k += run;
            if (k > 63) { bad = 3000; break; }
            blk[s_zigzag[k]] = (int16_t)v;
          } else if (run == 15) {
            k += 15;
          } else {
            break;
          }
        }
        if (bad) break;
      }
      if (bad) break;
    }
  }
  if (!bad && reader_position(r) > end_bit) bad = 4000;                // read past the segment's data
  if (bad) atomicExch(status, bad + (int)blockIdx.y);
}

// ---------------------------------------------------------------------------------------------
// kernel 2b: Huffman decoding of scans with few (or no) restart markers: one CTA per image, one thread per
// SUBSEQUENCE of the clean stream (a run of B bytes inside one restart segment; B is chosen so that the image has
// at most 1024 of them).  A thread that does not start a segment does not know where a codeword begins, which
// block coefficient comes next or which component it belongs to, so the decoder state (bit position u, zigzag
// index z, block-in-MCU c) at every subsequence boundary is found by iteration (self-synchronising Huffman
// decoding, Klein & Wiseman 2003; on GPUs: Weissenberger & Schmidt 2018 / 2021): every thread first decodes its
// subsequence from a guessed state, then again from the exit state of its predecessor, until no exit state
// changes -- Huffman streams re-synchronise within a few dozen symbols, so two rounds settle almost every
// boundary, and the loop is exact in any case (the first thread of a segment starts from the true state, and
// each round extends the settled prefix).  Then the blocks completed per subsequence are prefix-summed into
// block indices, a last pass writes the coefficients (DC as differences), and the DC predictions are resolved
// with a prefix sum over the per-thread sums of differences.
constexpr int kSyncThreads = 1024;

struct SyncShared {
  HuffDev dc[2], ac[2];
  int slot_dc[4], slot_ac[4];
  uint8_t zigzag[64];
  uint8_t comp_of[8], bx_of[8], by_of[8], dc_of[8], ac_of[8];   // per block of an MCU
  uint32_t seg_first[kSyncMaxSegments + 1];   // first subsequence of each segment
  uint32_t count[kSyncThreads];               // scans: completed blocks, sums of DC differences
  unsigned long long state[kSyncThreads];     // exit state (u << 16 | z << 4 | c), u = bit position in the clean stream
  uint8_t changed[2][kSyncThreads];
  int warp_sums[kSyncThreads / 32];
};

__device__ __forceinline__ int block_scan_excl(int v, int* warp_sums, int* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) warp_sums[warp] = incl;
  __syncthreads();
  int before = 0, all = 0;
  for (int w = 0; w < (int)blockDim.x / 32; w++) {
    const int s = warp_sums[w];
    before += w < warp ? s : 0;
    all += s;
  }
  __syncthreads();
  *total = all;
  return before + incl - v;
}

// Decodes from the state `entry` until the next codeword starts at or after bit `limit` (WRITE: or block
// `g_limit` is reached).  -> exit state; *blocks = blocks completed.  WRITE stores the coefficients of blocks
// gi, gi + 1, ... (DC as the difference to the previous block of the component).
template <bool WRITE>
__device__ __forceinline__ unsigned long long decode_span(const SyncShared& S, const JpegGeometry& g,
                                                         const uint32_t* words_in, uint32_t limit,
                                                         unsigned long long entry, int* blocks, int16_t* image_coef,
                                                         int bpm, int gi, int g_limit, int dsum[3], int* dc_first,
                                                         int* dc_count, int* bad) {
  uint32_t u = (uint32_t)(entry >> 16);
  int z = (int)((entry >> 4) & 63u), c = (int)(entry & 15u);
  const uint32_t* words = words_in;
  asm volatile("" : "+l"(words));            // keep the stream's base address in registers across the loop
  CleanReader r;
  reader_init(r, words, u);
  int done = 0;
  int16_t* blk = nullptr;
  int mx = 0, my = 0;
  const HuffDev* hd = &S.dc[S.dc_of[c]];     // the tables of the block being decoded
  const HuffDev* ha = &S.ac[S.ac_of[c]];
  if (WRITE) {
    const int m = gi / bpm;
    my = m / g.mcux; mx = m - my * g.mcux;
    *dc_first = gi + (z > 0 ? 1 : 0);
  }
  auto block_ptr = [&](int q) -> int16_t* {
    const int comp = S.comp_of[q];
    return image_coef + g.coef_offset[comp] +
           ((size_t)(my * g.v[comp] + S.by_of[q]) * g.bw[comp] + (mx * g.h[comp] + S.bx_of[q])) * 64;
  };
  if (WRITE && gi < g_limit) blk = block_ptr(c);
  for (;;) {
    reader_refill(r);
    u = reader_position(r);
    if (u >= limit) break;
    if (WRITE && gi >= g_limit) break;
    int len;
    const int sym = peek_symbol(r, z == 0 ? *hd : *ha, &len);
    if (sym < 0 || (z == 0 && sym > 15)) {    // no such code / DC category: only in a wrong state or a corrupt file
      r.acc <<= 1; r.n -= 1;
      if (WRITE) *bad = 2000;
      continue;
    }
    const int run = z == 0 ? 0 : sym >> 4, size = z == 0 ? sym : sym & 15;
    const int v = take_code_and_value(r, len, size);
    if (size) {
      const int k = z + run;
      if (k > 63) {
        if (WRITE) *bad = 3000;
        z = 64;
      } else {
        if (WRITE) {
          if (z == 0) {
            const int comp = S.comp_of[c];
            dsum[0] += comp == 0 ? v : 0; dsum[1] += comp == 1 ? v : 0; dsum[2] += comp == 2 ? v : 0;
            ++*dc_count;
          }
          blk[S.zigzag[k]] = (int16_t)v;
        }
        z = k + 1;
      }
    } else if (z == 0) {
      if (WRITE) ++*dc_count;
      z = 1;
    } else {
      z = run == 15 ? z + 16 : 64;
    }
    if (z >= 64) {
      z = 0;
      done++;
      if (++c == bpm) {
        c = 0;
        if (WRITE && ++mx == g.mcux) { mx = 0; my++; }
      }
      hd = &S.dc[S.dc_of[c]];
      ha = &S.ac[S.ac_of[c]];
      if (WRITE) {
        gi++;
        if (gi < g_limit) blk = block_ptr(c);
      }
    }
  }
  *blocks = done;
  return ((unsigned long long)u << 16) | (unsigned long long)((z << 4) | c);
}

__global__ void __launch_bounds__(kSyncThreads) jpeg_sync_huffman_kernel(const uint8_t* __restrict__ clean_all,
                                                                         const JpegImage* __restrict__ images,
                                                                         const JpegTables* __restrict__ tables,
                                                                         const uint32_t* __restrict__ seg_offsets,
                                                                         int seg_pitch, const JpegGeometry g,
                                                                         int16_t* __restrict__ coef,
                                                                         int* __restrict__ status) {
  extern __shared__ __align__(16) uint8_t sync_smem[];
  SyncShared& S = *reinterpret_cast<SyncShared*>(sync_smem);
  const JpegImage im = images[blockIdx.x];
  if (im.n_segments > kSyncMaxSegments) return;            // (uniform) jpeg_huffman_kernel's image
  const JpegTables& t = tables[im.table_set];
  const int tid = (int)threadIdx.x;
  if (tid < 64) S.zigzag[tid] = kZigzag[tid];
  if (tid < 4) { S.slot_dc[tid] = -1; S.slot_ac[tid] = -1; }
  __syncthreads();
  int bpm = 0;
  for (int c = 0; c < g.ncomp; c++) bpm += g.h[c] * g.v[c];
  if (tid == 0) {
    int nd = 0, na = 0, q = 0;
    for (int c = 0; c < g.ncomp; c++) {
      if (S.slot_dc[im.td[c]] < 0 && nd < 2) S.slot_dc[im.td[c]] = nd++;
      if (S.slot_ac[im.ta[c]] < 0 && na < 2) S.slot_ac[im.ta[c]] = na++;
      for (int by = 0; by < g.v[c]; by++)
        for (int bx = 0; bx < g.h[c]; bx++, q++) {
          S.comp_of[q] = (uint8_t)c; S.bx_of[q] = (uint8_t)bx; S.by_of[q] = (uint8_t)by;
          S.dc_of[q] = (uint8_t)max(S.slot_dc[im.td[c]], 0); S.ac_of[q] = (uint8_t)max(S.slot_ac[im.ta[c]], 0);
        }
    }
  }
  __syncthreads();
  bool table_ok = true;
  for (int c = 0; c < g.ncomp; c++) table_ok &= S.slot_dc[im.td[c]] >= 0 && S.slot_ac[im.ta[c]] >= 0;
  if (!table_ok) {                                          // (uniform) three different tables
    if (tid == 0) atomicExch(status, 1000 + (int)blockIdx.x);
    return;
  }
  for (int id = 0; id < 4; id++) {
    if (S.slot_dc[id] >= 0) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&t.dc[id]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&S.dc[S.slot_dc[id]]);
      for (int i = tid; i < (int)(sizeof(HuffDev) / 4); i += kSyncThreads) dst[i] = __ldg(src + i);
    }
    if (S.slot_ac[id] >= 0) {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(&t.ac[id]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(&S.ac[S.slot_ac[id]]);
      for (int i = tid; i < (int)(sizeof(HuffDev) / 4); i += kSyncThreads) dst[i] = __ldg(src + i);
    }
  }
  // ---- the subsequences: B clean bytes each, never across a segment boundary
  const uint32_t* offs = seg_offsets + (size_t)blockIdx.x * seg_pitch;
  const uint32_t* words = reinterpret_cast<const uint32_t*>(clean_all + im.scan_offset);
  const int n_seg = im.n_segments;
  const uint32_t clean_bytes = offs[n_seg];
  const uint32_t B = max(((clean_bytes + (uint32_t)(kSyncThreads - n_seg) - 1) / (uint32_t)(kSyncThreads - n_seg) + 63u) & ~63u, 64u);
  int total;
  {
    int mine = 0;
    if (tid < n_seg) mine = (int)((offs[tid + 1] - offs[tid] + B - 1) / B);
    const int first = block_scan_excl(mine, S.warp_sums, &total);
    if (tid < n_seg) S.seg_first[tid] = (uint32_t)first;
    if (tid == 0) S.seg_first[n_seg] = (uint32_t)total;
  }
  __syncthreads();
  const int n_sub = total;                                  // <= kSyncThreads by the choice of B
  const bool active = tid < n_sub;
  int seg = 0;
  uint32_t first_bit = 0, limit = 0;
  bool first_of_segment = false;
  if (active) {
    int lo = 0, hi = n_seg - 1;                             // last segment whose first subsequence is <= tid
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if ((int)S.seg_first[mid] <= tid) lo = mid; else hi = mid - 1;
    }
    seg = lo;
    const uint32_t seg_lo = offs[seg], seg_hi = offs[seg + 1];
    const uint32_t i = (uint32_t)tid - S.seg_first[seg];
    first_of_segment = i == 0;
    const uint32_t sub_lo = seg_lo + i * B;
    first_bit = 8u * sub_lo;
    limit = 8u * min(sub_lo + B, seg_hi);
  }
  // ---- the exit state of every subsequence, by iteration
  int16_t* image_coef = coef + (size_t)blockIdx.x * g.coef_per_image;
  int dsum[3] = {0, 0, 0};
  int dc_first = 0, dc_count = 0;
  int bad = 0;
  int blocks = 0;
  int cur = 0;
  for (int iter = 0;; iter++) {
    bool need = false;
    unsigned long long entry = 0;
    if (active) {
      if (iter == 0) {
        need = true;
        entry = (unsigned long long)first_bit << 16;       // guess: a block begins here (exact for a segment's first)
      } else if (!first_of_segment && S.changed[cur ^ 1][tid - 1]) {
        need = true;
        entry = S.state[tid - 1];
      }
    }
    unsigned long long exit_state = 0;
    if (need)
      exit_state = decode_span<false>(S, g, words, limit, entry, &blocks, nullptr, bpm, 0, 0, dsum, &dc_first, &dc_count, &bad);
    __syncthreads();
    bool ch = false;
    if (need) {
      ch = iter == 0 || exit_state != S.state[tid];
      S.state[tid] = exit_state;
    }
    if (active) S.changed[cur][tid] = ch ? 1 : 0;
    const int any = __syncthreads_or(ch ? 1 : 0);
    cur ^= 1;
    if (!any) {
      if (tid == 0) { atomicMax(status + 1, iter); atomicAdd(status + 2, iter); }    // diagnostics: rounds needed
      break;
    }
  }
  // ---- block index of every subsequence, then the pass that writes
  {
    const int before = block_scan_excl(active ? blocks : 0, S.warp_sums, &total);
    if (active) S.count[tid] = (uint32_t)before;
  }
  __syncthreads();
  const int total_mcus = g.mcux * g.mcuy;
  if (active) {
    const int interval = im.restart_interval > 0 ? im.restart_interval : total_mcus;
    const int g0 = seg * interval * bpm + (int)(S.count[tid] - S.count[S.seg_first[seg]]);
    const int g_limit = min((seg + 1) * interval, total_mcus) * bpm;
    const unsigned long long entry = first_of_segment ? (unsigned long long)first_bit << 16 : S.state[tid - 1];
    const unsigned long long exit_state = decode_span<true>(S, g, words, limit, entry, &blocks, image_coef, bpm, g0, g_limit,
                                                            dsum, &dc_first, &dc_count, &bad);
    const bool last_of_segment = tid + 1 == (int)S.seg_first[seg + 1];
    if (last_of_segment && g0 + blocks != g_limit) bad = 4000;        // the segment ended early (or ran over)
    // (cannot happen: the iteration settled; a thread that stopped at the segment's last block skipped the pad bits)
    if (!last_of_segment && g0 + blocks < g_limit && exit_state != S.state[tid]) bad = 5000;
  }
  if (bad) atomicExch(status, bad + (int)blockIdx.x);
  // ---- DC predictions: prefix sums of the differences per component, restarted at every segment
  int pred[3];
#pragma unroll
  for (int c = 0; c < 3; c++) {
    __syncthreads();
    const int before = block_scan_excl(active ? dsum[c] : 0, S.warp_sums, &total);
    if (active) S.count[tid] = (uint32_t)before;
    __syncthreads();
    pred[c] = active ? (int)(S.count[tid] - S.count[S.seg_first[seg]]) : 0;
  }
  if (active && dc_count > 0) {
    const int m = dc_first / bpm;
    int q = dc_first - m * bpm;
    int my = m / g.mcux, mx = m - my * g.mcux;
    for (int i = 0; i < dc_count; i++) {
      const int comp = S.comp_of[q];
      int16_t* blk = image_coef + g.coef_offset[comp] +
                     ((size_t)(my * g.v[comp] + S.by_of[q]) * g.bw[comp] + (mx * g.h[comp] + S.bx_of[q])) * 64;
      const int v = (comp == 0 ? pred[0] : comp == 1 ? pred[1] : pred[2]) + (int)blk[0];
      if (comp == 0) pred[0] = v; else if (comp == 1) pred[1] = v; else pred[2] = v;
      blk[0] = (int16_t)v;
      if (++q == bpm) {
        q = 0;
        if (++mx == g.mcux) { mx = 0; my++; }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 3: dequantisation + the accurate integer inverse DCT (jidctint.c, CONST_BITS 13, PASS1_BITS 2),
// one thread per 8x8 block.
#define AIKO_IDCT_1D(i0, i1, i2, i3, i4, i5, i6, i7, o0, o1, o2, o3, o4, o5, o6, o7, SHIFT)          \
  {                                                                                                    \
    int z1, z2, z3, z4, z5, t0, t1, t2, t3, t10, t11, t12, t13;                                        \
    z2 = (i2); z3 = (i6);                                                                              \
    z1 = (z2 + z3) * 4433;                                                                             \
    t2 = z1 + z3 * (-15137);                                                                           \
    t3 = z1 + z2 * 6270;                                                                               \
    z2 = (i0); z3 = (i4);                                                                              \
    t0 = (z2 + z3) * 8192;                                                                             \
    t1 = (z2 - z3) * 8192;                                                                             \
    t10 = t0 + t3; t13 = t0 - t3; t11 = t1 + t2; t12 = t1 - t2;                                        \
    t0 = (i7); t1 = (i5); t2 = (i3); t3 = (i1);                                                        \
    z1 = t0 + t3; z2 = t1 + t2; z3 = t0 + t2; z4 = t1 + t3;                                            \
    z5 = (z3 + z4) * 9633;                                                                             \
    t0 *= 2446; t1 *= 16819; t2 *= 25172; t3 *= 12299;                                                 \
    z1 *= -7373; z2 *= -20995; z3 *= -16069; z4 *= -3196;                                              \
    z3 += z5; z4 += z5;                                                                                \
    t0 += z1 + z3; t1 += z2 + z4; t2 += z2 + z3; t3 += z1 + z4;                                        \
    (o0) = (t10 + t3 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o7) = (t10 - t3 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o1) = (t11 + t2 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o6) = (t11 - t2 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o2) = (t12 + t1 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o5) = (t12 - t1 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o3) = (t13 + t0 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
    (o4) = (t13 - t0 + (1 << ((SHIFT)-1))) >> (SHIFT);                                                 \
  }

__device__ __forceinline__ unsigned clamp_sample(int v) {
  v += 128;
  return (unsigned)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

constexpr int kIdctThreads = 128;

__global__ void __launch_bounds__(kIdctThreads) jpeg_idct_kernel(const int16_t* __restrict__ coef,
                                                                 const JpegImage* __restrict__ images,
                                                                 const JpegTables* __restrict__ tables,
                                                                 const JpegGeometry g, uint8_t* __restrict__ planes) {
  __shared__ uint16_t s_q[3][64];
  const JpegImage im = images[blockIdx.y];
  for (int i = threadIdx.x; i < g.ncomp * 64; i += kIdctThreads) s_q[i >> 6][i & 63] = tables[im.table_set].q[im.tq[i >> 6]][i & 63];
  __syncthreads();
  const long long blocks_per_image = g.coef_per_image / 64;
  const long long blk = (long long)blockIdx.x * kIdctThreads + threadIdx.x;
  if (blk >= blocks_per_image) return;
  int c = 0;
  if (g.ncomp == 3) c = blk * 64 >= g.coef_offset[2] ? 2 : (blk * 64 >= g.coef_offset[1] ? 1 : 0);
  const long long local = blk - g.coef_offset[c] / 64;
  const int by = (int)(local / g.bw[c]), bx = (int)(local - (long long)by * g.bw[c]);
  const uint4* src = reinterpret_cast<const uint4*>(coef + (size_t)blockIdx.y * g.coef_per_image + blk * 64);
  int x[64];
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const uint4 v = __ldg(src + r);                       // eight coefficients of row r
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int k = 0; k < 4; k++) {
      x[8 * r + 2 * k] = (int)(short)(w[k] & 0xffffu) * (int)s_q[c][8 * r + 2 * k];
      x[8 * r + 2 * k + 1] = (int)(short)(w[k] >> 16) * (int)s_q[c][8 * r + 2 * k + 1];
    }
  }
#pragma unroll
  for (int col = 0; col < 8; col++)                       // pass 1: columns
    AIKO_IDCT_1D(x[col], x[8 + col], x[16 + col], x[24 + col], x[32 + col], x[40 + col], x[48 + col], x[56 + col],
                 x[col], x[8 + col], x[16 + col], x[24 + col], x[32 + col], x[40 + col], x[48 + col], x[56 + col], 11)
  const int pitch = g.bw[c] * 8;
  uint8_t* out = planes + (size_t)blockIdx.y * g.plane_per_image + g.plane_offset[c] + (size_t)by * 8 * pitch + bx * 8;
#pragma unroll
  for (int r = 0; r < 8; r++) {                           // pass 2: rows
    int o0, o1, o2, o3, o4, o5, o6, o7;
    AIKO_IDCT_1D(x[8 * r], x[8 * r + 1], x[8 * r + 2], x[8 * r + 3], x[8 * r + 4], x[8 * r + 5], x[8 * r + 6], x[8 * r + 7],
                 o0, o1, o2, o3, o4, o5, o6, o7, 18)
    uint2 px;
    px.x = clamp_sample(o0) | (clamp_sample(o1) << 8) | (clamp_sample(o2) << 16) | (clamp_sample(o3) << 24);
    px.y = clamp_sample(o4) | (clamp_sample(o5) << 8) | (clamp_sample(o6) << 16) | (clamp_sample(o7) << 24);
    *reinterpret_cast<uint2*>(out + (size_t)r * pitch) = px;
  }
}

// ---------------------------------------------------------------------------------------------
// kernel 4: fancy upsampling (jdsample.c h2v1 / h2v2 triangle filters; rows above / below the image
// repeat the edge row) + YCbCr -> RGB (jdcolor.c, 16-bit fixed point).  One thread per 8 output pixels
// of one row; MODE: 0 gray, 1 = 4:4:4, 2 = 4:2:2 (h2v1), 3 = 4:2:0 (h2v2).
__device__ __forceinline__ unsigned clamp8(int v) { return (unsigned)(v < 0 ? 0 : (v > 255 ? 255 : v)); }

// Six neighbouring chroma samples c[0..5] = columns sx0 - 1 .. sx0 + 4 of one row (sx0 % 4 == 0; columns outside
// the padded plane read as 0 and are never used by the edge rules below).
__device__ __forceinline__ void chroma_six(const uint8_t* __restrict__ row, int sx0, int pitch, int c[6]) {
  const uint32_t w = *reinterpret_cast<const uint32_t*>(row + sx0);
  c[0] = sx0 > 0 ? (int)row[sx0 - 1] : 0;
  c[1] = (int)(w & 0xffu); c[2] = (int)((w >> 8) & 0xffu); c[3] = (int)((w >> 16) & 0xffu); c[4] = (int)(w >> 24);
  c[5] = sx0 + 4 < pitch ? (int)row[sx0 + 4] : 0;
}

template <int MODE>
__device__ __forceinline__ void chroma_eight(const uint8_t* __restrict__ plane, int pitch, int dw, int dh, int x0, int y,
                                             int out[8]) {
  if (MODE == 1) {
    const uint2 v = *reinterpret_cast<const uint2*>(plane + (size_t)y * pitch + x0);
#pragma unroll
    for (int k = 0; k < 4; k++) { out[k] = (int)((v.x >> (8 * k)) & 0xffu); out[4 + k] = (int)((v.y >> (8 * k)) & 0xffu); }
    return;
  }
  const int sx0 = x0 >> 1;
  if (dw <= 2) {
    // jdsample.c jinit_upsampler installs the fancy filters only for downsampled_width > 2: narrower components
    // are upsampled by replication (h2v1_upsample / h2v2_upsample); such an image is at most 4 pixels wide
    const uint8_t* row = plane + (size_t)(MODE == 3 ? y >> 1 : y) * pitch;
#pragma unroll
    for (int k = 0; k < 8; k++) out[k] = row[min(sx0 + (k >> 1), pitch - 1)];
    return;
  }
  int c[6];
  if (MODE == 2) {
    chroma_six(plane + (size_t)y * pitch, sx0, pitch, c);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int j = 1 + (k >> 1), v = c[j];
      if (k & 1) out[k] = (sx0 + (k >> 1) == dw - 1) ? v : (v * 3 + c[j + 1] + 2) >> 2;
      else out[k] = (x0 + k == 0) ? v : (v * 3 + c[j - 1] + 1) >> 2;
    }
    return;
  }
  const int sy = y >> 1;
  const int fy = min(max((y & 1) ? sy + 1 : sy - 1, 0), dh - 1);
  int f[6];
  chroma_six(plane + (size_t)sy * pitch, sx0, pitch, c);
  chroma_six(plane + (size_t)fy * pitch, sx0, pitch, f);
#pragma unroll
  for (int j = 0; j < 6; j++) c[j] = c[j] * 3 + f[j];     // jdsample.c: thiscolsum
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const int j = 1 + (k >> 1), v = c[j];
    if (k & 1) out[k] = (sx0 + (k >> 1) == dw - 1) ? (v * 4 + 7) >> 4 : (v * 3 + c[j + 1] + 7) >> 4;
    else out[k] = (x0 + k == 0) ? (v * 4 + 8) >> 4 : (v * 3 + c[j - 1] + 8) >> 4;
  }
}

template <int MODE>
__global__ void __launch_bounds__(256) jpeg_colour_kernel(const uint8_t* __restrict__ planes, const JpegGeometry g,
                                                          uint8_t* __restrict__ dst, long long frame_stride) {
  const int groups = (g.W + 7) >> 3;
  const int item = (int)blockIdx.x * 256 + (int)threadIdx.x;
  if (item >= groups * g.H) return;
  const int y = item / groups, x0 = (item - y * groups) * 8;
  const uint8_t* base = planes + (size_t)blockIdx.y * g.plane_per_image;
  const uint2 yw = *reinterpret_cast<const uint2*>(base + g.plane_offset[0] + (size_t)y * (g.bw[0] * 8) + x0);
  uint32_t o[6];
  if (MODE == 0) {
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const unsigned yy = ((k < 4 ? yw.x : yw.y) >> (8 * (k & 3))) & 0xffu;
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const int at = 3 * k + q;
        if ((at & 3) == 0) o[at >> 2] = yy; else o[at >> 2] |= yy << (8 * (at & 3));
      }
    }
  } else {
    int cb[8], cr[8];
    const int pitch_c = g.bw[1] * 8;
    chroma_eight<MODE>(base + g.plane_offset[1], pitch_c, g.dw[1], g.dh[1], x0, y, cb);
    chroma_eight<MODE>(base + g.plane_offset[2], pitch_c, g.dw[2], g.dh[2], x0, y, cr);
#pragma unroll
    for (int k = 0; k < 8; k++) {
      const int yy = (int)(((k < 4 ? yw.x : yw.y) >> (8 * (k & 3))) & 0xffu);
      const int u = cb[k] - 128, v = cr[k] - 128;
      const unsigned px[3] = {clamp8(yy + ((91881 * v + 32768) >> 16)),
                              clamp8(yy + ((-22554 * u + 32768 - 46802 * v) >> 16)),
                              clamp8(yy + ((116130 * u + 32768) >> 16))};
#pragma unroll
      for (int q = 0; q < 3; q++) {
        const int at = 3 * k + q;
        if ((at & 3) == 0) o[at >> 2] = px[q]; else o[at >> 2] |= px[q] << (8 * (at & 3));
      }
    }
  }
  uint8_t* out = dst + (size_t)blockIdx.y * frame_stride + ((size_t)y * g.W + x0) * 3;
  const int npx = min(8, g.W - x0);
  const uintptr_t addr = reinterpret_cast<uintptr_t>(out);
  if (npx == 8 && (addr & 7) == 0) {
    uint2* p = reinterpret_cast<uint2*>(out);
    p[0] = make_uint2(o[0], o[1]); p[1] = make_uint2(o[2], o[3]); p[2] = make_uint2(o[4], o[5]);
  } else if (npx == 8 && (addr & 3) == 0) {
    uint32_t* p = reinterpret_cast<uint32_t*>(out);
#pragma unroll
    for (int q = 0; q < 6; q++) p[q] = o[q];
  } else {
    for (int q = 0; q < npx * 3; q++) out[q] = (uint8_t)(o[q >> 2] >> (8 * (q & 3)));
  }
}

}  // namespace aiko

// ---------------------------------------------------------------------------------------------
// Host side of the C ABI.
using namespace aiko;

namespace {

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct WorkspaceLayout {
  size_t images, tables, head_bytes;       // uploaded from the staging buffer in one copy
  size_t seg_offsets, status, compressed, clean, coef, planes, total;
  int seg_pitch;
};

WorkspaceLayout layout_for(int n, const JpegGeometry& g, size_t compressed_capacity, int max_segments) {
  WorkspaceLayout L{};
  size_t at = 0;
  L.images = at; at = align_up(at + (size_t)n * sizeof(JpegImage), 256);
  L.tables = at; at = align_up(at + (size_t)n * sizeof(JpegTables), 256);
  L.head_bytes = at;
  L.seg_pitch = max_segments + 1;
  L.seg_offsets = at; at = align_up(at + (size_t)n * L.seg_pitch * sizeof(uint32_t), 256);
  L.status = at; at += 256;
  L.compressed = at; at = align_up(at + compressed_capacity + 16 * (size_t)n + 64, 256);
  // the scans without stuffing and markers; a corrupt segment is read up to one block (64 x 31 bits) past its end
  L.clean = at; at = align_up(at + compressed_capacity + 16 * (size_t)n + 512, 256);
  L.coef = at; at = align_up(at + (size_t)n * g.coef_per_image * sizeof(int16_t), 256);
  L.planes = at; at = align_up(at + (size_t)n * g.plane_per_image, 256);
  L.total = at;
  return L;
}

int geometry_from_args(int width, int height, int ncomp, int hmax, int vmax, JpegGeometry* g) {
  ParsedImage im;
  im.W = width; im.H = height; im.ncomp = ncomp;
  for (int i = 0; i < 3; i++) { im.h[i] = im.v[i] = 1; }
  im.h[0] = hmax; im.v[0] = vmax;
  if (ncomp != 1 && ncomp != 3) { set_last_error("jpeg: %d components", ncomp); return AIKO_B200_ERR_UNSUPPORTED; }
  return make_geometry(im, g);
}

}  // namespace

extern "C" int aiko_b200_jpeg_gpu_info_host(const uint8_t* jpeg_host, size_t length, int* width_host, int* height_host,
                                        int* components_host, int* hmax_host, int* vmax_host,
                                        int* restart_interval_host) {
  AIKO_CHECK_ARG(jpeg_host && width_host && height_host, "jpeg_info: null argument");
  ParsedImage im;
  const int rc = parse_jpeg(jpeg_host, length, &im, false);
  if (rc != AIKO_B200_OK) return rc;
  *width_host = im.W; *height_host = im.H;
  if (components_host) *components_host = im.ncomp;
  if (hmax_host) *hmax_host = im.ncomp == 3 ? im.h[0] : 1;
  if (vmax_host) *vmax_host = im.ncomp == 3 ? im.v[0] : 1;
  if (restart_interval_host) *restart_interval_host = im.restart;
  return AIKO_B200_OK;
}

extern "C" size_t aiko_b200_jpeg_gpu_staging_bytes(int n_images) {
  if (n_images <= 0) return 0;
  return align_up((size_t)n_images * sizeof(JpegImage), 256) + align_up((size_t)n_images * sizeof(JpegTables), 256);
}

extern "C" size_t aiko_b200_jpeg_gpu_workspace_bytes(int n_images, int width, int height, int components, int hmax, int vmax,
                                                 size_t compressed_capacity, int max_segments) {
  JpegGeometry g;
  if (n_images <= 0 || max_segments <= 0 || max_segments > kMaxSegments ||
      geometry_from_args(width, height, components, hmax, vmax, &g) != AIKO_B200_OK)
    return 0;
  return layout_for(n_images, g, compressed_capacity, max_segments).total;
}

extern "C" int aiko_b200_jpeg_gpu_decode_rgb_u8(const uint8_t* const* jpegs_host, const size_t* lengths_host, int n_images,
                                            uint8_t* dst, int height, int width, int64_t frame_stride,
                                            void* staging_host, size_t staging_bytes, void* workspace,
                                            size_t workspace_bytes, size_t compressed_capacity, int max_segments,
                                            void* stream) {
  AIKO_CHECK_ARG(n_images >= 0, "jpeg_decode: n_images < 0");
  if (n_images == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(jpegs_host && lengths_host && dst && staging_host && workspace, "jpeg_decode: null pointer");
  AIKO_CHECK_ARG(frame_stride >= (int64_t)height * width * 3, "jpeg_decode: frame stride smaller than a frame");
  AIKO_CHECK_ARG(staging_bytes >= aiko_b200_jpeg_gpu_staging_bytes(n_images), "jpeg_decode: staging buffer too small");
  AIKO_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "jpeg_decode: workspace must be 256-byte aligned");
  cudaStream_t st = static_cast<cudaStream_t>(stream);

  // ---- parse every image; all must share one geometry
  std::vector<ParsedImage> parsed((size_t)n_images);
  JpegGeometry g{};
  size_t scan_total = 0;
  int needed_segments = 1;
  bool any_few = false, any_many = false;   // restart segments per image: which entropy kernels run
  for (int i = 0; i < n_images; i++) {
    const int rc = parse_jpeg(jpegs_host[i], lengths_host[i], &parsed[i], false);
    if (rc != AIKO_B200_OK) return rc;
    JpegGeometry gi;
    const int rg = make_geometry(parsed[i], &gi);
    if (rg != AIKO_B200_OK) return rg;
    if (i == 0) g = gi;
    if (gi.W != g.W || gi.H != g.H || gi.ncomp != g.ncomp || gi.hmax != g.hmax || gi.vmax != g.vmax) {
      set_last_error("jpeg_decode: image %d is %dx%d (%d components, %dx%d sampling), the batch is %dx%d (%d, %dx%d)", i,
                     gi.W, gi.H, gi.ncomp, gi.hmax, gi.vmax, g.W, g.H, g.ncomp, g.hmax, g.vmax);
      return AIKO_B200_ERR_INVALID_ARGUMENT;
    }
    const int segs = parsed[i].restart > 0 ? (g.mcux * g.mcuy + parsed[i].restart - 1) / parsed[i].restart : 1;
    any_few |= segs <= kSyncMaxSegments;
    any_many |= segs > kSyncMaxSegments;
    needed_segments = segs > needed_segments ? segs : needed_segments;
    scan_total += align_up(parsed[i].scan_bytes, 16);
  }
  AIKO_CHECK_ARG(g.W == width && g.H == height, "jpeg_decode: images are %dx%d, the destination is %dx%d", g.W, g.H, width, height);
  if (needed_segments > max_segments || scan_total > compressed_capacity + 16 * (size_t)n_images) {
    set_last_error("jpeg_decode: workspace sized for %d segments / %zu compressed bytes, the batch needs %d / %zu",
                   max_segments, compressed_capacity, needed_segments, scan_total);
    return AIKO_B200_ERR_WORKSPACE;
  }
  const WorkspaceLayout L = layout_for(n_images, g, compressed_capacity, max_segments);
  if (workspace_bytes < L.total) {
    set_last_error("jpeg_decode: workspace of %zu bytes, need %zu", workspace_bytes, L.total);
    return AIKO_B200_ERR_WORKSPACE;
  }

  // ---- descriptors and (de-duplicated) tables into the staging buffer
  uint8_t* stage = static_cast<uint8_t*>(staging_host);
  JpegImage* images = reinterpret_cast<JpegImage*>(stage + L.images);
  JpegTables* tables = reinterpret_cast<JpegTables*>(stage + L.tables);
  int n_sets = 0;
  size_t scan_at = 0;
  for (int i = 0; i < n_images; i++) {
    const ParsedImage& im = parsed[i];
    int set = -1;
    for (int k = n_sets - 1; k >= 0 && set < 0; k--)
      if (memcmp(&tables[k], &im.tables, sizeof(JpegTables)) == 0) set = k;
    if (set < 0) {
      set = n_sets++;
      memcpy(&tables[set], &im.tables, sizeof(JpegTables));
    }
    JpegImage& d = images[i];
    memset(&d, 0, sizeof(d));
    d.scan_offset = (uint32_t)scan_at;
    d.scan_bytes = (uint32_t)im.scan_bytes;
    d.restart_interval = im.restart > 0 ? im.restart : g.mcux * g.mcuy;
    d.n_segments = (g.mcux * g.mcuy + d.restart_interval - 1) / d.restart_interval;
    d.table_set = set;
    for (int c = 0; c < g.ncomp; c++) { d.td[c] = (uint8_t)im.td[c]; d.ta[c] = (uint8_t)im.ta[c]; d.tq[c] = (uint8_t)im.tq[c]; }
    scan_at += align_up(im.scan_bytes, 16);
  }
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  const size_t head = L.tables + (size_t)n_sets * sizeof(JpegTables);
  AIKO_CUDA(cudaMemcpyAsync(ws, stage, head, cudaMemcpyHostToDevice, st));
  for (int i = 0; i < n_images; i++)              // the compressed scans travel as they are (asynchronously when page-locked)
    AIKO_CUDA(cudaMemcpyAsync(ws + L.compressed + images[i].scan_offset, parsed[i].scan, parsed[i].scan_bytes,
                              cudaMemcpyHostToDevice, st));
  // diagnostic: AIKO_B200_JPEG_STAGES = bit mask of the stages to run (1 segments, 2 entropy decode, 4 IDCT, 8 colour)
  const char* stage_env = getenv("AIKO_B200_JPEG_STAGES");
  const int stages = stage_env ? atoi(stage_env) : 15;
  AIKO_CUDA(cudaMemsetAsync(ws + L.status, 0, 256, st));
  if (stages & 2) AIKO_CUDA(cudaMemsetAsync(ws + L.coef, 0, (size_t)n_images * g.coef_per_image * sizeof(int16_t), st));

  const JpegImage* d_images = reinterpret_cast<const JpegImage*>(ws + L.images);
  const JpegTables* d_tables = reinterpret_cast<const JpegTables*>(ws + L.tables);
  uint32_t* d_seg = reinterpret_cast<uint32_t*>(ws + L.seg_offsets);
  int* d_status = reinterpret_cast<int*>(ws + L.status);
  int16_t* d_coef = reinterpret_cast<int16_t*>(ws + L.coef);
  uint8_t* d_planes = ws + L.planes;

  if (stages & 1)
    jpeg_unstuff_kernel<<<(unsigned)n_images, kSegThreads, 0, st>>>(ws + L.compressed, d_images, ws + L.clean, d_seg, L.seg_pitch, d_status);
  AIKO_CUDA(cudaGetLastError());
  if ((stages & 2) && any_many)
    jpeg_huffman_kernel<<<dim3((unsigned)((needed_segments + kHuffThreads - 1) / kHuffThreads), (unsigned)n_images),
                          kHuffThreads, 0, st>>>(ws + L.clean, d_images, d_tables, d_seg, L.seg_pitch, g, d_coef, d_status);
  AIKO_CUDA(cudaGetLastError());
  if ((stages & 2) && any_few)
    jpeg_sync_huffman_kernel<<<(unsigned)n_images, kSyncThreads, sizeof(SyncShared), st>>>(
        ws + L.clean, d_images, d_tables, d_seg, L.seg_pitch, g, d_coef, d_status);
  AIKO_CUDA(cudaGetLastError());
  const long long blocks = g.coef_per_image / 64;
  if (stages & 4) jpeg_idct_kernel<<<dim3((unsigned)((blocks + kIdctThreads - 1) / kIdctThreads), (unsigned)n_images), kIdctThreads, 0, st>>>(
      d_coef, d_images, d_tables, g, d_planes);
  AIKO_CUDA(cudaGetLastError());
  const long long items = (long long)((g.W + 7) / 8) * g.H;
  const dim3 colour_grid((unsigned)((items + 255) / 256), (unsigned)n_images);
  if (getenv("AIKO_B200_JPEG_DEBUG")) {                 // diagnostic: rounds of the state iteration (synchronises)
    int dbg[4] = {0, 0, 0, 0};
    AIKO_CUDA(cudaMemcpyAsync(dbg, ws + L.status, sizeof(dbg), cudaMemcpyDeviceToHost, st));
    AIKO_CUDA(cudaStreamSynchronize(st));
    fprintf(stderr, "jpeg_decode: status %d, state iteration rounds: max %d, mean %.2f over %d images\n", dbg[0], dbg[1],
            (double)dbg[2] / n_images, n_images);
  }
  if (!(stages & 8)) return AIKO_B200_OK;
  if (g.ncomp == 1) jpeg_colour_kernel<0><<<colour_grid, 256, 0, st>>>(d_planes, g, dst, frame_stride);
  else if (g.hmax == 1) jpeg_colour_kernel<1><<<colour_grid, 256, 0, st>>>(d_planes, g, dst, frame_stride);
  else if (g.vmax == 1) jpeg_colour_kernel<2><<<colour_grid, 256, 0, st>>>(d_planes, g, dst, frame_stride);
  else jpeg_colour_kernel<3><<<colour_grid, 256, 0, st>>>(d_planes, g, dst, frame_stride);
  return cuda_status(cudaGetLastError(), "jpeg_colour launch");
}

// 0 when the last decode that used `workspace` on `stream` met no corrupt entropy data; synchronises the stream.
extern "C" int aiko_b200_jpeg_gpu_status(const void* workspace, int n_images, int width, int height, int components, int hmax,
                                     int vmax, size_t compressed_capacity, int max_segments, void* stream) {
  JpegGeometry g;
  AIKO_CHECK_ARG(workspace && n_images > 0, "jpeg_status: bad argument");
  const int rg = geometry_from_args(width, height, components, hmax, vmax, &g);
  if (rg != AIKO_B200_OK) return rg;
  const WorkspaceLayout L = layout_for(n_images, g, compressed_capacity, max_segments);
  int status = 0;
  AIKO_CUDA(cudaMemcpyAsync(&status, static_cast<const uint8_t*>(workspace) + L.status, sizeof(int), cudaMemcpyDeviceToHost,
                            static_cast<cudaStream_t>(stream)));
  AIKO_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  if (status != 0) {
    set_last_error("jpeg: corrupt entropy-coded data (code %d: 1+i restart markers of image i do not match its restart "
                   "interval, 1000+i more than two Huffman tables, 2000+i invalid code, 3000+i coefficient index past 63, "
                   "4000+i a segment with too few or too many blocks)", status);
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  return AIKO_B200_OK;
}
