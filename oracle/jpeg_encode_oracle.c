/* oracle/jpeg_encode_oracle.c -- TEST INFRASTRUCTURE ONLY (never linked or imported by the product).
 *
 * CPU restatement of the baseline JPEG ENCODER the reference reaches through
 * ImageWriteFile: image.save(path) for a ".jpg" path (R/elements/media/image_io.py:250), i.e.
 * Pillow's JpegEncode.c driving libjpeg-turbo with its defaults: quality 75 (any quality 1..100
 * here), 4:2:0 chroma subsampling for RGB images (4:4:4 and 4:2:2 also restated), gray for mode L,
 * standard Huffman tables (optimize=False), no restart markers, JFIF 1.01 header with density 1:1.
 * Restated from the published algorithm of libjpeg / libjpeg-turbo 3.x:
 *   jccolor.c  rgb_ycc_convert     16-bit fixed-point RGB -> YCbCr
 *   jcsample.c h2v1 / h2v2_downsample  box filter with the alternating bias, right edge replicated
 *   jcprepct.c / jccoefct.c        bottom edge replicated; dummy blocks: zero AC, DC of the block before
 *   jfdctint.c jpeg_fdct_islow     accurate integer forward DCT (CONST_BITS 13, PASS1_BITS 2)
 *   jcdctmgr.c quantize            sign(x) * ((|x| + q/2) / q) on the 8x-scaled DCT output
 *   jchuff.c   encode_one_block    Huffman coding, FF stuffing, final padding with 1 bits
 *   jcmarker.c                     SOI, APP0, DQT, SOF0, DHT, SOS, EOI
 * Pinned: the bytes equal Pillow's for random and natural images of every layout and size
 * (tests/test_oracle_golden.py). */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "jpeg_std_tables.inc"

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

typedef struct {
  uint8_t* p;
  size_t cap, n;
  uint32_t acc;      /* bits not yet written, right-aligned */
  int nbits;
} Out;

static void put_byte(Out* o, int b) {
  if (o->n < o->cap) o->p[o->n] = (uint8_t)b;
  o->n++;
}
static void put16(Out* o, int v) { put_byte(o, v >> 8); put_byte(o, v & 255); }

static void emit_bits(Out* o, unsigned code, int size) {           /* jchuff.c emit_bits */
  if (size == 0) return;
  o->acc = (o->acc << size) | (code & ((1u << size) - 1u));
  o->nbits += size;
  while (o->nbits >= 8) {
    const int b = (int)((o->acc >> (o->nbits - 8)) & 255u);
    put_byte(o, b);
    if (b == 0xff) put_byte(o, 0);
    o->nbits -= 8;
  }
}

typedef struct { unsigned code[256]; int size[256]; } HuffEnc;

static void build_enc(HuffEnc* h, const uint8_t* bits, const uint8_t* vals) {   /* jchuff.c jpeg_make_c_derived_tbl */
  memset(h, 0, sizeof(*h));
  unsigned code = 0;
  int k = 0;
  for (int len = 1; len <= 16; len++) {
    for (int i = 0; i < bits[len - 1]; i++, k++, code++) { h->code[vals[k]] = code; h->size[vals[k]] = len; }
    code <<= 1;
  }
}

static int bit_length(int v) { int n = 0; while (v) { n++; v >>= 1; } return n; }

#define FIX_0_298631336 2446
#define FIX_0_390180644 3196
#define FIX_0_541196100 4433
#define FIX_0_765366865 6270
#define FIX_0_899976223 7373
#define FIX_1_175875602 9633
#define FIX_1_501321110 12299
#define FIX_1_847759065 15137
#define FIX_1_961570560 16069
#define FIX_2_053119869 16819
#define FIX_2_562915447 20995
#define FIX_3_072711026 25172
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))

static void fdct_1d(const int* d, int stride, int* out, int first_pass) {
  const int d0 = d[0], d1 = d[stride], d2 = d[2 * stride], d3 = d[3 * stride], d4 = d[4 * stride], d5 = d[5 * stride],
            d6 = d[6 * stride], d7 = d[7 * stride];
  int tmp0 = d0 + d7, tmp7 = d0 - d7, tmp1 = d1 + d6, tmp6 = d1 - d6, tmp2 = d2 + d5, tmp5 = d2 - d5, tmp3 = d3 + d4,
      tmp4 = d3 - d4;
  const int tmp10 = tmp0 + tmp3, tmp13 = tmp0 - tmp3, tmp11 = tmp1 + tmp2, tmp12 = tmp1 - tmp2;
  const int shift = first_pass ? 13 - 2 : 13 + 2;
  if (first_pass) {
    out[0] = (tmp10 + tmp11) * 4;
    out[4 * stride] = (tmp10 - tmp11) * 4;
  } else {
    out[0] = DESCALE(tmp10 + tmp11, 2);
    out[4 * stride] = DESCALE(tmp10 - tmp11, 2);
  }
  int z1 = (tmp12 + tmp13) * FIX_0_541196100;
  out[2 * stride] = DESCALE(z1 + tmp13 * FIX_0_765366865, shift);
  out[6 * stride] = DESCALE(z1 + tmp12 * (-FIX_1_847759065), shift);
  z1 = tmp4 + tmp7;
  int z2 = tmp5 + tmp6, z3 = tmp4 + tmp6, z4 = tmp5 + tmp7;
  const int z5 = (z3 + z4) * FIX_1_175875602;
  tmp4 *= FIX_0_298631336; tmp5 *= FIX_2_053119869; tmp6 *= FIX_3_072711026; tmp7 *= FIX_1_501321110;
  z1 *= -FIX_0_899976223; z2 *= -FIX_2_562915447; z3 *= -FIX_1_961570560; z4 *= -FIX_0_390180644;
  z3 += z5; z4 += z5;
  out[7 * stride] = DESCALE(tmp4 + z1 + z3, shift);
  out[5 * stride] = DESCALE(tmp5 + z2 + z4, shift);
  out[3 * stride] = DESCALE(tmp6 + z2 + z3, shift);
  out[1 * stride] = DESCALE(tmp7 + z1 + z4, shift);
}

/* 8x8 samples (already level-shifted) -> quantised coefficients in natural order */
static void fdct_quantise(const int* samples, const uint8_t* q, int16_t* coef) {
  int w[64];
  for (int r = 0; r < 8; r++) fdct_1d(samples + 8 * r, 1, w + 8 * r, 1);
  for (int c = 0; c < 8; c++) fdct_1d(w + c, 8, w + c, 0);
  for (int i = 0; i < 64; i++) {
    const int div = q[i] * 8;
    const int t = w[i];
    coef[i] = (int16_t)(t < 0 ? -((-t + (div >> 1)) / div) : (t + (div >> 1)) / div);
  }
}

static void scaled_quant(const uint8_t* base, int quality, uint8_t* out) {       /* jcparam.c */
  if (quality < 1) quality = 1;
  if (quality > 100) quality = 100;
  const int scale = quality < 50 ? 5000 / quality : 200 - quality * 2;
  for (int i = 0; i < 64; i++) {
    long t = ((long)base[i] * scale + 50) / 100;
    if (t < 1) t = 1;
    if (t > 255) t = 255;
    out[i] = (uint8_t)t;
  }
}

static void put_dht(Out* o, int id, const uint8_t* bits, const uint8_t* vals) {
  int total = 0;
  for (int i = 0; i < 16; i++) total += bits[i];
  put_byte(o, 0xff); put_byte(o, 0xc4); put16(o, 19 + total);
  put_byte(o, id);
  for (int i = 0; i < 16; i++) put_byte(o, bits[i]);
  for (int i = 0; i < total; i++) put_byte(o, vals[i]);
}

/* rgb: h x w x channels (3 = RGB, 1 = gray), tight rows.  hs, vs: luma sampling factors (2,2 = 4:2:0,
 * 2,1 = 4:2:2, 1,1 = 4:4:4; ignored for gray).  -> bytes of the file (may exceed cap: nothing is written
 * past it). */
size_t oracle_jpeg_encode(const uint8_t* rgb, int h, int w, int channels, int quality, int hs, int vs, uint8_t* out,
                          size_t cap) {
  Out o;
  o.p = out; o.cap = cap; o.n = 0; o.acc = 0; o.nbits = 0;
  const int ncomp = channels == 1 ? 1 : 3;
  if (ncomp == 1) { hs = 1; vs = 1; }
  uint8_t q[2][64];
  scaled_quant(kStdLuminanceQuant, quality, q[0]);
  scaled_quant(kStdChrominanceQuant, quality, q[1]);
  /* ---- headers (jcmarker.c) */
  put_byte(&o, 0xff); put_byte(&o, 0xd8);
  {
    static const uint8_t app0[] = {0xff, 0xe0, 0, 16, 'J', 'F', 'I', 'F', 0, 1, 1, 0, 0, 1, 0, 1, 0, 0};
    for (size_t i = 0; i < sizeof(app0); i++) put_byte(&o, app0[i]);
  }
  for (int t = 0; t < (ncomp == 1 ? 1 : 2); t++) {
    put_byte(&o, 0xff); put_byte(&o, 0xdb); put16(&o, 67); put_byte(&o, t);
    for (int k = 0; k < 64; k++) put_byte(&o, q[t][kZigzag[k]]);
  }
  put_byte(&o, 0xff); put_byte(&o, 0xc0); put16(&o, 8 + 3 * ncomp); put_byte(&o, 8); put16(&o, h); put16(&o, w);
  put_byte(&o, ncomp);
  for (int c = 0; c < ncomp; c++) {
    put_byte(&o, c + 1);
    put_byte(&o, c == 0 ? (hs << 4) | vs : 0x11);
    put_byte(&o, c == 0 ? 0 : 1);
  }
  put_dht(&o, 0x00, kStdDcLuminanceBits, kStdDcLuminanceVals);
  put_dht(&o, 0x10, kStdAcLuminanceBits, kStdAcLuminanceVals);
  if (ncomp == 3) {
    put_dht(&o, 0x01, kStdDcChrominanceBits, kStdDcChrominanceVals);
    put_dht(&o, 0x11, kStdAcChrominanceBits, kStdAcChrominanceVals);
  }
  put_byte(&o, 0xff); put_byte(&o, 0xda); put16(&o, 6 + 2 * ncomp); put_byte(&o, ncomp);
  for (int c = 0; c < ncomp; c++) { put_byte(&o, c + 1); put_byte(&o, c == 0 ? 0x00 : 0x11); }
  put_byte(&o, 0); put_byte(&o, 63); put_byte(&o, 0);
  /* ---- component planes, padded to whole blocks */
  const int mcux = (w + 8 * hs - 1) / (8 * hs), mcuy = (h + 8 * vs - 1) / (8 * vs);
  int cw[3], ch[3], bw[3], bh[3], ph[3], pv[3];
  uint8_t* plane[3] = {0, 0, 0};
  for (int c = 0; c < ncomp; c++) {
    ph[c] = c == 0 ? hs : 1; pv[c] = c == 0 ? vs : 1;
    cw[c] = (w * ph[c] + hs - 1) / hs; ch[c] = (h * pv[c] + vs - 1) / vs;      /* downsampled size */
    bw[c] = (cw[c] + 7) / 8; bh[c] = (ch[c] + 7) / 8;                           /* real blocks */
    plane[c] = (uint8_t*)malloc((size_t)bw[c] * 8 * bh[c] * 8);
  }
  for (int c = 0; c < ncomp; c++) {
    const int pw = bw[c] * 8, fx = hs / ph[c], fy = vs / pv[c];                  /* source pixels per sample */
    for (int y = 0; y < bh[c] * 8; y++) {
      const int sy = y < ch[c] ? y : ch[c] - 1;                                 /* bottom edge: the last sample row again */
      for (int x = 0; x < pw; x++) {
        int sum = 0;
        for (int dy = 0; dy < fy; dy++)
          for (int dx = 0; dx < fx; dx++) {
            int yy = sy * fy + dy, xx = x * fx + dx;
            if (yy > h - 1) yy = h - 1;                                          /* jcprepct.c expand_bottom_edge of the colour rows */
            if (xx > w - 1) xx = w - 1;                                          /* jcsample.c expand_right_edge */
            const uint8_t* px = rgb + ((size_t)yy * w + xx) * channels;
            int v;
            if (ncomp == 1) {
              v = px[0];
            } else {
              const int r = px[0], g = px[1], b = px[2];
              if (c == 0) v = (19595 * r + 38470 * g + 7471 * b + 32768) >> 16;
              else if (c == 1) v = (-11059 * r - 21709 * g + 32768 * b + (128 << 16) + 32767) >> 16;
              else v = (32768 * r - 27439 * g - 5329 * b + (128 << 16) + 32767) >> 16;
            }
            sum += v;
          }
        int s;
        if (fx == 1 && fy == 1) s = sum;
        else if (fy == 1) s = (sum + (x & 1)) >> 1;                              /* h2v1: bias 0, 1, 0, 1 */
        else s = (sum + ((x & 1) ? 2 : 1)) >> 2;                                 /* h2v2: bias 1, 2, 1, 2 */
        plane[c][(size_t)y * pw + x] = (uint8_t)s;
      }
    }
  }
  /* ---- blocks in MCU order */
  HuffEnc dc[2], ac[2];
  build_enc(&dc[0], kStdDcLuminanceBits, kStdDcLuminanceVals);
  build_enc(&ac[0], kStdAcLuminanceBits, kStdAcLuminanceVals);
  build_enc(&dc[1], kStdDcChrominanceBits, kStdDcChrominanceVals);
  build_enc(&ac[1], kStdAcChrominanceBits, kStdAcChrominanceVals);
  int last_dc[3] = {0, 0, 0};
  for (int my = 0; my < mcuy; my++)
    for (int mx = 0; mx < mcux; mx++)
      for (int c = 0; c < ncomp; c++) {
        int16_t prev_in_mcu_dc = 0;        /* DC of the block before in MCU_buffer order (jccoefct.c dummy blocks) */
        int have_prev = 0;
        for (int by = 0; by < pv[c]; by++)
          for (int bx = 0; bx < ph[c]; bx++) {
            int16_t coef[64];
            const int gx = mx * ph[c] + bx, gy = my * pv[c] + by;
            if (gx < bw[c] && gy < bh[c]) {
              int samples[64];
              for (int y = 0; y < 8; y++)
                for (int x = 0; x < 8; x++) samples[8 * y + x] = plane[c][(size_t)(gy * 8 + y) * bw[c] * 8 + gx * 8 + x] - 128;
              fdct_quantise(samples, q[c == 0 ? 0 : 1], coef);
            } else {
              memset(coef, 0, sizeof(coef));
              coef[0] = have_prev ? prev_in_mcu_dc : 0;
            }
            prev_in_mcu_dc = coef[0];
            have_prev = 1;
            /* jchuff.c encode_one_block */
            const HuffEnc* hd = &dc[c == 0 ? 0 : 1];
            const HuffEnc* ha = &ac[c == 0 ? 0 : 1];
            int temp = coef[0] - last_dc[c], temp2 = temp;
            last_dc[c] = coef[0];
            if (temp < 0) { temp = -temp; temp2--; }
            int nbits = bit_length(temp);
            emit_bits(&o, hd->code[nbits], hd->size[nbits]);
            if (nbits) emit_bits(&o, (unsigned)temp2, nbits);
            int r = 0;
            for (int k = 1; k < 64; k++) {
              temp = coef[kZigzag[k]];
              if (temp == 0) { r++; continue; }
              while (r > 15) { emit_bits(&o, ha->code[0xf0], ha->size[0xf0]); r -= 16; }
              temp2 = temp;
              if (temp < 0) { temp = -temp; temp2--; }
              nbits = bit_length(temp);
              emit_bits(&o, ha->code[(r << 4) + nbits], ha->size[(r << 4) + nbits]);
              emit_bits(&o, (unsigned)temp2, nbits);
              r = 0;
            }
            if (r > 0) emit_bits(&o, ha->code[0], ha->size[0]);
          }
      }
  emit_bits(&o, 0x7f, 7);                  /* jchuff.c flush_bits: pad the last byte with ones */
  o.nbits = 0;
  put_byte(&o, 0xff); put_byte(&o, 0xd9);
  for (int c = 0; c < ncomp; c++) free(plane[c]);
  return o.n;
}
