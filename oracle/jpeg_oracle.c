/*
 * jpeg_oracle.c -- CPU restatement of baseline JPEG decoding as libjpeg(-turbo) does it with its
 * default settings, which is what PIL.Image.open(...).load() runs for the reference's ImageReadFile /
 * ImageResize (R/elements/media/image_io.py:184,218): Huffman entropy decoding, dequantisation,
 * the accurate integer inverse DCT (jidctint.c "islow", CONST_BITS 13, PASS1_BITS 2), "fancy"
 * triangle-filter chroma upsampling (jdsample.c h2v1 / h2v2) and the 16-bit fixed-point YCbCr -> RGB
 * tables (jdcolor.c).  libjpeg-turbo is a dependency of Pillow and is not vendored by the reference;
 * this follows its published algorithm.
 *
 * TEST INFRASTRUCTURE: the checker of the CUDA decoder (csrc/jpeg_kernels.cu).  Pinned against Pillow
 * itself in tests/test_oracle_golden.py (bit-exact on random images of every supported layout).
 *
 * Subset: SOF0 (baseline, 8 bit), one interleaved scan, 1 component (gray) or 3 (YCbCr) with luma
 * sampling 1x1, 2x1 or 2x2 and chroma 1x1; restart intervals allowed.  Anything else: return -2.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int bits[17];           /* bits[k] = number of codes of length k */
  uint8_t vals[256];
  int mincode[17], maxcode[18], valptr[17];
  int present;
} Huff;

typedef struct {
  int id, h, v, tq, td, ta;
  int bw, bh;             /* blocks per row / column (padded to whole MCUs) */
  int pw, ph;             /* plane size in samples = bw * 8, bh * 8 */
  int dw, dh;             /* downsampled (true) size: ceil(image * h / hmax) */
  int16_t* coef;          /* bw * bh blocks of 64, natural order, not dequantised */
  uint8_t* plane;
} Comp;

typedef struct {
  const uint8_t* p;
  const uint8_t* end;
  uint32_t acc;
  int nbits;
  int marker;             /* a marker was reached: feed zeros */
} Bits;

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

static void huff_build(Huff* h) {
  int code = 0, k = 0;
  for (int len = 1; len <= 16; len++) {
    h->valptr[len] = k;
    h->mincode[len] = code;
    code += h->bits[len];
    k += h->bits[len];
    h->maxcode[len] = h->bits[len] ? code - 1 : -1;
    code <<= 1;
  }
  h->maxcode[17] = 0x7fffffff;
  h->present = 1;
}

static void fill(Bits* b) {
  while (b->nbits <= 24) {
    int byte = 0;
    if (!b->marker && b->p < b->end) {
      byte = *b->p;
      if (byte == 0xff) {
        if (b->p + 1 < b->end && b->p[1] == 0) b->p += 2;      /* stuffed zero */
        else { b->marker = 1; byte = 0; }                        /* a marker: stay on it */
      } else {
        b->p++;
      }
    }
    b->acc |= (uint32_t)byte << (24 - b->nbits);
    b->nbits += 8;
  }
}

static int get_bits(Bits* b, int n) {
  if (n == 0) return 0;
  fill(b);
  int v = (int)(b->acc >> (32 - n));
  b->acc <<= n;
  b->nbits -= n;
  return v;
}

static int decode_symbol(Bits* b, const Huff* h) {
  fill(b);
  int code = 0;
  for (int len = 1; len <= 16; len++) {
    code = (int)(b->acc >> (32 - len));
    if (h->maxcode[len] >= 0 && code <= h->maxcode[len] && code >= h->mincode[len]) {
      b->acc <<= len;
      b->nbits -= len;
      return h->vals[h->valptr[len] + code - h->mincode[len]];
    }
  }
  return -1;
}

static int extend(int v, int s) { return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v; }

/* ---- jidctint.c: jpeg_idct_islow ---------------------------------------------------------------- */
#define CONST_BITS 13
#define PASS1_BITS 2
#define DESCALE(x, n) (((x) + (1 << ((n)-1))) >> (n))
enum {
  F_0_298631336 = 2446, F_0_390180644 = 3196, F_0_541196100 = 4433, F_0_765366865 = 6270,
  F_0_899976223 = 7373, F_1_175875602 = 9633, F_1_501321110 = 12299, F_1_847759065 = 15137,
  F_1_961570560 = 16069, F_2_053119869 = 16819, F_2_562915447 = 20995, F_3_072711026 = 25172
};

static void idct_1d(const int* in, int stride, int* out, int ostride, int shift) {
  int z1, z2, z3, z4, z5, tmp0, tmp1, tmp2, tmp3, tmp10, tmp11, tmp12, tmp13;
  z2 = in[2 * stride]; z3 = in[6 * stride];
  z1 = (z2 + z3) * F_0_541196100;
  tmp2 = z1 + z3 * (-F_1_847759065);
  tmp3 = z1 + z2 * F_0_765366865;
  z2 = in[0]; z3 = in[4 * stride];
  tmp0 = (z2 + z3) * (1 << CONST_BITS);
  tmp1 = (z2 - z3) * (1 << CONST_BITS);
  tmp10 = tmp0 + tmp3; tmp13 = tmp0 - tmp3; tmp11 = tmp1 + tmp2; tmp12 = tmp1 - tmp2;
  tmp0 = in[7 * stride]; tmp1 = in[5 * stride]; tmp2 = in[3 * stride]; tmp3 = in[1 * stride];
  z1 = tmp0 + tmp3; z2 = tmp1 + tmp2; z3 = tmp0 + tmp2; z4 = tmp1 + tmp3;
  z5 = (z3 + z4) * F_1_175875602;
  tmp0 *= F_0_298631336; tmp1 *= F_2_053119869; tmp2 *= F_3_072711026; tmp3 *= F_1_501321110;
  z1 *= -F_0_899976223; z2 *= -F_2_562915447; z3 *= -F_1_961570560; z4 *= -F_0_390180644;
  z3 += z5; z4 += z5;
  tmp0 += z1 + z3; tmp1 += z2 + z4; tmp2 += z2 + z3; tmp3 += z1 + z4;
  out[0 * ostride] = DESCALE(tmp10 + tmp3, shift);
  out[7 * ostride] = DESCALE(tmp10 - tmp3, shift);
  out[1 * ostride] = DESCALE(tmp11 + tmp2, shift);
  out[6 * ostride] = DESCALE(tmp11 - tmp2, shift);
  out[2 * ostride] = DESCALE(tmp12 + tmp1, shift);
  out[5 * ostride] = DESCALE(tmp12 - tmp1, shift);
  out[3 * ostride] = DESCALE(tmp13 + tmp0, shift);
  out[4 * ostride] = DESCALE(tmp13 - tmp0, shift);
}

static void idct_block(const int16_t* coef, const uint16_t* q, uint8_t* out, int pitch) {
  int in[64], ws[64], res[64];
  for (int i = 0; i < 64; i++) in[i] = coef[i] * (int)q[i];
  for (int c = 0; c < 8; c++) idct_1d(in + c, 8, ws + c, 8, CONST_BITS - PASS1_BITS);       /* columns */
  for (int r = 0; r < 8; r++) idct_1d(ws + 8 * r, 1, res + 8 * r, 1, CONST_BITS + PASS1_BITS + 3); /* rows */
  for (int r = 0; r < 8; r++)
    for (int c = 0; c < 8; c++) {
      int v = res[8 * r + c] + 128;
      out[r * pitch + c] = (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
    }
}

/* ---- jdsample.c: fancy upsampling of one chroma plane to full resolution --------------------------- */
static uint8_t row_at(const Comp* c, int y, int x) {      /* rows above / below the image repeat the edge row */
  if (y < 0) y = 0;
  if (y > c->dh - 1) y = c->dh - 1;
  return c->plane[(size_t)y * c->pw + x];
}

/* out: W x H samples of the component at full resolution */
static void upsample(const Comp* c, int hmax, int vmax, int W, int H, uint8_t* out) {
  const int hs = hmax / c->h, vs = vmax / c->v;
  if (hs == 1 && vs == 1) {
    for (int y = 0; y < H; y++) memcpy(out + (size_t)y * W, c->plane + (size_t)y * c->pw, (size_t)W);
    return;
  }
  const int dw = c->dw;
  if (dw <= 2) {
    /* jdsample.c jinit_upsampler: the fancy filters are only installed for downsampled_width > 2; narrower
     * components are upsampled by replication (h2v1_upsample / h2v2_upsample) */
    for (int y = 0; y < H; y++)
      for (int x = 0; x < W; x++) out[(size_t)y * W + x] = c->plane[(size_t)(vs == 2 ? y >> 1 : y) * c->pw + (x >> 1)];
    return;
  }
  int* sums = (int*)malloc(sizeof(int) * (size_t)dw);
  for (int y = 0; y < H; y++) {
    const int sy = vs == 2 ? y >> 1 : y;
    /* vertical: 3/4 nearer row + 1/4 further row (h2v2), or the row itself times 4 (h2v1) */
    for (int x = 0; x < dw; x++) {
      if (vs == 2) sums[x] = row_at(c, sy, x) * 3 + row_at(c, (y & 1) ? sy + 1 : sy - 1, x);
      else sums[x] = row_at(c, sy, x);
    }
    uint8_t* o = out + (size_t)y * W;
    for (int x = 0; x < W; x++) {
      const int sx = x >> 1;
      int v;
      if (vs == 2) {
        if (dw == 1) v = (sums[0] * 4 + ((x & 1) ? 7 : 8)) >> 4;
        else if (x == 0) v = (sums[0] * 4 + 8) >> 4;
        else if (sx == dw - 1 && (x & 1)) v = (sums[sx] * 4 + 7) >> 4;
        else if (x & 1) v = (sums[sx] * 3 + sums[sx + 1] + 7) >> 4;
        else v = (sums[sx] * 3 + sums[sx - 1] + 8) >> 4;
      } else {          /* h2v1 */
        if (dw == 1) v = sums[0];
        else if (x == 0) v = sums[0];
        else if (sx == dw - 1 && (x & 1)) v = sums[sx];
        else if (x & 1) v = (sums[sx] * 3 + sums[sx + 1] + 2) >> 2;
        else v = (sums[sx] * 3 + sums[sx - 1] + 1) >> 2;
      }
      o[x] = (uint8_t)v;
    }
  }
  free(sums);
}

static int be16(const uint8_t* p) { return (p[0] << 8) | p[1]; }

/* Decodes `data` into rgb (interleaved R, G, B; gray images give R = G = B).  rgb may be NULL to only
 * read the size.  Returns 0, -1 (malformed) or -2 (outside the supported subset). */
int orc_jpeg_decode_rgb(const uint8_t* data, size_t len, uint8_t* rgb, int* width, int* height) {
  Huff dc[4], ac[4];
  uint16_t qt[4][64];
  Comp comp[3];
  int ncomp = 0, W = 0, H = 0, restart = 0, have_sof = 0, rc = -1;
  memset(dc, 0, sizeof dc); memset(ac, 0, sizeof ac); memset(comp, 0, sizeof comp); memset(qt, 0, sizeof qt);
  if (len < 4 || data[0] != 0xff || data[1] != 0xd8) return -1;
  size_t pos = 2;
  const uint8_t* scan = NULL;
  while (pos + 4 <= len) {
    if (data[pos] != 0xff) return -1;
    const int m = data[pos + 1];
    if (m == 0xff) { pos++; continue; }
    const int seg = be16(data + pos + 2);
    const uint8_t* s = data + pos + 4;
    if (pos + 2 + (size_t)seg > len) return -1;
    if (m == 0xdb) {                                   /* DQT */
      const uint8_t* e = data + pos + 2 + seg;
      while (s < e) {
        const int pq = s[0] >> 4, tq = s[0] & 15;
        if (tq > 3) return -1;
        s++;
        for (int i = 0; i < 64; i++) {
          qt[tq][kZigzag[i]] = (uint16_t)(pq ? be16(s) : s[0]);
          s += pq ? 2 : 1;
        }
      }
    } else if (m == 0xc4) {                            /* DHT */
      const uint8_t* e = data + pos + 2 + seg;
      while (s < e) {
        const int tc = s[0] >> 4, th = s[0] & 15;
        if (th > 3 || tc > 1) return -1;
        Huff* h = tc ? &ac[th] : &dc[th];
        int total = 0;
        for (int k = 1; k <= 16; k++) { h->bits[k] = s[k]; total += s[k]; }
        if (total > 256) return -1;
        memcpy(h->vals, s + 17, (size_t)total);
        huff_build(h);
        s += 17 + total;
      }
    } else if (m == 0xc0) {                            /* SOF0 */
      if (s[0] != 8) return -2;
      H = be16(s + 1); W = be16(s + 3); ncomp = s[5];
      if ((ncomp != 1 && ncomp != 3) || W <= 0 || H <= 0) return -2;
      for (int i = 0; i < ncomp; i++) {
        comp[i].id = s[6 + 3 * i]; comp[i].h = s[7 + 3 * i] >> 4; comp[i].v = s[7 + 3 * i] & 15;
        comp[i].tq = s[8 + 3 * i];
      }
      have_sof = 1;
    } else if (m >= 0xc1 && m <= 0xcf && m != 0xc4 && m != 0xc8 && m != 0xcc) {
      return -2;                                       /* progressive, extended, arithmetic, ... */
    } else if (m == 0xdd) {                            /* DRI */
      restart = be16(s);
    } else if (m == 0xda) {                            /* SOS */
      if (!have_sof || s[0] != ncomp) return -2;
      for (int i = 0; i < ncomp; i++) {
        if (s[1 + 2 * i] != comp[i].id) return -2;
        comp[i].td = s[2 + 2 * i] >> 4; comp[i].ta = s[2 + 2 * i] & 15;
      }
      scan = data + pos + 2 + seg;
      break;
    }
    pos += 2 + (size_t)seg;
  }
  if (!scan) return -1;
  if (width) *width = W;
  if (height) *height = H;
  if (!rgb) return 0;
  int hmax = comp[0].h, vmax = comp[0].v;
  if (ncomp == 1) { comp[0].h = comp[0].v = hmax = vmax = 1; }        /* a single component is never subsampled */
  if (ncomp == 3 && (comp[1].h != 1 || comp[1].v != 1 || comp[2].h != 1 || comp[2].v != 1)) return -2;
  if (!((hmax == 1 && vmax == 1) || (hmax == 2 && vmax == 1) || (hmax == 2 && vmax == 2))) return -2;
  const int mcux = (W + 8 * hmax - 1) / (8 * hmax), mcuy = (H + 8 * vmax - 1) / (8 * vmax);
  for (int i = 0; i < ncomp; i++) {
    Comp* c = &comp[i];
    c->bw = mcux * c->h; c->bh = mcuy * c->v; c->pw = c->bw * 8; c->ph = c->bh * 8;
    c->dw = (W * c->h + hmax - 1) / hmax; c->dh = (H * c->v + vmax - 1) / vmax;
    c->coef = (int16_t*)calloc((size_t)c->bw * c->bh * 64, sizeof(int16_t));
    c->plane = (uint8_t*)malloc((size_t)c->pw * c->ph);
    if (!dc[c->td].present || !ac[c->ta].present) goto done;
  }
  /* ---- entropy decoding (jdhuff.c) */
  {
    Bits b = {scan, data + len, 0, 0, 0};
    int pred[3] = {0, 0, 0}, todo = restart, next_rst = 0;
    for (int my = 0; my < mcuy; my++)
      for (int mx = 0; mx < mcux; mx++) {
        if (restart && todo == 0) {                     /* expect RSTn: byte-align, skip it, reset predictors */
          const uint8_t* p = b.p;
          while (p + 1 < b.end && !(p[0] == 0xff && p[1] >= 0xd0 && p[1] <= 0xd7)) p++;
          if (p + 1 >= b.end || p[1] != 0xd0 + next_rst) goto done;
          b.p = p + 2; b.acc = 0; b.nbits = 0; b.marker = 0;
          next_rst = (next_rst + 1) & 7;
          pred[0] = pred[1] = pred[2] = 0;
          todo = restart;
        }
        for (int i = 0; i < ncomp; i++) {
          Comp* c = &comp[i];
          for (int by = 0; by < c->v; by++)
            for (int bx = 0; bx < c->h; bx++) {
              int16_t* blk = c->coef + ((size_t)(my * c->v + by) * c->bw + (mx * c->h + bx)) * 64;
              int s = decode_symbol(&b, &dc[c->td]);
              if (s < 0 || s > 11) goto done;
              if (s) pred[i] += extend(get_bits(&b, s), s);
              blk[0] = (int16_t)pred[i];
              for (int k = 1; k < 64; k++) {
                const int rs = decode_symbol(&b, &ac[c->ta]);
                if (rs < 0) goto done;
                const int r = rs >> 4, sz = rs & 15;
                if (sz) {
                  k += r;
                  if (k > 63) goto done;
                  blk[kZigzag[k]] = (int16_t)extend(get_bits(&b, sz), sz);
                } else if (r == 15) {
                  k += 15;
                } else {
                  break;
                }
              }
            }
        }
        todo--;
      }
  }
  /* ---- inverse DCT into the component planes */
  for (int i = 0; i < ncomp; i++) {
    Comp* c = &comp[i];
    for (int by = 0; by < c->bh; by++)
      for (int bx = 0; bx < c->bw; bx++)
        idct_block(c->coef + ((size_t)by * c->bw + bx) * 64, qt[c->tq], c->plane + (size_t)by * 8 * c->pw + bx * 8, c->pw);
  }
  /* ---- upsampling + colour conversion (jdcolor.c, SCALEBITS 16) */
  {
    uint8_t* full[3] = {NULL, NULL, NULL};
    for (int i = 0; i < ncomp; i++) {
      full[i] = (uint8_t*)malloc((size_t)W * H);
      upsample(&comp[i], hmax, vmax, W, H, full[i]);
    }
    for (size_t p = 0; p < (size_t)W * H; p++) {
      const int y = full[0][p];
      if (ncomp == 1) { rgb[3 * p] = rgb[3 * p + 1] = rgb[3 * p + 2] = (uint8_t)y; continue; }
      const int cb = full[1][p] - 128, cr = full[2][p] - 128;
      int r = y + ((91881 * cr + 32768) >> 16);
      int g = y + ((-22554 * cb + 32768 - 46802 * cr) >> 16);
      int bl = y + ((116130 * cb + 32768) >> 16);
      rgb[3 * p] = (uint8_t)(r < 0 ? 0 : (r > 255 ? 255 : r));
      rgb[3 * p + 1] = (uint8_t)(g < 0 ? 0 : (g > 255 ? 255 : g));
      rgb[3 * p + 2] = (uint8_t)(bl < 0 ? 0 : (bl > 255 ? 255 : bl));
    }
    for (int i = 0; i < ncomp; i++) free(full[i]);
  }
  rc = 0;
done:
  for (int i = 0; i < ncomp; i++) { free(comp[i].coef); free(comp[i].plane); }
  return rc;
}
