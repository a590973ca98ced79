/*
 * oracle/image_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, single-threaded CPU restatement of the per-pixel arithmetic behind
 * the reference's image PipelineElements.  The reference (pure Python) delegates
 * that arithmetic to un-vendored third-party libraries, so each function below
 * restates the PUBLISHED algorithm of the library call the reference makes, and
 * cites the reference call site it stands for:
 *
 *   orc_resize_linear_u8      cv2.resize(image, (W, H))  [INTER_LINEAR]
 *                             R/elements/media/image_io.py:216
 *   orc_resize_bicubic_aa_u8  PIL.Image.resize((W, H))   [BICUBIC, antialias]
 *                             R/elements/media/image_io.py:218
 *   orc_swap_rb_u8            cv2.cvtColor(RGB2BGR / BGR2RGB)
 *                             R/elements/media/image_io.py:117,160
 *                             R/elements/media/video_io.py:129,299
 *                             R/elements/media/webcam_io.py:122
 *   orc_bgr2gray_u8           cv2.cvtColor(BGR2GRAY)
 *                             R/elements/media/image_io.py:158, webcam_io.py:124
 *   orc_gray2bgr_u8           cv2.cvtColor(GRAY2BGR)   image_io.py:115
 *   orc_flip_u8               cv2.flip                 webcam_io.py:127,129
 *   orc_rectangle_u8          cv2.rectangle(..., thickness=2)  image_io.py:154-155
 *   orc_text_lut_u8           cv2.putText(..., LINE_AA) replayed from per-pixel
 *                             transfer tables          image_io.py:151-153
 *   orc_resize_norm_f32       extension (BASELINE config 4): resize_linear then
 *                             (x/255 - mean)/std, HWC -> CHW
 *
 * (R/ = /root/reference/src/aiko_services/.)
 *
 * Parity pinning: the reference ships no golden vectors or tests for this path
 * (SURVEY.md section 4), so this restatement is pinned against outputs of the
 * reference's own elements run in the build container (tests/golden/, made by
 * tests/golden/make_golden.py) and, where cv2 / Pillow are importable, directly
 * against those libraries in tests/test_oracle_vs_libs.py.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------ */
/* cv2.resize INTER_LINEAR, 8-bit.  Fixed point: 11-bit coefficients.        */

static inline int orc_round_half_even_f(float v) { return (int)lrintf(v); }

static void linear_coeffs(int src, int dst, int vertical, int *ofs, short *c0, short *c1) {
  double inv_scale = (double)dst / (double)src;
  double scale = 1.0 / inv_scale;
  for (int d = 0; d < dst; d++) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = (int)floorf(f);
    f -= (float)s;
    if (!vertical) {
      if (s < 0) { s = 0; f = 0.f; }
      if (s >= src - 1) { s = src - 1; f = 0.f; }
    }
    ofs[d] = s;
    c0[d] = (short)orc_round_half_even_f((1.f - f) * 2048.f);
    c1[d] = (short)orc_round_half_even_f(f * 2048.f);
  }
}

static inline int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

/* src: [sh, sw, ch] u8 tight; dst: [dh, dw, ch] u8 tight. returns 0 / -1 */
int orc_resize_linear_u8(const uint8_t *src, int sh, int sw, int ch, uint8_t *dst, int dh, int dw) {
  if (sh <= 0 || sw <= 0 || dh <= 0 || dw <= 0 || ch <= 0) return -1;
  int *xo = (int *)malloc(sizeof(int) * dw), *yo = (int *)malloc(sizeof(int) * dh);
  short *a0 = (short *)malloc(2 * dw), *a1 = (short *)malloc(2 * dw);
  short *b0 = (short *)malloc(2 * dh), *b1 = (short *)malloc(2 * dh);
  int32_t *r0 = (int32_t *)malloc(sizeof(int32_t) * dw * ch);
  int32_t *r1 = (int32_t *)malloc(sizeof(int32_t) * dw * ch);
  linear_coeffs(sw, dw, 0, xo, a0, a1);
  linear_coeffs(sh, dh, 1, yo, b0, b1);
  for (int y = 0; y < dh; y++) {
    int y0 = clampi(yo[y], 0, sh - 1), y1 = clampi(yo[y] + 1, 0, sh - 1);
    const uint8_t *p0 = src + (size_t)y0 * sw * ch, *p1 = src + (size_t)y1 * sw * ch;
    for (int x = 0; x < dw; x++) {
      int s0 = xo[x], s1 = s0 + 1 < sw ? s0 + 1 : sw - 1;
      for (int c = 0; c < ch; c++) {
        r0[x * ch + c] = p0[s0 * ch + c] * a0[x] + p0[s1 * ch + c] * a1[x];
        r1[x * ch + c] = p1[s0 * ch + c] * a0[x] + p1[s1 * ch + c] * a1[x];
      }
    }
    uint8_t *o = dst + (size_t)y * dw * ch;
    for (int i = 0; i < dw * ch; i++) {
      int v = (((b0[y] * (r0[i] >> 4)) >> 16) + ((b1[y] * (r1[i] >> 4)) >> 16) + 2) >> 2;
      o[i] = (uint8_t)clampi(v, 0, 255);
    }
  }
  free(xo); free(yo); free(a0); free(a1); free(b0); free(b1); free(r0); free(r1);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* PIL.Image.resize default (BICUBIC, a = -0.5, antialiased support),        */
/* 8 bits per channel: horizontal pass, uint8 intermediate, vertical pass.   */

#define ORC_PRECISION_BITS (32 - 8 - 2)

static double bicubic_kernel(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}

/* returns ksize; bounds[2*i] = first tap, bounds[2*i+1] = tap count; kk[i*ksize + t] int */
static int bicubic_coeffs(int in_size, int out_size, int **bounds_out, int32_t **kk_out) {
  double scale = (double)in_size / out_size, filterscale = scale;
  if (filterscale < 1.0) filterscale = 1.0;
  double support = 2.0 * filterscale;
  int ksize = (int)ceil(support) * 2 + 1;
  int *bounds = (int *)malloc(sizeof(int) * 2 * out_size);
  int32_t *kk = (int32_t *)calloc((size_t)out_size * ksize, sizeof(int32_t));
  double *k = (double *)malloc(sizeof(double) * ksize);
  for (int xx = 0; xx < out_size; xx++) {
    double center = 0.0 + (xx + 0.5) * scale, ww = 0.0, ss = 1.0 / filterscale;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    for (int x = 0; x < xmax; x++) {
      double w = bicubic_kernel((x + xmin - center + 0.5) * ss);
      k[x] = w;
      ww += w;
    }
    for (int x = 0; x < xmax; x++) {
      if (ww != 0.0) k[x] /= ww;
      double v = k[x] * (double)(1 << ORC_PRECISION_BITS);
      kk[(size_t)xx * ksize + x] = v < 0 ? (int32_t)(-0.5 + v) : (int32_t)(0.5 + v);
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
  }
  free(k);
  *bounds_out = bounds;
  *kk_out = kk;
  return ksize;
}

static inline uint8_t clip8_fixed(int32_t v) {
  v >>= ORC_PRECISION_BITS;
  return (uint8_t)(v < 0 ? 0 : (v > 255 ? 255 : v));
}

int orc_resize_bicubic_aa_u8(const uint8_t *src, int sh, int sw, int ch, uint8_t *dst, int dh, int dw) {
  if (sh <= 0 || sw <= 0 || dh <= 0 || dw <= 0 || ch <= 0) return -1;
  const uint8_t *hsrc = src;
  uint8_t *tmp = NULL;
  if (dw != sw) { /* horizontal pass -> tmp [sh, dw, ch] */
    int *bounds; int32_t *kk;
    int ksize = bicubic_coeffs(sw, dw, &bounds, &kk);
    tmp = (uint8_t *)malloc((size_t)sh * dw * ch);
    for (int y = 0; y < sh; y++)
      for (int x = 0; x < dw; x++) {
        int xmin = bounds[2 * x], n = bounds[2 * x + 1];
        const int32_t *k = kk + (size_t)x * ksize;
        for (int c = 0; c < ch; c++) {
          int32_t ss = 1 << (ORC_PRECISION_BITS - 1);
          for (int t = 0; t < n; t++) ss += src[((size_t)y * sw + xmin + t) * ch + c] * k[t];
          tmp[((size_t)y * dw + x) * ch + c] = clip8_fixed(ss);
        }
      }
    free(bounds); free(kk);
    hsrc = tmp;
  }
  if (dh != sh) {
    int *bounds; int32_t *kk;
    int ksize = bicubic_coeffs(sh, dh, &bounds, &kk);
    for (int y = 0; y < dh; y++) {
      int ymin = bounds[2 * y], n = bounds[2 * y + 1];
      const int32_t *k = kk + (size_t)y * ksize;
      for (int i = 0; i < dw * ch; i++) {
        int32_t ss = 1 << (ORC_PRECISION_BITS - 1);
        for (int t = 0; t < n; t++) ss += hsrc[(size_t)(ymin + t) * dw * ch + i] * k[t];
        dst[(size_t)y * dw * ch + i] = clip8_fixed(ss);
      }
    }
    free(bounds); free(kk);
  } else {
    memcpy(dst, hsrc, (size_t)dh * dw * ch);
  }
  free(tmp);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* Colour conversions and flips (cv2.cvtColor / cv2.flip), 8-bit.            */

int orc_swap_rb_u8(const uint8_t *src, uint8_t *dst, size_t pixels) {
  for (size_t i = 0; i < pixels; i++) {
    uint8_t a = src[3 * i], b = src[3 * i + 1], c = src[3 * i + 2];
    dst[3 * i] = c; dst[3 * i + 1] = b; dst[3 * i + 2] = a;
  }
  return 0;
}

/* 15-bit fixed point: Y = (B*3735 + G*19235 + R*9798 + 16384) >> 15.
 * rgb_order = 0: input is B,G,R; 1: input is R,G,B. */
int orc_bgr2gray_u8(const uint8_t *src, uint8_t *dst, size_t pixels, int rgb_order) {
  for (size_t i = 0; i < pixels; i++) {
    int b = src[3 * i + (rgb_order ? 2 : 0)], g = src[3 * i + 1], r = src[3 * i + (rgb_order ? 0 : 2)];
    dst[i] = (uint8_t)((b * 3735 + g * 19235 + r * 9798 + 16384) >> 15);
  }
  return 0;
}

int orc_gray2bgr_u8(const uint8_t *src, uint8_t *dst, size_t pixels) {
  for (size_t i = 0; i < pixels; i++) dst[3 * i] = dst[3 * i + 1] = dst[3 * i + 2] = src[i];
  return 0;
}

/* flip_code as cv2.flip: 0 = around x axis (vertical), >0 = around y axis
 * (horizontal), <0 = both. */
int orc_flip_u8(const uint8_t *src, uint8_t *dst, int h, int w, int ch, int flip_code) {
  for (int y = 0; y < h; y++) {
    int sy = (flip_code <= 0) ? h - 1 - y : y;
    for (int x = 0; x < w; x++) {
      int sx = (flip_code != 0) ? w - 1 - x : x;
      memcpy(dst + ((size_t)y * w + x) * ch, src + ((size_t)sy * w + sx) * ch, ch);
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------ */
/* cv2.rectangle(img, (x0,y0), (x1,y1), colour, thickness=2), 8-connected.   */
/* Closed form: the outline covers                        */
/*   [xa-1 .. xb+1] x [ya-1 .. yb+1]  minus  [xa+2 .. xb-2] x [ya+2 .. yb-2] */
/* minus the four extreme corner pixels, clipped to the image, where         */
/* xa=min(x0,x1) etc.                                                        */

int orc_rect_covers(int x, int y, int x0, int y0, int x1, int y1) {
  int xa = x0 < x1 ? x0 : x1, xb = x0 < x1 ? x1 : x0;
  int ya = y0 < y1 ? y0 : y1, yb = y0 < y1 ? y1 : y0;
  if (x < xa - 1 || x > xb + 1 || y < ya - 1 || y > yb + 1) return 0;
  if (x >= xa + 2 && x <= xb - 2 && y >= ya + 2 && y <= yb - 2) return 0;
  if ((x == xa - 1 || x == xb + 1) && (y == ya - 1 || y == yb + 1)) return 0;
  return 1;
}

int orc_rectangle_u8(uint8_t *img, int h, int w, int ch, int x0, int y0, int x1, int y1,
                     const uint8_t *colour) {
  int xa = x0 < x1 ? x0 : x1, xb = x0 < x1 ? x1 : x0;
  int ya = y0 < y1 ? y0 : y1, yb = y0 < y1 ? y1 : y0;
  for (int y = clampi(ya - 1, 0, h); y <= yb + 1 && y < h; y++)
    for (int x = clampi(xa - 1, 0, w); x <= xb + 1 && x < w; x++)
      if (orc_rect_covers(x, y, x0, y0, x1, y1)) memcpy(img + ((size_t)y * w + x) * ch, colour, ch);
  return 0;
}

/* ------------------------------------------------------------------------ */
/* Anti-aliased text replay.  A glyph patch is a [ph, pw] map of uint16      */
/* transfer-function ids (0 = untouched) into luts[id][256]; one patch per   */
/* colour component value.  The patch is stamped with its top-left at        */
/* (px, py); pixels outside the image are skipped.  comp_patch[c] selects    */
/* which patch/lut pair channel c uses.                                      */

int orc_text_lut_u8(uint8_t *img, int h, int w, int ch, int px, int py, int ph, int pw,
                    const uint16_t *const *patch_per_ch, const uint8_t *const *luts_per_ch) {
  for (int y = 0; y < ph; y++) {
    int iy = py + y;
    if (iy < 0 || iy >= h) continue;
    for (int x = 0; x < pw; x++) {
      int ix = px + x;
      if (ix < 0 || ix >= w) continue;
      for (int c = 0; c < ch; c++) {
        uint16_t id = patch_per_ch[c][(size_t)y * pw + x];
        if (id) {
          uint8_t *p = img + ((size_t)iy * w + ix) * ch + c;
          *p = luts_per_ch[c][(size_t)id * 256 + *p];
        }
      }
    }
  }
  return 0;
}

/* ------------------------------------------------------------------------ */
/* Extension (BASELINE config 4): bilinear resize then normalise to fp32 CHW */
/* dst[c][y][x] = ((float)u8 / 255.f - mean[c]) / std[c]  (numpy float32 ops)*/

int orc_resize_norm_f32(const uint8_t *src, int sh, int sw, float *dst, int dh, int dw,
                        const float *mean, const float *stdv) {
  uint8_t *tmp = (uint8_t *)malloc((size_t)dh * dw * 3);
  int rc = orc_resize_linear_u8(src, sh, sw, 3, tmp, dh, dw);
  if (rc == 0)
    for (int c = 0; c < 3; c++)
      for (size_t i = 0; i < (size_t)dh * dw; i++) {
        volatile float v = (float)tmp[3 * i + c] / 255.f;
        volatile float u = v - mean[c];
        dst[(size_t)c * dh * dw + i] = u / stdv[c];
      }
  free(tmp);
  return rc;
}
