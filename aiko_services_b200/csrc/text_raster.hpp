// Host-side restatement of the rasteriser behind
//   cv2.putText(img, text, org, FONT_HERSHEY_SIMPLEX, 0.75, colour, 2, cv2.LINE_AA)
// (the call ImageOverlay makes, R/elements/media/image_io.py:151-153).  OpenCV is not vendored by
// the reference; what follows is the published algorithm of its drawing module as the installed
// library (4.13) executes it, pinned bit for bit against cv2 in tests/test_text_raster.py:
//   putText    : glyph polylines in 16.16 fixed point, one ThickLine per segment
//   ThickLine  : (thickness 2) the segment's 4-point rectangle, then a 4-point "circle" of radius
//                one pixel on its end points, all through FillConvexPoly(LINE_AA)
//   FillConvexPoly : anti-aliased edges with LineAA, then solid scan-line fill
//   LineAA     : clip to the image, step along the major axis, three pixels per step weighted by a
//                Gaussian-like filter table, slope and end-point corrections; every touched pixel
//                is blended TWICE with  d += ((c - d) * a + 127) >> 8
// Nothing here touches pixels: the rasteriser reports, per pixel, the ORDER of "blend with a" and
// "store the colour" events to a sink.  The planner composes them into 256-entry transfer tables
// that the overlay kernel replays on the GPU (include/aiko_b200.h, glyph atlas).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <vector>

namespace aiko_text {

constexpr int XY_SHIFT = 16;
constexpr int64_t XY_ONE = 1 << XY_SHIFT;

struct P2 {
  int64_t x, y;
};

static const uint8_t kSlopeCorr[32] = {181, 181, 181, 182, 182, 183, 184, 185, 187, 188, 190, 192, 194, 196, 198, 201,
                                       203, 206, 209, 211, 214, 218, 221, 224, 227, 231, 235, 238, 242, 246, 250, 254};
static const int kFilter[64] = {168, 177, 185, 194, 202, 210, 218, 224, 231, 236, 241, 246, 249, 252, 254, 254,
                                254, 254, 252, 249, 246, 241, 236, 231, 224, 218, 210, 202, 194, 185, 177, 168,
                                158, 149, 140, 131, 122, 114, 105, 97,  89,  82,  75,  68,  62,  56,  50,  45,
                                40,  36,  32,  28,  25,  22,  19,  16,  14,  12,  11,  9,   8,   7,   5,   5};

static const char* const kHersheySimplex[95] = {
#include "hershey_simplex.inc"
};

inline int cv_round(double v) { return (int)std::nearbyint(v); }   // round half to even, as cvRound

// cv::clipLine on 64-bit points, image = [0, w) x [0, h)
inline bool clip_line(int64_t w, int64_t h, P2& p1, P2& p2) {
  int64_t right = w - 1, bottom = h - 1;
  if (w <= 0 || h <= 0) return false;
  int64_t &x1 = p1.x, &y1 = p1.y, &x2 = p2.x, &y2 = p2.y;
  int c1 = (x1 < 0) + (x1 > right) * 2 + (y1 < 0) * 4 + (y1 > bottom) * 8;
  int c2 = (x2 < 0) + (x2 > right) * 2 + (y2 < 0) * 4 + (y2 > bottom) * 8;
  if ((c1 & c2) == 0 && (c1 | c2) != 0) {
    int64_t a;
    if (c1 & 12) {
      a = c1 < 8 ? 0 : bottom;
      x1 += (int64_t)((double)(a - y1) * (x2 - x1) / (y2 - y1));
      y1 = a;
      c1 = (x1 < 0) + (x1 > right) * 2;
    }
    if (c2 & 12) {
      a = c2 < 8 ? 0 : bottom;
      x2 += (int64_t)((double)(a - y2) * (x2 - x1) / (y2 - y1));
      y2 = a;
      c2 = (x2 < 0) + (x2 > right) * 2;
    }
    if ((c1 & c2) == 0 && (c1 | c2) != 0) {
      if (c1) {
        a = c1 == 1 ? 0 : right;
        y1 += (int64_t)((double)(a - x1) * (y2 - y1) / (x2 - x1));
        x1 = a;
        c1 = 0;
      }
      if (c2) {
        a = c2 == 1 ? 0 : right;
        y2 += (int64_t)((double)(a - x2) * (y2 - y1) / (x2 - x1));
        x2 = a;
        c2 = 0;
      }
    }
  }
  return (c1 | c2) == 0;
}

// Sink concept:  int width() / height()  (image size: the clip rectangle)
//                void blend(int x, int y, int a)   d += ((c-d)*a+127)>>8, twice
//                void fill(int y, int x0, int x1)  store the colour on [x0, x1]
template <class Sink>
void line_aa(Sink& s, P2 pt1, P2 pt2) {
  const int W = s.width(), H = s.height();
  if (!clip_line((int64_t)W << XY_SHIFT, (int64_t)H << XY_SHIFT, pt1, pt2)) return;
  int64_t dx = pt2.x - pt1.x, dy = pt2.y - pt1.y;
  int64_t j = dx < 0 ? -1 : 0;
  int64_t ax = (dx ^ j) - j;
  int64_t i = dy < 0 ? -1 : 0;
  int64_t ay = (dy ^ i) - i;
  int64_t x_step, y_step;
  int ecount, scount = 0, slope;
  if (ax > ay) {
    dy = (dy ^ j) - j;
    pt1.x ^= pt2.x & j; pt2.x ^= pt1.x & j; pt1.x ^= pt2.x & j;
    pt1.y ^= pt2.y & j; pt2.y ^= pt1.y & j; pt1.y ^= pt2.y & j;
    x_step = XY_ONE;
    y_step = (dy * XY_ONE) / (ax | 1);
    pt2.x += XY_ONE;
    ecount = (int)((pt2.x >> XY_SHIFT) - (pt1.x >> XY_SHIFT));
    j = -(pt1.x & (XY_ONE - 1));
    pt1.y += ((y_step * j) >> XY_SHIFT) + (XY_ONE >> 1);
    slope = (int)((y_step >> (XY_SHIFT - 5)) & 0x3f);
    slope ^= (y_step < 0 ? 0x3f : 0);
    i = (pt1.x >> (XY_SHIFT - 7)) & 0x78;
    j = (pt2.x >> (XY_SHIFT - 7)) & 0x78;
  } else {
    dx = (dx ^ i) - i;
    pt1.x ^= pt2.x & i; pt2.x ^= pt1.x & i; pt1.x ^= pt2.x & i;
    pt1.y ^= pt2.y & i; pt2.y ^= pt1.y & i; pt1.y ^= pt2.y & i;
    x_step = (dx * XY_ONE) / (ay | 1);
    y_step = XY_ONE;
    pt2.y += XY_ONE;
    ecount = (int)((pt2.y >> XY_SHIFT) - (pt1.y >> XY_SHIFT));
    j = -(pt1.y & (XY_ONE - 1));
    pt1.x += ((x_step * j) >> XY_SHIFT) + (XY_ONE >> 1);
    slope = (int)((x_step >> (XY_SHIFT - 5)) & 0x3f);
    slope ^= (x_step < 0 ? 0x3f : 0);
    i = (pt1.y >> (XY_SHIFT - 7)) & 0x78;
    j = (pt2.y >> (XY_SHIFT - 7)) & 0x78;
  }
  slope = (slope & 0x20) ? 0x100 : kSlopeCorr[slope];
  int ep[9];
  {
    int t0 = slope << 7;
    int t1 = ((0x78 - (int)i) | 4) * slope;
    int t2 = ((int)j | 4) * slope;
    ep[0] = 0;
    ep[8] = slope;
    ep[1] = ep[3] = (int)((((((j - i) & 0x78) | 4) * slope) >> 8) & 0x1ff);
    ep[2] = (t1 >> 8) & 0x1ff;
    ep[4] = (int)((((((j - i) + 0x80) | 4) * slope) >> 8) & 0x1ff);
    ep[5] = ((t1 + t0) >> 8) & 0x1ff;
    ep[6] = (t2 >> 8) & 0x1ff;
    ep[7] = ((t2 + t0) >> 8) & 0x1ff;
  }
  if (ax > ay) {
    int x = (int)(pt1.x >> XY_SHIFT);
    for (; ecount >= 0; x++, pt1.y += y_step, scount++, ecount--) {
      if ((unsigned)x >= (unsigned)W) continue;
      int y = (int)((pt1.y >> XY_SHIFT) - 1);
      int ep_corr = ep[(((scount >= 2) + 1) & (scount | 2)) * 3 + (((ecount >= 2) + 1) & (ecount | 2))];
      int dist = (int)((pt1.y >> (XY_SHIFT - 5)) & 31);
      int a = (ep_corr * kFilter[dist + 32] >> 8) & 0xff;
      if ((unsigned)y < (unsigned)H) s.blend(x, y, a);
      a = (ep_corr * kFilter[dist] >> 8) & 0xff;
      if ((unsigned)(y + 1) < (unsigned)H) s.blend(x, y + 1, a);
      a = (ep_corr * kFilter[63 - dist] >> 8) & 0xff;
      if ((unsigned)(y + 2) < (unsigned)H) s.blend(x, y + 2, a);
    }
  } else {
    int y = (int)(pt1.y >> XY_SHIFT);
    for (; ecount >= 0; y++, pt1.x += x_step, scount++, ecount--) {
      if ((unsigned)y >= (unsigned)H) continue;
      int x = (int)((pt1.x >> XY_SHIFT) - 1);
      int ep_corr = ep[(((scount >= 2) + 1) & (scount | 2)) * 3 + (((ecount >= 2) + 1) & (ecount | 2))];
      int dist = (int)((pt1.x >> (XY_SHIFT - 5)) & 31);
      int a = (ep_corr * kFilter[dist + 32] >> 8) & 0xff;
      if ((unsigned)x < (unsigned)W) s.blend(x, y, a);
      a = (ep_corr * kFilter[dist] >> 8) & 0xff;
      if ((unsigned)(x + 1) < (unsigned)W) s.blend(x + 1, y, a);
      a = (ep_corr * kFilter[63 - dist] >> 8) & 0xff;
      if ((unsigned)(x + 2) < (unsigned)W) s.blend(x + 2, y, a);
    }
  }
}

// cv::FillConvexPoly(img, v, npts, colour, LINE_AA, XY_SHIFT)
template <class Sink>
void fill_convex_poly_aa(Sink& s, const P2* v, int npts) {
  struct Edge {
    int idx, di;
    int64_t x, dx;
    int ye;
  } edge[2];
  const int shift = XY_SHIFT;
  const int delta = 1 << shift >> 1;
  const int W = s.width(), H = s.height();
  int imin = 0;
  int edges = npts;
  int64_t xmin, xmax, ymin, ymax;
  const int delta1 = (int)XY_ONE - 1, delta2 = 0;       // LINE_AA
  P2 p0 = v[npts - 1];
  xmin = xmax = v[0].x;
  ymin = ymax = v[0].y;
  for (int i = 0; i < npts; i++) {
    P2 p = v[i];
    if (p.y < ymin) {
      ymin = p.y;
      imin = i;
    }
    ymax = p.y > ymax ? p.y : ymax;
    xmax = p.x > xmax ? p.x : xmax;
    xmin = p.x < xmin ? p.x : xmin;
    line_aa(s, p0, p);
    p0 = p;
  }
  xmin = (xmin + delta) >> shift;
  xmax = (xmax + delta) >> shift;
  ymin = (ymin + delta) >> shift;
  ymax = (ymax + delta) >> shift;
  if (npts < 3 || (int)xmax < 0 || (int)ymax < 0 || (int)xmin >= W || (int)ymin >= H) return;
  ymax = ymax < H - 1 ? ymax : H - 1;
  edge[0].idx = edge[1].idx = imin;
  int y;
  edge[0].ye = edge[1].ye = y = (int)ymin;
  edge[0].di = 1;
  edge[1].di = npts - 1;
  edge[0].x = edge[1].x = -XY_ONE;
  edge[0].dx = edge[1].dx = 0;
  do {
    if (y < (int)ymax || y == (int)ymin) {
      for (int i = 0; i < 2; i++) {
        if (y >= edge[i].ye) {
          int idx0 = edge[i].idx, di = edge[i].di;
          int idx = idx0 + di;
          if (idx >= npts) idx -= npts;
          int ty = 0;
          for (; edges-- > 0;) {
            ty = (int)((v[idx].y + delta) >> shift);
            if (ty > y) {
              int64_t xs = v[idx0].x;
              int64_t xe = v[idx].x;
              edge[i].ye = ty;
              edge[i].dx = ((xe - xs) * 2 + (ty - y)) / (2 * (ty - y));
              edge[i].x = xs;
              edge[i].idx = idx;
              break;
            }
            idx0 = idx;
            idx += di;
            if (idx >= npts) idx -= npts;
          }
        }
      }
    }
    if (edges < 0) break;
    if (y >= 0) {
      int left = 0, right = 1;
      if (edge[0].x > edge[1].x) {
        left = 1;
        right = 0;
      }
      int xx1 = (int)((edge[left].x + delta1) >> XY_SHIFT);
      int xx2 = (int)((edge[right].x + delta2) >> XY_SHIFT);
      if (xx2 >= 0 && xx1 < W) {
        if (xx1 < 0) xx1 = 0;
        if (xx2 >= W) xx2 = W - 1;
        s.fill(y, xx1, xx2);
      }
    }
    edge[0].x += edge[0].dx;
    edge[1].x += edge[1].dx;
  } while (++y <= (int)ymax);
}

// cv::ThickLine(img, p0, p1, colour, thickness = 2, LINE_AA, flags, XY_SHIFT)
template <class Sink>
void thick_line2(Sink& s, P2 p0, P2 p1, int flags) {
  const double INV_XY_ONE = 1. / XY_ONE;
  const int64_t thickness = 2 << (XY_SHIFT - 1);
  P2 pt[4];
  double dx = (p0.x - p1.x) * INV_XY_ONE, dy = (p1.y - p0.y) * INV_XY_ONE;
  double r = dx * dx + dy * dy;
  if (std::fabs(r) > 2.220446049250313e-16) {
    r = (double)thickness / std::sqrt(r);
    int64_t dpx = cv_round(dy * r), dpy = cv_round(dx * r);
    pt[0] = {p0.x + dpx, p0.y + dpy};
    pt[1] = {p0.x - dpx, p0.y - dpy};
    pt[2] = {p1.x - dpx, p1.y - dpy};
    pt[3] = {p1.x + dpx, p1.y + dpy};
    fill_convex_poly_aa(s, pt, 4);
  }
  for (int i = 0; i < 2; i++) {
    if (flags & (i + 1)) {
      // EllipseEx(center p0, axes (thickness, thickness), 0..360): radius 1 px -> 90-degree steps
      P2 c[5] = {{p0.x + thickness, p0.y}, {p0.x, p0.y + thickness}, {p0.x - thickness, p0.y},
                 {p0.x, p0.y - thickness}, {p0.x + thickness, p0.y}};
      fill_convex_poly_aa(s, c, 5);
    }
    p0 = p1;
  }
}

inline int glyph_index(unsigned char c) { return (c < 32 || c > 126) ? '?' - 32 : c - 32; }

// Advance of a character in glyph units (pixels = units * font scale).
inline int glyph_advance(unsigned char c) {
  const char* g = kHersheySimplex[glyph_index(c)];
  return (int)(unsigned char)g[1] - (int)(unsigned char)g[0];
}

// The strokes of ONE character with the pen at view_x (16.16) and baseline row org_y:
// what cv::putText does for that character.  hscale = cvRound(font_scale * 65536).
template <class Sink>
void draw_glyph(Sink& s, unsigned char c, int64_t view_x, int org_y, int hscale) {
  const char* ptr = kHersheySimplex[glyph_index(c)];
  const int64_t view_y = ((int64_t)org_y << XY_SHIFT) - 9 * (int64_t)hscale;     // base line of the simplex face
  view_x -= ((int)(unsigned char)ptr[0] - 'R') * (int64_t)hscale;
  std::vector<P2> pts;
  for (ptr += 2;;) {
    if (*ptr == ' ' || !*ptr) {
      if (pts.size() > 1) {
        int flags = 3;
        for (size_t k = 1; k < pts.size(); k++) {
          thick_line2(s, pts[k - 1], pts[k], flags);
          flags = 2;
        }
      }
      if (!*ptr++) break;
      pts.clear();
    } else {
      int px = (int)(unsigned char)ptr[0] - 'R', py = (int)(unsigned char)ptr[1] - 'R';
      ptr += 2;
      pts.push_back({px * (int64_t)hscale + view_x, py * (int64_t)hscale + view_y});
    }
  }
}

}  // namespace aiko_text
