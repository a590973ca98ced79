// Stand-alone ImageOverlay on device-resident frames (R/elements/media/image_io.py:107-163).
// One CTA per image row: the row is staged in shared memory as 3-channel pixels, the
// frame's draw list is replayed on it in order, and the row is written back.  Grayscale
// frames take the reference's GRAY->BGR, draw, BGR->GRAY round trip (image_io.py:115,158).
#include <stdlib.h>
#include <string.h>

#include "overlay.cuh"
#include "overlay_marks.h"

namespace aiko {

static_assert(sizeof(aiko_b200_overlay) == 96 && sizeof(aiko_b200_overlay_op) == 32, "ABI layout (tests/test_abi.py)");

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
    const int li = overlay_list(p.ov, n);
    const int ob = p.ov.op_offsets[li];
    overlay_apply_row(p.ov.ops + ob, p.ov.op_offsets[li + 1] - ob, p.ov, y, p.w, row, kThreads);
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

// ---------------------------------------------------------------------------------------------
// Tile-parallel overlay, in place.  Used right after the resize kernel wrote the frames, while
// they are still in L2: one CTA per 64 x 16 pixel tile that the frame's draw list touches (the
// host lists those tiles with the ops that can touch each).  Every thread owns 4 pixels, holds
// them in registers, applies the tile's ops to them in list order and writes them back once;
// ownership makes barriers between ops unnecessary.  Without a tile list from the host the
// kernel runs over every tile of every frame and culls the draw list itself.
constexpr int kTileW = 64, kTileH = 16;     // 16 groups of 4 pixels x 16 rows = 256 threads

struct TileParams {
  uint8_t* frames;
  int n, h, w;
  long long row_stride, frame_stride;
  int tiles_x, tiles_y;
  int cells_per_cta;       // overlay_cells_kernel: consecutive cells of one frame handled by a CTA
  OverlayDev ov;
};

struct Pixels4 {         // 12 bytes = 4 RGB pixels of one thread
  uint32_t w0, w1, w2;
};

__device__ __forceinline__ Pixels4 load_pixels(const uint8_t* pix, bool words, int count) {
  Pixels4 r{0, 0, 0};
  if (words) {
    r.w0 = reinterpret_cast<const uint32_t*>(pix)[0];
    r.w1 = reinterpret_cast<const uint32_t*>(pix)[1];
    r.w2 = reinterpret_cast<const uint32_t*>(pix)[2];
  } else {
    uint32_t v[3] = {0, 0, 0};
    for (int q = 0; q < count * 3; q++) v[q >> 2] |= (uint32_t)pix[q] << (8 * (q & 3));
    r.w0 = v[0]; r.w1 = v[1]; r.w2 = v[2];
  }
  return r;
}

// Applies one op to the thread's pixels mlo .. mhi of row y.  Returns true if anything changed.
__device__ __forceinline__ bool apply_op_regs(const int4 lo4, const int4 hi4, const OverlayDev& ov, int psize, int y,
                                              int mlo, int mhi, Pixels4& px) {
  const int kind = lo4.x, x0 = lo4.y, y0 = lo4.z, x1 = lo4.w, y1 = hi4.x;
  const int grow = (kind == AIKO_B200_OP_RECT) ? 1 : 0;
  if (y < y0 - grow || y > y1 + grow || mhi < x0 - grow || mlo > x1 + grow) return false;
  if (kind == AIKO_B200_OP_RECT) {
    bool cover[4], any = false;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      cover[k] = (mlo + k <= mhi) && rect_covers(mlo + k, y, x0, y0, x1, y1);
      any |= cover[k];
    }
    if (!any) return false;
    const uint32_t c = (uint32_t)hi4.y & 0xffffffu;     // colour bytes 0..2; pixel k = bytes 3k .. 3k+2
    if (cover[0]) px.w0 = (px.w0 & 0xff000000u) | c;
    if (cover[1]) { px.w0 = (px.w0 & 0x00ffffffu) | (c << 24); px.w1 = (px.w1 & 0xffff0000u) | (c >> 8); }
    if (cover[2]) { px.w1 = (px.w1 & 0x0000ffffu) | (c << 16); px.w2 = (px.w2 & 0xffffff00u) | (c >> 16); }
    if (cover[3]) px.w2 = (px.w2 & 0x000000ffu) | (c << 8);
    return true;
  }
  const int xs = max(mlo, x0), xe = min(mhi, x1);
  const uint32_t p0 = (uint32_t)hi4.y, p1 = (uint32_t)hi4.z, p2 = (uint32_t)hi4.w;
  const int opx = x0 - int(p2 & 0xffff), opy = y0 - int(p2 >> 16);
  const uint16_t* q = ov.patch_ids + (y - opy) * ov.patch_w - opx;
  const uint16_t* qc[3] = {q + (size_t)(p0 & 0xffff) * psize, q + (size_t)(p0 >> 16) * psize,
                           q + (size_t)(p1 & 0xffff) * psize};
  uint16_t id[4][3];
  bool any = false;
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const int x = mlo + k;
      id[k][c] = (x >= xs && x <= xe) ? __ldg(qc[c] + x) : (uint16_t)0;
      any |= id[k][c] != 0;
    }
  if (!any) return false;
  // All twelve transfer-table look-ups read the ORIGINAL bytes, so they are issued together and only
  // then merged: one L2 round trip for the glyph instead of a chain of up to twelve.
  const uint32_t v[3] = {px.w0, px.w1, px.w2};
  uint32_t nv[12];
#pragma unroll
  for (int k = 0; k < 4; k++)
#pragma unroll
    for (int c = 0; c < 3; c++) {
      const int byte = 3 * k + c;                       // compile-time after unrolling
      const uint32_t old = (v[byte >> 2] >> (8 * (byte & 3))) & 0xffu;
      nv[byte] = id[k][c] ? (uint32_t)__ldg(ov.luts + (size_t)id[k][c] * 256 + old) : old;
    }
  px.w0 = nv[0] | (nv[1] << 8) | (nv[2] << 16) | (nv[3] << 24);
  px.w1 = nv[4] | (nv[5] << 8) | (nv[6] << 16) | (nv[7] << 24);
  px.w2 = nv[8] | (nv[9] << 8) | (nv[10] << 16) | (nv[11] << 24);
  return true;
}

__global__ void __launch_bounds__(kThreads, 6) overlay_tiles_kernel(const TileParams p) {
  __shared__ uint32_t hitmask[kThreads / 32];
  __shared__ int4 sops[kThreads * 2];         // staged op records (32 bytes each), one batch
  const int tiles_per_frame = p.tiles_x * p.tiles_y;
  const int frame = blockIdx.x / tiles_per_frame;
  const int t = blockIdx.x - frame * tiles_per_frame;
  const int ty = t / p.tiles_x, tx = t - ty * p.tiles_x;
  const int li = overlay_list(p.ov, frame);
  const int ob = __ldg(p.ov.op_offsets + li);
  const int nops = __ldg(p.ov.op_offsets + li + 1) - ob;
  const int tile_x0 = tx * kTileW, tile_y0 = ty * kTileH;
  const int tile_x1 = min(tile_x0 + kTileW, p.w) - 1, tile_y1 = min(tile_y0 + kTileH, p.h) - 1;
  const aiko_b200_overlay_op* ops = p.ov.ops + ob;

  // this thread's pixels
  const int y = tile_y0 + (threadIdx.x >> 4);
  const int mlo = tile_x0 + (threadIdx.x & 15) * 4, mhi = min(mlo + 3, p.w - 1);
  const bool mine = y <= tile_y1 && mlo <= tile_x1;
  uint8_t* pix = p.frames + (long long)frame * p.frame_stride + (long long)y * p.row_stride + (long long)mlo * 3;
  const bool words = mine && mlo + 3 < p.w && ((reinterpret_cast<uintptr_t>(pix) & 3) == 0);
  Pixels4 px{0, 0, 0};
  bool loaded = false, dirty = false;
  const int psize = p.ov.patch_w * p.ov.patch_h;

  for (int base = 0; base < nops; base += kThreads) {
    // ---- which ops of this batch touch the tile
    {
      const int i = base + threadIdx.x;
      bool hit = false;
      if (i < nops) {
        const int4 lo4 = __ldg(reinterpret_cast<const int4*>(ops + i));      // kind, x0, y0, x1
        const int y1 = __ldg(&ops[i].y1);
        const int grow = (lo4.x == AIKO_B200_OP_RECT) ? 1 : 0;
        hit = lo4.y - grow <= tile_x1 && lo4.w + grow >= tile_x0 && lo4.z - grow <= tile_y1 && y1 + grow >= tile_y0;
      }
      const uint32_t m = __ballot_sync(0xffffffffu, hit);
      if ((threadIdx.x & 31) == 0) hitmask[threadIdx.x >> 5] = m;
      if (!__syncthreads_or(hit)) continue;               // nothing in this batch touches the tile
    }
    // ---- stage the hit ops, in list order, in shared memory: thread t fetches the t-th hit
    int total = 0, my_word = -1, before = 0;
#pragma unroll
    for (int wd = 0; wd < kThreads / 32; wd++) {
      const int c = __popc(hitmask[wd]);
      if (my_word < 0 && (int)threadIdx.x < total + c) { my_word = wd; before = total; }
      total += c;
    }
    if ((int)threadIdx.x < total) {
      uint32_t bits = hitmask[my_word];
      for (int k = (int)threadIdx.x - before; k > 0; k--) bits &= bits - 1;
      const int4* src = reinterpret_cast<const int4*>(ops + base + my_word * 32 + (__ffs(bits) - 1));
      sops[2 * threadIdx.x] = __ldg(src);
      sops[2 * threadIdx.x + 1] = __ldg(src + 1);
    }
    __syncthreads();
    if (mine) {
      for (int i = 0; i < total; i++) {
        const int4 lo4 = sops[2 * i], hi4 = sops[2 * i + 1];
        const int grow = (lo4.x == AIKO_B200_OP_RECT) ? 1 : 0;
        if (y < lo4.z - grow || y > hi4.x + grow || mhi < lo4.y - grow || mlo > lo4.w + grow) continue;
        if (!loaded) { px = load_pixels(pix, words, mhi - mlo + 1); loaded = true; }
        dirty |= apply_op_regs(lo4, hi4, p.ov, psize, y, mlo, mhi, px);
      }
    }
    __syncthreads();                              // hitmask / sops are rewritten by the next batch
  }
  if (dirty) {
    if (words) {
      reinterpret_cast<uint32_t*>(pix)[0] = px.w0;
      reinterpret_cast<uint32_t*>(pix)[1] = px.w1;
      reinterpret_cast<uint32_t*>(pix)[2] = px.w2;
    } else {
      const uint32_t v[3] = {px.w0, px.w1, px.w2};
      for (int q = 0; q < (mhi - mlo + 1) * 3; q++) pix[q] = (uint8_t)(v[q >> 2] >> (8 * (q & 3)));
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Overlay by cells: one thread per 4-pixel group that a draw list can touch, found ON THE DEVICE
// from the 32-byte op records alone (the host ships nothing but the ops).
//   * Each op has up to four marks (overlay_marks.h): blocks of groups it can touch.  A CTA builds
//     the marks of its frame's list in shared memory together with two bit tables -- for every row
//     and for every group column, the set of marks that cover it -- so "which marks cover cell
//     (row, group)" is one AND of two bit words.
//   * Threads enumerate the cells of all marks (prefix sums over the marks, binary search per
//     thread).  A cell covered by several marks would be enumerated once per mark: only the thread
//     that comes from the FIRST covering mark owns it, the others leave.  The owner applies the
//     ops of all covering marks in list order to its 12 bytes, held in registers, and stores them
//     once; ownership makes synchronisation between ops unnecessary.
// A CTA handles kCellsPerCta consecutive cells of one frame (blockIdx.y); the grid's x extent
// comes from the host's upper bound on cells per list.
constexpr int kCellsPerCta = 256;           // default: one cell per thread (measured: see DESIGN.md)
constexpr int kCellOps = 256;              // the cell kernel takes lists of up to this many ops and marks

__device__ __forceinline__ int block_scan_exclusive(int v, int* warp_sums, int* total) {
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
#pragma unroll
  for (int w = 0; w < kThreads / 32; w++) {
    const int s = warp_sums[w];
    if (w < warp) before += s;
    all += s;
  }
  __syncthreads();                          // warp_sums may be reused right away
  *total = all;
  return before + incl - v;
}

template <int MW>                           // mask words: lists of up to 32 * MW marks
__global__ void __launch_bounds__(kThreads) overlay_cells_kernel(const TileParams p, int ops_cap) {
  extern __shared__ __align__(16) uint8_t cell_smem[];
  __shared__ int warp_sums[kThreads / 32];
  constexpr int kMarks = 32 * MW;
  const int G = (p.w + 3) >> 2;
  int4* sops = reinterpret_cast<int4*>(cell_smem);                               // [2 * ops_cap]
  OverlayMark* marks = reinterpret_cast<OverlayMark*>(sops + 2 * ops_cap);       // [kMarks]
  int* prefix = reinterpret_cast<int*>(marks + kMarks);                          // [kMarks + 1]
  uint32_t* rowmask = reinterpret_cast<uint32_t*>(prefix + kMarks + 1);          // [h][MW]
  uint32_t* gmask = rowmask + (size_t)p.h * MW;                                  // [G][MW]
  unsigned short* mark_op = reinterpret_cast<unsigned short*>(gmask + (size_t)G * MW);   // [kMarks]

  const int frame = blockIdx.y;
  const int li = overlay_list(p.ov, frame);
  const int ob = __ldg(p.ov.op_offsets + li);
  const int nops = __ldg(p.ov.op_offsets + li + 1) - ob;
  if (nops <= 0 || nops > ops_cap || nops > kThreads) return;     // (longer lists: the host picks the tile kernel)
  const int4* src = reinterpret_cast<const int4*>(p.ov.ops + ob);
  for (int i = threadIdx.x; i < nops * 2; i += kThreads) sops[i] = __ldg(src + i);
  for (int i = threadIdx.x; i < (p.h + G) * MW; i += kThreads) rowmask[i] = 0;
  // marks of op threadIdx.x (same registers as the staged copy: no barrier needed yet)
  OverlayMarks4 mine;
  mine.valid = 0;
  if ((int)threadIdx.x < nops) {
    const int4 lo4 = __ldg(src + 2 * threadIdx.x);
    const int y1 = __ldg(&p.ov.ops[ob + threadIdx.x].y1);
    mine = overlay_op_marks4(lo4.x, lo4.y, lo4.z, lo4.w, y1, p.w, p.h);
  }
  const int nmine = __popc(mine.valid);
  int nmarks;
  const int mfirst = block_scan_exclusive(nmine, warp_sums, &nmarks);
  if (nmarks > kMarks) return;                                     // uniform: the host's mark bound was wrong
  {
    int at = mfirst;
#pragma unroll
    for (int k = 0; k < kMaxMarksPerOp; k++)
      if (mine.valid >> k & 1) {
        marks[at] = mine.m[k];
        mark_op[at] = (unsigned short)threadIdx.x;
        at++;
      }
  }
  __syncthreads();
  // bit tables: warps take marks round-robin, lanes stride over the mark's rows, then its groups
  for (int m = threadIdx.x >> 5; m < nmarks; m += kThreads / 32) {
    const OverlayMark mk = marks[m];
    const uint32_t bit = 1u << (m & 31);
    const int wd = m >> 5;
    for (int r = mk.y0 + (threadIdx.x & 31); r <= mk.y1; r += 32) atomicOr(&rowmask[r * MW + wd], bit);
    for (int g = mk.g0 + (threadIdx.x & 31); g <= mk.g1; g += 32) atomicOr(&gmask[g * MW + wd], bit);
  }
  int total;
  {
    const int cells = ((int)threadIdx.x < nmarks) ? overlay_mark_cells(marks[threadIdx.x]) : 0;
    const int before = block_scan_exclusive(cells, warp_sums, &total);      // (its barriers also publish the bit tables)
    if ((int)threadIdx.x < nmarks) prefix[threadIdx.x] = before;
    if (threadIdx.x == 0) prefix[nmarks] = total;
  }
  __syncthreads();
  const int cell_end = min(total, (int)(blockIdx.x + 1) * p.cells_per_cta);
  const int psize = p.ov.patch_w * p.ov.patch_h;
  for (int c = blockIdx.x * p.cells_per_cta + threadIdx.x; c < cell_end; c += kThreads) {
    int lo = 0, hi = nmarks - 1;                        // last mark whose first cell is <= c
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (prefix[mid] <= c) lo = mid; else hi = mid - 1;
    }
    const OverlayMark mk = marks[lo];
    const int local = c - prefix[lo], ng = mk.g1 - mk.g0 + 1;
    const int r = local / ng;
    const int g = mk.g0 + (local - r * ng), y = mk.y0 + r;
    uint32_t cover[MW];
    int firstmark = -1;
#pragma unroll
    for (int w = MW - 1; w >= 0; w--) {
      cover[w] = rowmask[y * MW + w] & gmask[g * MW + w];
      if (cover[w]) firstmark = w * 32 + __ffs(cover[w]) - 1;
    }
    if (firstmark != lo) continue;                       // another thread owns this cell
    const int mlo = g * 4, mhi = min(mlo + 3, p.w - 1);
    uint8_t* pix = p.frames + (long long)frame * p.frame_stride + (long long)y * p.row_stride + (long long)mlo * 3;
    const bool words = mlo + 3 < p.w && ((reinterpret_cast<uintptr_t>(pix) & 3) == 0);
    Pixels4 px = load_pixels(pix, words, mhi - mlo + 1);
    bool dirty = false;
    int last_op = -1;
#pragma unroll
    for (int w = 0; w < MW; w++) {
      uint32_t bits = cover[w];
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        const int op = mark_op[w * 32 + b];
        if (op == last_op) continue;                     // two marks of one op
        last_op = op;
        dirty |= apply_op_regs(sops[2 * op], sops[2 * op + 1], p.ov, psize, y, mlo, mhi, px);
      }
    }
    if (dirty) {
      if (words) {
        reinterpret_cast<uint32_t*>(pix)[0] = px.w0;
        reinterpret_cast<uint32_t*>(pix)[1] = px.w1;
        reinterpret_cast<uint32_t*>(pix)[2] = px.w2;
      } else {
        const uint32_t v[3] = {px.w0, px.w1, px.w2};
        for (int q = 0; q < (mhi - mlo + 1) * 3; q++) pix[q] = (uint8_t)(v[q >> 2] >> (8 * (q & 3)));
      }
    }
  }
}

template <int MW>
static int launch_cells(const TileParams& p, int ops_cap, size_t smem, dim3 grid, cudaStream_t stream) {
  if (smem > 48 * 1024)
    AIKO_CUDA(cudaFuncSetAttribute(overlay_cells_kernel<MW>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  overlay_cells_kernel<MW><<<grid, kThreads, smem, stream>>>(p, ops_cap);
  return cuda_status(cudaGetLastError(), "overlay_cells launch");
}

// frames: uint8 [n, h, w, 3] with the given strides, modified in place
int launch_overlay_tiles(uint8_t* frames, int n, int h, int w, long long row_stride, long long frame_stride,
                         const OverlayDev& ov, cudaStream_t stream) {
  TileParams p{};
  p.frames = frames; p.n = n; p.h = h; p.w = w; p.row_stride = row_stride; p.frame_stride = frame_stride;
  p.tiles_x = (w + kTileW - 1) / kTileW;
  p.tiles_y = (h + kTileH - 1) / kTileH;
  p.ov = ov;
  const int G = (w + 3) >> 2;
  if (ov.max_cells > 0 && ov.max_marks > 0 && ov.max_marks <= kCellOps && ov.max_ops > 0 && ov.max_ops <= kCellOps &&
      ov.frame_w == w && ov.frame_h == h && h <= kMarkLimit && G <= kMarkLimit && n <= 65535 &&
      !getenv("AIKO_B200_OVERLAY_TILES")) {
    const int mw = ov.max_marks <= 32 ? 1 : ov.max_marks <= 64 ? 2 : ov.max_marks <= 128 ? 4 : 8;
    const int ops_cap = (ov.max_ops + 3) & ~3;
    const size_t smem = (size_t)ops_cap * 32 + (size_t)mw * 32 * sizeof(OverlayMark) + ((size_t)mw * 32 + 1) * 4 +
                        (size_t)(h + G) * mw * 4 + (size_t)mw * 32 * 2 + 16;
    if (smem <= (size_t)device_max_smem_optin()) {
      int per_cta = kCellsPerCta;
      if (const char* e = getenv("AIKO_B200_OVERLAY_CELLS_PER_CTA")) per_cta = atoi(e) >= kThreads ? atoi(e) / kThreads * kThreads : kCellsPerCta;
      p.cells_per_cta = per_cta;
      const dim3 grid((unsigned)((ov.max_cells + per_cta - 1) / per_cta), (unsigned)n);
      switch (mw) {
        case 1: return launch_cells<1>(p, ops_cap, smem, grid, stream);
        case 2: return launch_cells<2>(p, ops_cap, smem, grid, stream);
        case 4: return launch_cells<4>(p, ops_cap, smem, grid, stream);
        default: return launch_cells<8>(p, ops_cap, smem, grid, stream);
      }
    }
  }
  const long long grid = (long long)n * p.tiles_x * p.tiles_y;
  if (grid >= ((long long)1 << 31)) {
    set_last_error("overlay: too many tiles (%lld)", grid);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  // (Programmatic dependent launch was measured slower here, twice: with 6 resize CTAs per SM, and
  // again with 3 -- 50.5 against 48.9 us per step -- although the dependents then find free room.)
  overlay_tiles_kernel<<<(unsigned)grid, kThreads, 0, stream>>>(p);
  return cuda_status(cudaGetLastError(), "overlay_tiles launch");
}

// A planned buffer of a few KB travels as a kernel ARGUMENT (copied by the launch itself, so the host
// buffer is free again when the call returns) and is stored by one small CTA: cheaper on the stream
// than a copy-engine transfer between two kernels.
template <int WORDS>
struct PlanBlob {
  uint4 w[WORDS];
};
template <int WORDS>
__global__ void __launch_bounds__(kThreads) overlay_plan_store_kernel(const __grid_constant__ PlanBlob<WORDS> blob,
                                                                      uint4* dst, int words) {
  for (int i = threadIdx.x; i < words; i += kThreads) dst[i] = blob.w[i];
}
template <int WORDS>
static int store_plan(const void* host, size_t bytes, void* dst, cudaStream_t stream) {
  PlanBlob<WORDS> blob;
  memcpy(&blob, host, bytes);
  overlay_plan_store_kernel<WORDS><<<1, kThreads, 0, stream>>>(blob, static_cast<uint4*>(dst), (int)((bytes + 15) / 16));
  return cuda_status(cudaGetLastError(), "overlay_plan_store launch");
}

int overlay_from_abi(const aiko_b200_overlay* o, int n_frames, cudaStream_t stream, OverlayDev* out) {
  AIKO_CHECK_ARG(o && o->ops && o->op_offsets, "overlay: missing draw list (ops / op_offsets)");
  AIKO_CHECK_ARG(o->n_frames >= n_frames, "overlay: draw list made for %d frames, call has %d", o->n_frames, n_frames);
  AIKO_CHECK_ARG(o->list_of_frame || o->n_lists >= n_frames, "overlay: %d lists for %d frames and no list_of_frame",
                 o->n_lists, n_frames);
  AIKO_CHECK_ARG(o->patch_w >= 0 && o->patch_h >= 0, "overlay: bad patch size");
  if (o->upload_from_host && o->upload_bytes) {
    void* dst = const_cast<aiko_b200_overlay_op*>(o->ops);
    const size_t bytes = (size_t)o->upload_bytes;
    int rc = AIKO_B200_OK;
    if ((reinterpret_cast<uintptr_t>(dst) & 15) != 0 || bytes > 16384 || getenv("AIKO_B200_OVERLAY_UPLOAD_COPY"))
      rc = cuda_status(cudaMemcpyAsync(dst, o->upload_from_host, bytes, cudaMemcpyHostToDevice, stream), "overlay upload");
    else if (bytes <= 2048) rc = store_plan<128>(o->upload_from_host, bytes, dst, stream);
    else if (bytes <= 6144) rc = store_plan<384>(o->upload_from_host, bytes, dst, stream);
    else rc = store_plan<1024>(o->upload_from_host, bytes, dst, stream);
    if (rc != AIKO_B200_OK) return rc;
  }
  *out = OverlayDev{o->ops, o->op_offsets, o->list_of_frame, o->patch_ids, o->luts, o->patch_w, o->patch_h,
                    o->max_ops_per_list, o->max_marks_per_list, o->max_cells_per_list, o->frame_width, o->frame_height};
  return AIKO_B200_OK;
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_overlay_upload(void* plan_dev, const void* plan_host, size_t bytes, void* stream) {
  if (bytes == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(plan_dev && plan_host, "overlay_upload: null pointer");
  AIKO_CUDA(cudaMemcpyAsync(plan_dev, plan_host, bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_overlay_u8(const uint8_t* src, uint8_t* dst, int n_frames, int h, int w, int channels,
                                    int64_t row_stride, int64_t frame_stride, const aiko_b200_overlay* overlay,
                                    void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0 && h >= 0 && w >= 0, "overlay: negative size");
  if (n_frames == 0 || h == 0 || w == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "overlay: null frame pointer");
  AIKO_CHECK_ARG(channels == 1 || channels == 3, "overlay: channels %d not in {1,3}", channels);
  AIKO_CHECK_ARG(row_stride >= (int64_t)w * channels, "overlay: row stride smaller than a row");
  OverlayParams p{};
  p.src = src; p.dst = dst; p.n = n_frames; p.h = h; p.w = w; p.channels = channels;
  p.row_stride = row_stride; p.frame_stride = frame_stride;
  {
    const int rc = overlay_from_abi(overlay, n_frames, static_cast<cudaStream_t>(stream), &p.ov);
    if (rc != AIKO_B200_OK) return rc;
  }
  if (channels == 3 && !getenv("AIKO_B200_OVERLAY_ROWS")) {
    // RGB: a plain device copy (when the frames are not drawn in place), then the draw list on the
    // touched 4-pixel groups only -- the frames cross HBM once at copy speed instead of going
    // through a row kernel that culls the list for every row.
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (src != dst) {
      const size_t row_bytes = (size_t)w * 3;
      if (row_stride == (int64_t)row_bytes && frame_stride == (int64_t)row_bytes * h) {
        AIKO_CUDA(cudaMemcpyAsync(dst, src, row_bytes * h * n_frames, cudaMemcpyDeviceToDevice, st));
      } else {
        for (int f = 0; f < n_frames; f++)
          AIKO_CUDA(cudaMemcpy2DAsync(dst + (size_t)f * frame_stride, (size_t)row_stride, src + (size_t)f * frame_stride,
                                      (size_t)row_stride, row_bytes, (size_t)h, cudaMemcpyDeviceToDevice, st));
      }
    }
    return launch_overlay_tiles(dst, n_frames, h, w, row_stride, frame_stride, p.ov, st);
  }
  const size_t smem = ((size_t)w * 3 + 15) & ~(size_t)15;
  AIKO_CHECK_ARG(smem <= (size_t)device_max_smem_optin(), "overlay: row of %zu bytes exceeds shared memory", smem);
  AIKO_CUDA(cudaFuncSetAttribute(overlay_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long rows = (long long)n_frames * h;
  long long grid = (long long)sm_count() * 32;
  if (grid > rows) grid = rows;
  overlay_rows_kernel<<<(unsigned)grid, kThreads, smem, (cudaStream_t)stream>>>(p);
  return cuda_status(cudaGetLastError(), "overlay launch");
}
