// Antialiased bicubic resize (Pillow arithmetic, see resize_bicubic.cu) on the tensor cores.
//
// Why tensor cores for byte work: Pillow's two passes are ~35 exact integer multiply-adds per
// source pixel (1080p -> 640x480); on the integer pipe that is 10x more time than the frame takes
// to cross HBM.  Both passes are banded matrix products, and an integer MMA computes them exactly:
// a 22-bit weight is split into three byte planes, w = (w2 << 16) + (w1 << 8) + w0 with w2 signed,
// and  sum(w * p) = ((sum(w2 p) << 8) + sum(w1 p) << 8) + sum(w0 p)  in wrapping 32-bit arithmetic,
// which is the arithmetic Pillow's int accumulators use.  Each plane is one
// mma.sync.m16n8k32 (u8/s8 x u8 -> s32); the rounding constant 1 << 21 enters as the initial
// accumulator 32 of the top plane.  The result is bit-identical to the scalar kernel.
//
//   horizontal pass   D[16 dx][8 rows]  = A[16 dx][K px]   x B[K px][8 rows]
//                     A = weights of one 16-pixel tile, kept in registers for the whole kernel;
//                     B = source pixels of one channel, gathered from the interleaved rows the
//                     TMA engine staged in shared memory (3 words -> 3 channel registers).
//   vertical pass     D[16 bytes][8 dy] = A[16 bytes][K rows] x B[K rows][8 dy]
//                     A = the uint8 intermediate image, held transposed ([byte column][row]) in a
//                     shared-memory ring that the horizontal epilogue writes; B = weights.
//
// A work item is (frame, band of output rows, chunk of 16-pixel tiles); persistent CTAs stride over
// the items.  Inside a CTA every consumer warp owns a fixed run of tiles and does horizontal pass,
// ring and vertical pass for those columns alone, so warps synchronise only with the producer
// warp (full / empty mbarriers of the two staging buffers), never with each other.
//
// tcgen05 is not used on purpose: the weight matrices are banded (12 taps wide at scale 3), and
// the band fills 19 % of a 16-output x 64-pixel block but only 3-5 % of the 64/128-row blocks
// tcgen05.mma works on, which costs more than its higher rate returns: measured on B200
// (benchmarks/tcgen05_i8_rate.cu) kind::i8 runs at 2.1x (M = 64) / 4.2x (M = 128) the rate of
// mma.sync.m16n8k32, i.e. 50-55 us of tensor-pipe time for this batch against 31 us here.
#include "bicubic.cuh"

namespace aiko {

constexpr int kBatchRows = 8;        // source rows per staging buffer = N of the horizontal MMA
constexpr int kMmaStages = 2;

// ---------------------------------------------------------------------------------------------
// Table builders (run once per size pair by aiko_b200_resize_bicubic_prepare, after the
// coefficient kernels on the same stream).
__device__ __forceinline__ uint32_t weight_plane_byte(int w, int plane, int* bad) {
  if (plane == 0) return (uint32_t)w & 255u;
  if (plane == 1) return ((uint32_t)w >> 8) & 255u;
  const int top = w >> 16;
  if (top < -128 || top > 127) *bad = 1;
  return (uint32_t)top & 255u;
}

// weight of source index `src` for output `out`, 0 outside the output's taps
__device__ __forceinline__ int tap_weight(const int* bounds, const int* k, int ksize, int out, int n_out, int src) {
  if (out >= n_out) return 0;
  const int t = src - bounds[2 * out];
  return (t >= 0 && t < bounds[2 * out + 1]) ? k[(size_t)out * ksize + t] : 0;
}

__global__ void bicubic_x_frags_kernel(int* ws, int dw, int ksize, int off_bounds, int off_k, int kx, int ntx,
                                       int off_xa, int off_ax) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = id & 31, frag = id >> 5;   // frag = (tile * 3 + plane) * kx + ks
  if (frag >= ntx * 3 * kx) return;
  const int ks = frag % kx, plane = (frag / kx) % 3, tile = frag / (3 * kx);
  const int* bounds = ws + off_bounds;
  const int* k = ws + off_k;
  const int dx0 = tile * 16;
  const int xa = bounds[2 * dx0] & ~3;
  const int g = lane >> 2, q = lane & 3;
  int bad = 0;
  uint32_t regs[4];
  for (int r = 0; r < 4; r++) {      // a0: (g, 4q) a1: (g + 8, 4q) a2: (g, 16 + 4q) a3: (g + 8, 16 + 4q)
    const int dx = dx0 + g + 8 * (r & 1);
    const int px0 = xa + 32 * ks + 4 * q + 16 * (r >> 1);
    uint32_t v = 0;
    for (int i = 0; i < 4; i++)
      v |= weight_plane_byte(tap_weight(bounds, k, ksize, dx, dw, px0 + i), plane, &bad) << (8 * i);
    regs[r] = v;
  }
  reinterpret_cast<uint4*>(ws + off_ax)[(size_t)frag * 32 + lane] = make_uint4(regs[0], regs[1], regs[2], regs[3]);
  if (plane == 0 && ks == 0 && lane == 0) {
    ws[off_xa + tile] = xa;
    const int last = min(dx0 + 15, dw - 1);
    if (bounds[2 * last] + bounds[2 * last + 1] - xa > 32 * kx) bad = 1;   // window bound violated
  }
  if (bad) ws[1] = 1;
}

__global__ void bicubic_y_frags_kernel(int* ws, int dh, int ksize, int off_bounds, int off_k, int ky, int ngy,
                                       int off_ya, int off_yend, int off_by) {
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = id & 31, frag = id >> 5;   // frag = (group * 3 + plane) * ky + ks
  if (frag >= ngy * 3 * ky) return;
  const int ks = frag % ky, plane = (frag / ky) % 3, group = frag / (3 * ky);
  const int* bounds = ws + off_bounds;
  const int* k = ws + off_k;
  const int dy0 = group * 8;
  const int ya = bounds[2 * dy0] & ~3;
  const int g = lane >> 2, q = lane & 3;
  int bad = 0;
  uint32_t regs[2];
  for (int r = 0; r < 2; r++) {      // b0: (k = 4q, n = g)  b1: (k = 16 + 4q, n = g)
    const int row0 = ya + 32 * ks + 4 * q + 16 * r;
    uint32_t v = 0;
    for (int i = 0; i < 4; i++)
      v |= weight_plane_byte(tap_weight(bounds, k, ksize, dy0 + g, dh, row0 + i), plane, &bad) << (8 * i);
    regs[r] = v;
  }
  reinterpret_cast<uint2*>(ws + off_by)[(size_t)frag * 32 + lane] = make_uint2(regs[0], regs[1]);
  if (plane == 0 && ks == 0 && lane == 0) {
    const int last = min(dy0 + 7, dh - 1);
    const int yend = bounds[2 * last] + bounds[2 * last + 1];
    ws[off_ya + group] = ya;
    ws[off_yend + group] = yend;
    if (yend - ya > 32 * ky) bad = 1;
  }
  if (bad) ws[1] = 1;
}

int bicubic_mma_prepare(int* ws, const BicubicLayout& l, int sh, int sw, int dh, int dw, cudaStream_t stream) {
  (void)sh; (void)sw;
  const MmaPlan& m = l.mma;
  if (!m.ok) return AIKO_B200_OK;
  const int nx = m.ntx * 3 * m.kx * 32, ny = m.ngy * 3 * m.ky * 32;
  bicubic_x_frags_kernel<<<(nx + 127) / 128, 128, 0, stream>>>(ws, dw, l.x.ksize, l.x.off_bounds, l.x.off_k, m.kx,
                                                                m.ntx, m.off_xa, m.off_ax);
  bicubic_y_frags_kernel<<<(ny + 127) / 128, 128, 0, stream>>>(ws, dh, l.y.ksize, l.y.off_bounds, l.y.off_k, m.ky,
                                                                m.ngy, m.off_ya, m.off_yend, m.off_by);
  return cuda_status(cudaGetLastError(), "bicubic table launch");
}

// ---------------------------------------------------------------------------------------------
struct MmaParams {
  const uint8_t* src;
  uint8_t* dst;
  const int* ws;
  int n, sh, sw, dh, dw;
  int ntx, ngy;
  int off_xa, off_ax, off_ya, off_yend, off_by;
  int band_groups, bands_per_frame;   // output-row groups (8 rows) per CTA
  int chunks, tiles_per_chunk;        // 16-pixel tiles per CTA
  int raw_pitch;                      // bytes between staged source rows (= 48 mod 128: see gather)
  int items;                          // n * bands_per_frame * chunks
};

// D += A x B, m16n8k32, 8-bit operands, wrapping 32-bit accumulators.
__device__ __forceinline__ void mma_u8u8(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_s8u8(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.s8.u8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void mma_u8s8(int (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(d[0]), "+r"(d[1]), "+r"(d[2]), "+r"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// Three plane sums -> the 32-bit sum Pillow forms (wrapping), with the rounding constant already
// inside s0; two of them shifted down, saturated to bytes and packed (lo | hi << 8) by one
// cvt.pack.sat.
__device__ __forceinline__ int combine_planes(int s2, int s1, int s0) {
  return (int)((((((uint32_t)s2) << 8) + (uint32_t)s1) << 8) + (uint32_t)s0);
}
__device__ __forceinline__ uint32_t pack2_sat_u8(int lo, int hi) {
  uint32_t d;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi >> kPrecisionBits), "r"(lo >> kPrecisionBits), "r"(0));
  return d;
}

// Horizontal pass of one warp: 8 staged rows x NT tiles -> 8 rows of the ring.  NT is a template
// parameter so the tiles unroll without branches and their loads, MMAs and epilogues interleave.
// UA: the staged row starts `delta` (0..15) bytes into its buffer (rows that are not 16-byte
// aligned in global memory are copied from the boundary below them): four aligned words and a
// funnel shift replace the three aligned words of a gather.
template <int KX, int T, int CAP, int NT, bool UA, int C>
__device__ __forceinline__ void horizontal_tiles(const uint4 (&A)[T][3][KX], const int (&xoff)[T],
                                                 const uint8_t* __restrict__ rrow, uint8_t* __restrict__ ring_out,
                                                 int q, int delta) {
  constexpr int FIRST = 0;
  const uint8_t* rbase = rrow + (UA ? (delta & ~3) : 0);
  const int shift = UA ? (delta & 3) * 8 : 0;
  constexpr int kRound = 1 << (kPrecisionBits - 1);
#pragma unroll
  for (int t = FIRST; t < FIRST + NT; t++) {
    uint32_t B[KX][2][C];
#pragma unroll
    for (int ks = 0; ks < KX; ks++)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        // 4 pixels = 4 * C interleaved bytes = C words -> one register per channel.  (RGB: row pitch
        // 12 words mod 32 and 3 words per lane: the 32 lanes of each load hit 32 banks.)
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(rbase + xoff[t] + (32 * ks + 16 * h + 4 * q) * C);
        uint32_t w[C + 1];
#pragma unroll
        for (int i = 0; i < C; i++) w[i] = wp[i];
        if (UA) {
          w[C] = wp[C];
#pragma unroll
          for (int i = 0; i < C; i++) w[i] = __funnelshift_r(w[i], w[i + 1], shift);
        }
        if (C == 3) {
          B[ks][h][0] = __byte_perm(__byte_perm(w[0], w[1], 0x0630), w[2], 0x5210);
          B[ks][h][1] = __byte_perm(__byte_perm(w[0], w[1], 0x0741), w[2], 0x6210);
          B[ks][h][2 % C] = __byte_perm(__byte_perm(w[0], w[1], 0x0052), w[2 % C], 0x7410);
        } else if (C == 4) {
#pragma unroll
          for (int c = 0; c < C; c++)   // byte c of each of the four words
            B[ks][h][c] = __byte_perm(__byte_perm(w[0], w[1 % C], 0x0040 + 0x11 * c), __byte_perm(w[2 % C], w[3 % C], 0x0040 + 0x11 * c), 0x5410);
        } else {
          B[ks][h][0] = w[0];
        }
      }
#pragma unroll
    for (int c = 0; c < C; c++) {
      // one accumulator set per weight plane: three independent MMA chains
      int s2[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0}, s0[4] = {kRound, kRound, kRound, kRound};
#pragma unroll
      for (int ks = 0; ks < KX; ks++) {
        mma_s8u8(s2, A[t][2][ks].x, A[t][2][ks].y, A[t][2][ks].z, A[t][2][ks].w, B[ks][0][c], B[ks][1][c]);
        mma_u8u8(s1, A[t][1][ks].x, A[t][1][ks].y, A[t][1][ks].z, A[t][1][ks].w, B[ks][0][c], B[ks][1][c]);
        mma_u8u8(s0, A[t][0][ks].x, A[t][0][ks].y, A[t][0][ks].z, A[t][0][ks].w, B[ks][0][c], B[ks][1][c]);
      }
      // element 0, 1: pixel g, rows 2q, 2q + 1;  element 2, 3: pixel g + 8 (ring_out is at pixel g, row 2q)
      uint8_t* o = ring_out + (size_t)(t * 16 * C + c) * CAP;
      *reinterpret_cast<uint16_t*>(o) = (uint16_t)pack2_sat_u8(combine_planes(s2[0], s1[0], s0[0]),
                                                               combine_planes(s2[1], s1[1], s0[1]));
      *reinterpret_cast<uint16_t*>(o + 8 * C * CAP) = (uint16_t)pack2_sat_u8(combine_planes(s2[2], s1[2], s0[2]),
                                                                          combine_planes(s2[3], s1[3], s0[3]));
    }
  }
}

// A consumer warp keeps the weights of its T tiles in 12 * KX * T registers (96 for the settings
// used: 10 consumer warps + the producer warp at <= 168 registers).  More, thinner warps (T halved,
// 20 warps) and a producer-less variant with 12 evenly loaded warps were both measured slower
// (99 / 101 us against 87 us for 32 x 1080p): spills, and more per-warp overhead per batch.
template <int KX, int T>
struct WarpBudget {
  static constexpr int max_warps = 10;   // consumers
};

// What one work item (frame, band of output-row groups, chunk of tiles) covers.
struct MmaItem {
  int frame, j0, j1, tile0, tile1;
  int row_start, nbatches;     // first source row, batches of 8 rows
  int byte_lo, copy_bytes;     // source bytes per row the chunk needs
};
template <int KX, int C>
__device__ __forceinline__ MmaItem mma_item(const MmaParams& p, int item) {
  MmaItem it;
  const int chunk = item % p.chunks;
  item /= p.chunks;
  const int band = item % p.bands_per_frame;
  it.frame = item / p.bands_per_frame;
  it.tile0 = chunk * p.tiles_per_chunk;
  it.tile1 = min(p.ntx, it.tile0 + p.tiles_per_chunk);
  it.j0 = band * p.band_groups;
  it.j1 = min(p.ngy, it.j0 + p.band_groups);
  it.row_start = __ldg(p.ws + p.off_ya + it.j0);
  const int row_end = __ldg(p.ws + p.off_yend + it.j1 - 1);
  const int xa_lo = __ldg(p.ws + p.off_xa + it.tile0), xa_hi = __ldg(p.ws + p.off_xa + it.tile1 - 1);
  it.nbatches = (row_end - it.row_start + kBatchRows - 1) / kBatchRows;
  it.byte_lo = (xa_lo * C) & ~15;
  it.copy_bytes = min(p.sw * C, ((xa_hi + 32 * KX) * C + 15) & ~15) - it.byte_lo;
  return it;
}

// Persistent CTAs: CTA c takes work items c, c + gridDim.x, ...; the producer warp runs ahead of
// the consumers across item boundaries, so only the very first batch of a CTA waits for HBM with
// nothing else to do, and a warp loads its tile weights once (the chunk of a CTA never changes:
// the grid is a multiple of the chunk count).
template <int KX, int KY, int T, bool UA, int C>
__global__ void __launch_bounds__((WarpBudget<KX, T>::max_warps + 1) * 32, 1) bicubic_mma_kernel(const MmaParams p) {
  constexpr int CAP = 32 * KY + 16;   // ring rows: one K window + the rows a batch runs ahead, rounded;
                                      // as a pitch (bytes) it is 16 mod 32, which keeps the loads below conflict-free
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);      // per staging buffer: its rows have landed
  uint64_t* empty = full + kMmaStages;                     // per staging buffer: every consumer is done with it
  int* deltas = reinterpret_cast<int*>(smem + 64);         // per staging buffer and row: bytes before the row's first byte
  uint8_t* raw = smem + 128;
  uint8_t* ring = raw + (size_t)kMmaStages * kBatchRows * p.raw_pitch;   // [byte column][CAP rows]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, q = lane & 3;
  const int nwarps = (int)blockDim.x / 32 - 1;   // consumer warps; the last warp feeds the TMA ring
  // A weight or window outside what the tables can hold (cannot happen for the sizes MmaPlan.ok
  // admits): prepare left a non-zero word 1 in the workspace, which the host reports; the kernel
  // leaves the output alone instead of trapping (a trap would poison the whole CUDA context).
  if (__ldg(p.ws + 1) != 0) return;
  const int row_bytes = p.sw * C;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kMmaStages; s++) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], nwarps);     // every consumer warp arrives, with or without tiles in the chunk
    }
    fence_mbar_init();
  }
  __syncthreads();

  if (warp == nwarps) {
    // ---- producer: lanes 0..7 each ask the TMA engine for one source row of the batch ------------
    int gb = 0;   // batches issued by this CTA so far: staging buffer and barrier phase follow from it
    const uint64_t policy = l2_policy_evict_first();   // source rows are read once (plus the band halo)
    for (int item = blockIdx.x; item < p.items; item += gridDim.x) {
      const MmaItem it = mma_item<KX, C>(p, item);
      if (it.copy_bytes + (UA ? 32 : 0) > p.raw_pitch) continue;   // (consumers skip the same items)
      const uint8_t* fsrc = p.src + (size_t)it.frame * p.sh * row_bytes + it.byte_lo;
      for (int b = 0; b < it.nbatches; b++, gb++) {
        const int s = gb % kMmaStages, round = gb / kMmaStages;
        if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1), 1000, 100);
        const int row = min(it.row_start + b * kBatchRows + (lane & 7), p.sh - 1);   // rows past the frame carry no weight
        const uint8_t* a = fsrc + (size_t)row * row_bytes;
        const uint32_t d = UA ? (uint32_t)(reinterpret_cast<uintptr_t>(a) & 15) : 0u;
        const uint32_t nbytes = UA ? ((d + (uint32_t)it.copy_bytes + 15u) & ~15u) : (uint32_t)it.copy_bytes;
        if (UA && lane < kBatchRows) deltas[s * kBatchRows + lane] = (int)d;
        const uint32_t total = __reduce_add_sync(0xffffffffu, lane < kBatchRows ? nbytes : 0u);
        __syncwarp();
        if (lane == 0) mbar_expect_tx(&full[s], total);   // release: the row offsets above are visible with it
        __syncwarp();
        if (lane < kBatchRows)
          bulk_g2s_hint(raw + (size_t)(s * kBatchRows + lane) * p.raw_pitch, a - d, nbytes, &full[s], policy);
      }
    }
    return;
  }

  // ---- consumers: horizontal pass, ring and vertical pass of the warp's own tiles ---------------
  uint4 A[T][3][KX];
  int xoff[T];
  const int wcol0 = warp * T * 16 * C;                // first ring column (byte column inside the chunk)
  const uint2* byf = reinterpret_cast<const uint2*>(p.ws + p.off_by);
  const int* ya = p.ws + p.off_ya;
  const int* yend = p.ws + p.off_yend;
  const int out_row_bytes = p.dw * C;
  constexpr int kRound = 1 << (kPrecisionBits - 1);
  int cur_tile0 = -1, ntw = 0;
  int gb = 0;
  for (int item = blockIdx.x; item < p.items; item += gridDim.x) {
    const MmaItem it = mma_item<KX, C>(p, item);
    if (it.copy_bytes + (UA ? 32 : 0) > p.raw_pitch) continue;     // never true for admitted sizes; see the producer
    if (it.tile0 != cur_tile0) {     // first item of the CTA (and never again when the grid is a multiple of the chunks)
      cur_tile0 = it.tile0;
      const int wtile0 = it.tile0 + warp * T;
      ntw = max(0, min(T, it.tile1 - wtile0));
      const uint4* ax = reinterpret_cast<const uint4*>(p.ws + p.off_ax);
#pragma unroll
      for (int t = 0; t < T; t++) {
        const bool on = t < ntw;     // a slot without a tile gets zero weights and is filtered like the others
        xoff[t] = on ? __ldg(p.ws + p.off_xa + wtile0 + t) * C - it.byte_lo : 0;
#pragma unroll
        for (int pl = 0; pl < 3; pl++)
#pragma unroll
          for (int ks = 0; ks < KX; ks++)
            A[t][pl][ks] = on ? __ldg(ax + ((size_t)((wtile0 + t) * 3 + pl) * KX + ks) * 32 + lane) : make_uint4(0, 0, 0, 0);
      }
    }
    const int row_start = it.row_start, j1 = it.j1;
    uint8_t* fdst = p.dst + (size_t)it.frame * p.dh * out_row_bytes;
    const int gcol0 = it.tile0 * 16 * C + wcol0;      // first output byte column of this warp
    const bool cols_inside = gcol0 + ntw * 16 * C <= out_row_bytes;

    int jn = it.j0;       // next group of output rows to emit
    int slot = 0;         // ring row of the batch being filtered
    // the next group's record and weights are fetched a group ahead, so no batch waits on L2 for them
    int g_end = __ldg(yend + jn) - row_start, g_base = 0;
    uint2 W[3][KY];
#pragma unroll
    for (int pl = 0; pl < 3; pl++)
#pragma unroll
      for (int ks = 0; ks < KY; ks++) W[pl][ks] = __ldg(byf + ((size_t)(jn * 3 + pl) * KY + ks) * 32 + lane);
    for (int b = 0; b < it.nbatches; b++, gb++) {
      const int s = gb % kMmaStages, round = gb / kMmaStages;
      mbar_wait(&full[s], (uint32_t)(round & 1), 500, 0);
      if (ntw > 0) {
        // -- horizontal pass: 8 staged rows x this warp's tiles -> ring rows slot .. slot + 7
        const uint8_t* rrow = raw + (size_t)(s * kBatchRows + g) * p.raw_pitch;
        uint8_t* ring_out = ring + (size_t)(wcol0 + g * C) * CAP + slot + 2 * q;
        // slots without a tile: zero weights, spare columns
        horizontal_tiles<KX, T, CAP, T, UA, C>(A, xoff, rrow, ring_out, q, UA ? deltas[s * kBatchRows + g] : 0);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[s]);     // the staged rows are free again
      slot += kBatchRows;
      if (slot >= CAP) slot -= CAP;
      const int produced = (b + 1) * kBatchRows;

      // -- vertical pass for every group of 8 output rows whose last intermediate row now exists
      while (jn < j1 && g_end <= produced) {
        const int base = g_base % CAP;
        int o0[KY], o1[KY];
#pragma unroll
        for (int ks = 0; ks < KY; ks++) {
          o0[ks] = base + 32 * ks + 4 * q;
          if (o0[ks] >= CAP) o0[ks] -= CAP;
          o1[ks] = base + 32 * ks + 16 + 4 * q;
          if (o1[ks] >= CAP) o1[ks] -= CAP;
        }
        const int dy = jn * 8 + 2 * q;
        const bool inside = cols_inside && jn * 8 + 8 <= p.dh;   // no clipping of this group's stores
        uint8_t* orow = fdst + (size_t)dy * out_row_bytes + gcol0 + g;
#pragma unroll 3
        for (int mt = 0; mt < ntw * C; mt++) {
          const uint8_t* c0 = ring + (size_t)(wcol0 + mt * 16 + g) * CAP;
          const uint8_t* c1 = c0 + 8 * CAP;
          int s2[4] = {0, 0, 0, 0}, s1[4] = {0, 0, 0, 0}, s0[4] = {kRound, kRound, kRound, kRound};
#pragma unroll
          for (int ks = 0; ks < KY; ks++) {
            const uint32_t a0 = *reinterpret_cast<const uint32_t*>(c0 + o0[ks]);
            const uint32_t a1 = *reinterpret_cast<const uint32_t*>(c1 + o0[ks]);
            const uint32_t a2 = *reinterpret_cast<const uint32_t*>(c0 + o1[ks]);
            const uint32_t a3 = *reinterpret_cast<const uint32_t*>(c1 + o1[ks]);
            mma_u8s8(s2, a0, a1, a2, a3, W[2][ks].x, W[2][ks].y);
            mma_u8u8(s1, a0, a1, a2, a3, W[1][ks].x, W[1][ks].y);
            mma_u8u8(s0, a0, a1, a2, a3, W[0][ks].x, W[0][ks].y);
          }
          // element 0, 1: byte column g, output rows dy, dy + 1;  element 2, 3: byte column g + 8
          const uint32_t lo = pack2_sat_u8(combine_planes(s2[0], s1[0], s0[0]), combine_planes(s2[1], s1[1], s0[1]));
          const uint32_t hi = pack2_sat_u8(combine_planes(s2[2], s1[2], s0[2]), combine_planes(s2[3], s1[3], s0[3]));
          uint8_t* o = orow + mt * 16;
          if (inside) {
            o[0] = (uint8_t)lo;
            o[8] = (uint8_t)hi;
            o[out_row_bytes] = (uint8_t)(lo >> 8);
            o[out_row_bytes + 8] = (uint8_t)(hi >> 8);
          } else {
            const int col = gcol0 + mt * 16 + g;
            if (dy < p.dh) {
              if (col < out_row_bytes) o[0] = (uint8_t)lo;
              if (col + 8 < out_row_bytes) o[8] = (uint8_t)hi;
            }
            if (dy + 1 < p.dh) {
              if (col < out_row_bytes) o[out_row_bytes] = (uint8_t)(lo >> 8);
              if (col + 8 < out_row_bytes) o[out_row_bytes + 8] = (uint8_t)(hi >> 8);
            }
          }
        }
        jn++;
        if (jn < j1) {
          g_end = __ldg(yend + jn) - row_start;
          g_base = __ldg(ya + jn) - row_start;
#pragma unroll
          for (int pl = 0; pl < 3; pl++)
#pragma unroll
            for (int ks = 0; ks < KY; ks++) W[pl][ks] = __ldg(byf + ((size_t)(jn * 3 + pl) * KY + ks) * 32 + lane);
        }
      }
      __syncwarp();   // ring rows read above are rewritten by later batches
    }
  }
}

static int env_int(const char* name, int fallback) {
  const char* v = getenv(name);
  return v ? atoi(v) : fallback;
}

template <int KX, int KY, int T, bool UA, int C>
static int launch_mma(MmaParams p, int n_frames, double scale_y, cudaStream_t stream, bool* done) {
  constexpr int kMmaMaxWarps = WarpBudget<KX, T>::max_warps;
  constexpr int CAP = 32 * KY + 16;
  const int row_bytes = p.sw * C;
  const double scale_x = (double)p.sw / p.dw;
  const size_t max_smem = (size_t)device_max_smem_optin();
  int chunks = (p.ntx + kMmaMaxWarps * T - 1) / (kMmaMaxWarps * T);
  chunks = max(chunks, env_int("AIKO_B200_BICUBIC_CHUNKS", 1));
  size_t smem = 0;
  int nwarps = 0;
  for (;; chunks++) {
    if (chunks > p.ntx) return AIKO_B200_OK;   // does not fit: the scalar kernel takes it
    p.chunks = chunks;
    p.tiles_per_chunk = (p.ntx + chunks - 1) / chunks;
    p.chunks = (p.ntx + p.tiles_per_chunk - 1) / p.tiles_per_chunk;
    nwarps = (p.tiles_per_chunk + T - 1) / T;
    // source bytes a chunk touches per row (window starts are rounded down to 4 pixels, the copy to 16 bytes)
    const long long bound_px = (long long)(p.tiles_per_chunk * 16 * scale_x) + 32 * KX + 8;
    long long bound_bytes = (bound_px * C + 32 + 15) & ~15LL;
    if (bound_bytes > row_bytes) bound_bytes = (row_bytes + 15) & ~15;
    if (UA) bound_bytes += 32;                                // the 16-byte rounding at both ends of a row
    long long pitch = (bound_bytes + 127) / 128 * 128 + 48;   // 48 mod 128: see the gather in the kernel
    if (pitch - 128 >= bound_bytes) pitch -= 128;
    p.raw_pitch = (int)pitch;
    smem = 128 + (size_t)kMmaStages * kBatchRows * (size_t)pitch + (size_t)nwarps * T * 16 * C * CAP + 128;
    if (smem <= max_smem) break;
  }
  auto kernel = bicubic_mma_kernel<KX, KY, T, UA, C>;
  AIKO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int threads = (nwarps + 1) * 32;
  int per_sm = 0;
  AIKO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
  if (per_sm < 1) per_sm = 1;
  // bands of output-row groups: long enough that the rows filtered twice (the vertical support at
  // each band edge) stay few, and a CTA count that fills whole waves of the machine
  const long long slots = (long long)sm_count() * per_sm;
  const double rows_per_group = 8.0 * scale_y, halo = 4.0 * (scale_y < 1.0 ? 1.0 : scale_y) + kBatchRows;
  int best_g = 1;
  double best = -1.0;
  for (int gsz = 1; gsz <= p.ngy; gsz++) {
    const long long ctas = (long long)n_frames * p.chunks * ((p.ngy + gsz - 1) / gsz);
    const double wave_eff = (double)ctas / (double)(((ctas + slots - 1) / slots) * slots);
    const double halo_eff = gsz * rows_per_group / (gsz * rows_per_group + halo);
    if (wave_eff * halo_eff > best) { best = wave_eff * halo_eff; best_g = gsz; }
  }
  p.band_groups = env_int("AIKO_B200_BICUBIC_BAND_GROUPS", best_g);
  if (p.band_groups < 1) p.band_groups = 1;
  p.bands_per_frame = (p.ngy + p.band_groups - 1) / p.band_groups;
  const long long items = (long long)n_frames * p.chunks * p.bands_per_frame;
  AIKO_CHECK_ARG(items < ((long long)1 << 31), "resize_bicubic: too many bands");
  p.items = (int)items;
  long long grid = items < slots ? items : slots;       // persistent CTAs stride over the items
  if (grid > p.chunks) grid -= grid % p.chunks;         // ... and keep their chunk (= their tile weights)
  kernel<<<(unsigned)grid, threads, smem, stream>>>(p);
  *done = true;
  return cuda_status(cudaGetLastError(), "bicubic_mma launch");
}

int bicubic_mma_launch(const uint8_t* src, int n_frames, int sh, int sw, int channels, uint8_t* dst, int dh, int dw,
                       const int* ws, const BicubicLayout& l, cudaStream_t stream, bool* done) {
  *done = false;
  const MmaPlan& m = l.mma;
  // 1-, 3- and 4-channel frames (rows off 16-byte boundaries are copied from the boundary below them);
  // reductions beyond the tables (scale > ~5): scalar kernel
  if (!m.ok || env_int("AIKO_B200_BICUBIC_SCALAR", 0)) return AIKO_B200_OK;
  const bool unaligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)((size_t)sw * channels)) & 15) != 0;
  MmaParams p{};
  p.src = src; p.dst = dst; p.ws = ws;
  p.n = n_frames; p.sh = sh; p.sw = sw; p.dh = dh; p.dw = dw;
  p.ntx = m.ntx; p.ngy = m.ngy;
  p.off_xa = m.off_xa; p.off_ax = m.off_ax; p.off_ya = m.off_ya; p.off_yend = m.off_yend; p.off_by = m.off_by;
  const double scale_y = (double)sh / dh;
#define AIKO_MMA_CH(KX, KY, T, C)                                                                 \
  return unaligned ? launch_mma<KX, KY, T, true, C>(p, n_frames, scale_y, stream, done)              \
                   : launch_mma<KX, KY, T, false, C>(p, n_frames, scale_y, stream, done);
#define AIKO_MMA_CASE(KX, KY, T)                                                                   \
  if (m.kx == KX && m.ky == KY) {                                                                  \
    if (channels == 3) { AIKO_MMA_CH(KX, KY, T, 3) }                                               \
    if (channels == 1) { AIKO_MMA_CH(KX, KY, T, 1) }                                               \
    AIKO_MMA_CH(KX, KY, T, 4)                                                                      \
  }
  AIKO_MMA_CASE(1, 1, 8) AIKO_MMA_CASE(2, 1, 4) AIKO_MMA_CASE(3, 1, 2) AIKO_MMA_CASE(4, 1, 2)
  AIKO_MMA_CASE(1, 2, 8) AIKO_MMA_CASE(2, 2, 4) AIKO_MMA_CASE(3, 2, 2) AIKO_MMA_CASE(4, 2, 2)
#undef AIKO_MMA_CASE
#undef AIKO_MMA_CH
  return AIKO_B200_OK;
}

}  // namespace aiko
