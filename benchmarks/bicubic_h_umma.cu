// The horizontal pass of the Pillow antialiased bicubic resize (1920 -> 640, the first half of BASELINE config 1's
// ImageResize; reference: R/elements/media/image_io.py:218 -> Pillow ImagingResampleHorizontal_8bpc) written for the
// 5th-generation tensor cores, as a MEASUREMENT next to the product kernel (csrc/resize_bicubic_mma.cu, mma.sync):
// VERDICT r01 "next round" item 7 asks for exactly this experiment.  Not part of the library.
//
// Formulation.  The frames are interleaved RGB, so the reduction index that is contiguous in memory is the source
// BYTE.  D[128 image rows][144] += A[128 rows][K = 224 source bytes] * B[144][K]:
//   A = the image itself, K-major (a row's bytes are contiguous).  TMA tensor copies (this GPU's memory only, no
//       multicast) stage a stripe of 128 rows as 128-byte swizzle atoms ([atom][row][128 B], CU_TENSOR_MAP_SWIZZLE_128B);
//       a tile of 16 output pixels reads a 224-byte window of it that starts on a 32-byte boundary (32 bytes before
//       source byte 144 * tile for even tiles, 48 bytes before for odd ones), so that no K = 32 step straddles two
//       swizzle atoms; the UMMA descriptor's start address sits 0 / 32 / 64 / 96 bytes into an atom.
//   B = the banded weights of 16 output pixels x 3 channels against those 224 bytes, identical for every interior
//       tile (scale 3: shift-invariant; one set for even tiles, one for odd), as three SIGNED base-256 digit planes of
//       the 22-bit fixed-point weights (w = p0 + 256 p1 + 65536 p2, each in -128..127) stacked along N: one N = 144
//       instruction per K step reads A once for all three planes.  No-swizzle core-matrix layout, written once per CTA.
//   D = 144 s32 accumulator columns per tile in tensor memory (three slots), combined in the epilogue:
//       (d0 + (d1 << 8) + (d2 << 16) + 2^21) >> 22, clipped to 0..255 -- Pillow's arithmetic, bit for bit.
// The batch is treated as one image of 32 * 1080 rows (the pass never looks across rows).  The two output pixels at
// each end of a row have clamped windows / renormalised weights: the epilogue lanes that own them recompute them with
// scalar code.  Warps: 24 epilogue (TMEM lane quarter = warp % 4, column third, two groups taking alternate tiles),
// 1 MMA issuer (elected lane, descriptors on the uniform datapath), 1 loader (one elected lane, TMA).
//
// Measured on B200 (profiles/bicubic_h_umma_r02.jsonl): bit-exact, 77 us for this pass alone on 32 x 1080p; the product
// kernel does BOTH passes in 87 us, so it stays.  Versions on the way: 16-byte cp.async staging 152-170 us (LDGSTS
// issue-bound), 3-D TMA boxes with 16-byte rows 115 us, three N = 48 instructions per K step 92 us (shared-memory
// operand bandwidth: an M = 128 K = 32 A tile is 4 KB = 32 clocks of the 128 B/clk shared memory per instruction,
// whatever N is).  What is left is the epilogue's integer arithmetic (three 32-bit planes per output byte: ~1,560 warp
// instructions per tile at ~1.2 per clock), not the tensor pipe (37 k of ~125 k clocks): the issuer waits for a free
// accumulator slot about half of the time (-DEPILOGUE_CLOCKS prints the counters of one epilogue warp).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o benchmarks/_build/bicubic_h_umma benchmarks/bicubic_h_umma.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(x)                                                                                  \
  do {                                                                                            \
    cudaError_t e_ = (x);                                                                         \
    if (e_ != cudaSuccess) {                                                                      \
      fprintf(stderr, "%s:%d %s: %s\n", __FILE__, __LINE__, #x, cudaGetErrorString(e_));          \
      exit(2);                                                                                    \
    }                                                                                             \
  } while (0)

namespace {

constexpr int kSrcW = 1920, kDstW = 640, kRowBytes = kSrcW * 3, kOutRowBytes = kDstW * 3;
constexpr int kRowsPerBlock = 128;                    // M
constexpr int kTilePixels = 16, kTileN = 48;          // N = 16 pixels x 3 channels
constexpr int kTileK = 224;                           // bytes of a tile's window (32-byte aligned start: seven K = 32 steps)
#ifndef TILES_PER_ITEM
#define TILES_PER_ITEM 4                          /* even */
#endif
#ifndef STAGES
#define STAGES 2
#endif
constexpr int kTilesPerItem = TILES_PER_ITEM, kItemsPerRow = kDstW / kTilePixels / kTilesPerItem;
constexpr int kStages = STAGES;
constexpr int kStripeAtoms = (144 * kTilesPerItem + 64 + 127) / 128;      // 128-byte swizzle atoms in a staged stripe
constexpr int kAtomBytes = kRowsPerBlock * 128;
constexpr int kStageBytes = kStripeAtoms * kAtomBytes;
constexpr int kMmaN = 3 * kTileN;                         // the three byte planes of the weights side by side: one MMA
constexpr int kSetBytes = kMmaN * kTileK;                 // 32,256; two sets: even tiles, odd tiles
constexpr int kEpilogueWarps = 24, kLoaderWarps = 1, kThreads = (kEpilogueWarps + 1 + kLoaderWarps) * 32;
constexpr int kSlotColumns = kMmaN;                       // 144 accumulator columns per tile; three slots
constexpr size_t kSmemBytes = kStages * (size_t)kStageBytes + 2 * (size_t)kSetBytes + 1024 + 1024;
constexpr int kSlack = 256;                           // readable bytes before and after the frames (window overhang)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffffu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;
  return d;
}

// K-major, 128-byte swizzle (what a TMA copy with CU_TENSOR_MAP_SWIZZLE_128B writes): rows of 128 bytes, 8-row groups
// 1024 bytes apart; the start address may sit 32 / 64 / 96 bytes into the atom.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t addr) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffffu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= 1ull << 46;
  d |= 2ull << 61;
  return d;
}

// kind::i8, D = s32, A = u8, B = s8, K-major both, M = 128, N = 144.
__host__ __device__ constexpr uint32_t umma_idesc() {
  return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(kMmaN >> 3) << 17) | ((uint32_t)(kRowsPerBlock >> 4) << 24);
}

template <int ACCUMULATE>
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "n"(ACCUMULATE), "r"(0u)
      : "memory");
}

__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\telect.sync rx|px, 0xffffffff;\n\tselp.u32 %0, 1, 0, px;\n\t}" : "=r"(pred));
  return pred;
}

__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void bar_init(unsigned long long* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}

__device__ __forceinline__ void bar_arrive(unsigned long long* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}

// -> false when the wait was given up (about a second): the kernel then unwinds instead of hanging the GPU.
__device__ __forceinline__ bool bar_wait(unsigned long long* bar, uint32_t parity, int* gave_up) {
  uint32_t done = 0, polls = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    if (done) return true;
    if ((++polls & 1023u) == 0) {
      if (*(volatile int*)gave_up) return false;
      if (clock64() - t0 > 2000000000ll) { atomicExch(gave_up, 1); return false; }
    }
  }
}

struct EdgeTaps {                                       // output pixels 0, 1, 638, 639: clamped windows
  int xmin[4], count[4], k[4][13];
};

// (sat_u8(hi) << 8 | sat_u8(lo)) | c << 16: clip to 0..255 and pack, two values per instruction
__device__ __forceinline__ uint32_t pack_sat_u8(int hi, int lo, uint32_t c) {
  uint32_t d;
  asm("cvt.pack.sat.u8.s32.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(hi), "r"(lo), "r"(c));
  return d;
}

// All 13 taps, the ones past the window with weight 0 (they may read the first bytes of the next row): 13 independent
// loads instead of a dependent chain.
__device__ __forceinline__ int edge_pixel(const uint8_t* __restrict__ row, const EdgeTaps& edge, int which, int c) {
  const uint8_t* px = row + (size_t)edge.xmin[which] * 3 + c;
  int taps[13];
#pragma unroll
  for (int j = 0; j < 13; j++) taps[j] = (int)px[3 * j];
  int ss = 1 << 21;
#pragma unroll
  for (int j = 0; j < 13; j++) ss += taps[j] * edge.k[which][j];
  return ss >> 22;
}

constexpr int kSlots = 3;                                // accumulator slots in tensor memory (3 x 144 of 512 columns)

struct Barriers {
  unsigned long long data_full[kStages], data_empty[kStages], acc_full[kSlots], acc_empty[kSlots];
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(kThreads, 1)
bicubic_h_umma_kernel(const __grid_constant__ CUtensorMap src_map,   // (5760 bytes, rows) over the frames, box 128 x 128, 128-byte swizzle
                      const uint8_t* __restrict__ src,
                      const uint8_t* __restrict__ weights,  // [3 planes][12 chunks][48][16]: the B operand as staged
                      uint8_t* __restrict__ dst, int n_row_blocks, const EdgeTaps edge, int* __restrict__ gave_up,
                      long long* __restrict__ stats) {      // per CTA: clocks the issuer spent in total / waiting for data / waiting for a slot
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);      // swizzle atoms: 1024-byte aligned
  uint8_t* stage0 = smem;
  uint8_t* wsm = smem + kStages * kStageBytes;
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kStages * kStageBytes + 2 * kSetBytes);

  const int tid = (int)threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 2 * kSetBytes / 16; i += kThreads)
    reinterpret_cast<uint4*>(wsm)[i] = reinterpret_cast<const uint4*>(weights)[i];
  if (tid == 0) {
    for (int s = 0; s < kStages; s++) {
      bar_init(&bars->data_full[s], 1);
      bar_init(&bars->data_empty[s], 1);
    }
    for (int s = 0; s < kSlots; s++) {
      bar_init(&bars->acc_full[s], 1);
      bar_init(&bars->acc_empty[s], kEpilogueWarps / 2);
    }
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == kEpilogueWarps) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&bars->tmem_base)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem = bars->tmem_base;
  const int total_items = n_row_blocks * kItemsPerRow;

  if (warp < kEpilogueWarps) {
    // ---- epilogue: TMEM -> combine the byte planes -> 16 output bytes per lane per tile
    // two groups of 12 warps taking alternate tiles: a warp has two tile periods for the arithmetic of one tile
    const int quarter = warp & 3, third = (warp % 12) >> 2, group = warp / 12;
    const int my_items = ((int)blockIdx.x < total_items) ? (total_items - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
    const uint32_t lane_base = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(third * 16);
    for (int t = group; t < my_items * kTilesPerItem; t += 2) {          // t: tiles in this CTA's order
      const int item = (int)blockIdx.x + (t / kTilesPerItem) * (int)gridDim.x, q = t % kTilesPerItem;
      const int rb = item / kItemsPerRow, tile = (item % kItemsPerRow) * kTilesPerItem + q;
      const size_t row = (size_t)rb * kRowsPerBlock + quarter * 32 + lane;
      const int slot = t % kSlots;
#ifdef EPILOGUE_CLOCKS
      const long long e0 = clock64();
#endif
      if (!bar_wait(&bars->acc_full[slot], (t / kSlots) & 1, gave_up)) goto out;
#ifdef EPILOGUE_CLOCKS
      const long long e1 = clock64();
#endif
      asm volatile("tcgen05.fence::after_thread_sync;");
      uint32_t packed[4];
#pragma unroll
      for (int half = 0; half < 2; half++) {                               // eight columns at a time: 24 registers
        uint32_t d[3][8];
#pragma unroll
        for (int p = 0; p < 3; p++)
          asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                       : "=r"(d[p][0]), "=r"(d[p][1]), "=r"(d[p][2]), "=r"(d[p][3]), "=r"(d[p][4]), "=r"(d[p][5]), "=r"(d[p][6]), "=r"(d[p][7])
                       : "r"(lane_base + (uint32_t)(slot * kSlotColumns + p * kTileN + half * 8)));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        if (half == 1) {
          asm volatile("tcgen05.fence::before_thread_sync;");
          __syncwarp();
          if (lane == 0) bar_arrive(&bars->acc_empty[slot]);               // the accumulators are in registers: the slot is free
        }
        int v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = ((int)d[0][i] + ((int)d[1][i] << 8) + ((int)d[2][i] << 16) + (1 << 21)) >> 22;
        if (half == 0 && tile == 0 && third == 0) {
          // the two pixels at this end of the row: clamped window, renormalised taps (the window's bytes were loaded a
          // moment ago: L2 hits)
#pragma unroll
          for (int e = 0; e < 6; e++) v[e] = edge_pixel(src + row * kRowBytes, edge, e / 3, e % 3);
        } else if (half == 1 && tile == kDstW / kTilePixels - 1 && third == 2) {
#pragma unroll
          for (int e = 0; e < 6; e++) v[2 + e] = edge_pixel(src + row * kRowBytes, edge, 2 + e / 3, e % 3);
        }
        packed[2 * half] = pack_sat_u8(v[1], v[0], pack_sat_u8(v[3], v[2], 0u));
        packed[2 * half + 1] = pack_sat_u8(v[5], v[4], pack_sat_u8(v[7], v[6], 0u));
      }
      uint8_t* out = dst + row * kOutRowBytes + (size_t)tile * kTileN + third * 16;
      *reinterpret_cast<uint4*>(out) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
#ifdef EPILOGUE_CLOCKS
      if (stats && (warp == 5 || warp == 17) && lane == 0) {      // one warp of each group
        const long long e2 = clock64();
        atomicAdd((unsigned long long*)&stats[blockIdx.x * 8 + 3], (unsigned long long)(e1 - e0));
        atomicAdd((unsigned long long*)&stats[blockIdx.x * 8 + 4], (unsigned long long)(e2 - e1));
        atomicAdd((unsigned long long*)&stats[blockIdx.x * 8 + 5], 1ull);
      }
#endif
    }
  } else if (warp == kEpilogueWarps) {
    // ---- MMA issuer: the whole warp walks the loop (uniform control flow, so the descriptors are computed on the
    // uniform datapath); one elected lane issues.  Everything but the stage / slot base is a compile-time constant.
    const uint64_t b_base = umma_desc(smem_u32(wsm), kMmaN * 16, 128);
    int n = 0, t = 0;
    long long w_data = 0, w_slot = 0;
    const long long t_begin = clock64();
    for (int item = (int)blockIdx.x; item < total_items; item += (int)gridDim.x, n++) {
      const int stage = n % kStages;
      long long c0 = clock64();
      if (!bar_wait(&bars->data_full[stage], (n / kStages) & 1, gave_up)) goto out;
      w_data += clock64() - c0;
      asm volatile("tcgen05.fence::after_thread_sync;");
      const uint64_t a_base = umma_desc_sw128(smem_u32(stage0 + (size_t)stage * kStageBytes));
#pragma unroll
      for (int q = 0; q < kTilesPerItem; q++, t++) {
        const int slot = t % kSlots;
        c0 = clock64();
        if (t >= kSlots && !bar_wait(&bars->acc_empty[slot], ((t / kSlots) - 1) & 1, gave_up)) goto out;
        w_slot += clock64() - c0;
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t d_slot = tmem + (uint32_t)(slot * kSlotColumns);
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < kTileK / 32; k++) {
            // the tile's window starts 144 * (q & ~1) + 128 * (q & 1) bytes into the stripe: a multiple of 32, so a
            // K = 32 step never straddles two swizzle atoms
            const uint32_t o = (uint32_t)(144 * (q & ~1) + 128 * (q & 1) + 32 * k);
            const uint64_t da = a_base + (uint64_t)(((o >> 7) * kAtomBytes + (o & 127u)) >> 4);
            const uint64_t db = b_base + (uint64_t)(((q & 1) * kSetBytes + (2 * k) * (kMmaN * 16)) >> 4);
            if (k == 0) umma_i8<0>(d_slot, da, db, umma_idesc());
            else umma_i8<1>(d_slot, da, db, umma_idesc());
          }
          umma_commit(&bars->acc_full[slot]);
        }
        __syncwarp();
      }
      if (elect_one()) umma_commit(&bars->data_empty[stage]);  // arrives when the MMAs that read this stage are done
      __syncwarp();
    }
    if (stats && lane == 0) {
      stats[blockIdx.x * 8 + 0] = clock64() - t_begin;
      stats[blockIdx.x * 8 + 1] = w_data;
      stats[blockIdx.x * 8 + 2] = w_slot;
    }
  } else {
    // ---- loader: one thread, one tensor copy per item
    if (lane == 0) {
      int n = 0;
      for (int item = (int)blockIdx.x; item < total_items; item += (int)gridDim.x, n++) {
        const int stage = n % kStages;
        const int rb = item / kItemsPerRow, g = item % kItemsPerRow;
        if (n >= kStages && !bar_wait(&bars->data_empty[stage], ((n / kStages) - 1) & 1, gave_up)) break;
        const uint32_t bar = smem_u32(&bars->data_full[stage]);
#ifdef EXPERIMENT_LOAD_ONCE
        if (n >= kStages) { bar_arrive(&bars->data_full[stage]); continue; }       // timing experiment: stale data
#endif
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"((uint32_t)kStageBytes) : "memory");
        // bytes before the row's first and after its last are outside the tensor: filled with zeros
        for (int a = 0; a < kStripeAtoms; a++)
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::
                           "r"(smem_u32(stage0 + (size_t)stage * kStageBytes + (size_t)a * kAtomBytes)), "l"(&src_map),
                       "r"(144 * kTilesPerItem * g - 32 + 128 * a), "r"(rb * kRowsPerBlock), "r"(bar)
                       : "memory");
      }
    }
  }
out:
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == kEpilogueWarps) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem));
}

// Pillow's precompute_coeffs + normalize_coeffs_8bpc for the bicubic filter (libImaging/Resample.c).
double bicubic_filter(double x) {
  const double a = -0.5;
  if (x < 0.0) x = -x;
  if (x < 1.0) return ((a + 2.0) * x - (a + 3.0)) * x * x + 1;
  if (x < 2.0) return (((x - 5) * x + 8) * x - 4) * a;
  return 0.0;
}

int precompute(int in_size, int out_size, int** bounds_out, int** kk_out) {
  double scale = (double)in_size / out_size, filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = 2.0 * filterscale;
  const int ksize = (int)ceil(support) * 2 + 1;
  int* bounds = (int*)calloc((size_t)out_size * 2, sizeof(int));
  int* kk = (int*)calloc((size_t)out_size * ksize, sizeof(int));
  double* k = (double*)malloc(sizeof(double) * ksize);
  for (int xx = 0; xx < out_size; xx++) {
    const double center = (xx + 0.5) * scale, ss = 1.0 / filterscale;
    double ww = 0.0;
    int xmin = (int)(center - support + 0.5);
    if (xmin < 0) xmin = 0;
    int xmax = (int)(center + support + 0.5);
    if (xmax > in_size) xmax = in_size;
    xmax -= xmin;
    for (int x = 0; x < xmax; x++) { k[x] = bicubic_filter((x + xmin - center + 0.5) * ss); ww += k[x]; }
    for (int x = 0; x < xmax; x++) {
      if (ww != 0.0) k[x] /= ww;
      kk[xx * ksize + x] = k[x] < 0 ? (int)(-0.5 + k[x] * (1 << 22)) : (int)(0.5 + k[x] * (1 << 22));
    }
    bounds[2 * xx] = xmin;
    bounds[2 * xx + 1] = xmax;
  }
  free(k);
  *bounds_out = bounds;
  *kk_out = kk;
  return ksize;
}

}  // namespace

int main(int argc, char** argv) {
  const int n_frames = argc > 1 ? atoi(argv[1]) : 32, src_h = 1080;
  const long long n_rows = (long long)n_frames * src_h;
  if (n_rows % kRowsPerBlock) { fprintf(stderr, "frames x 1080 must be a multiple of 128 (use a multiple of 8 frames)\n"); return 2; }
  const int n_row_blocks = (int)(n_rows / kRowsPerBlock);
  cudaDeviceProp prop;
  CHECK(cudaGetDeviceProperties(&prop, 0));

  int *bounds, *kk;
  const int ksize = precompute(kSrcW, kDstW, &bounds, &kk);
  // the B operand of an interior tile (tile 1: no clamping anywhere), in the staged chunk order
  uint8_t* wplanes = (uint8_t*)calloc(2, kSetBytes);
  for (int xl = 0; xl < kTilePixels; xl++) {
    const int x = kTilePixels + xl;
    if (bounds[2 * x] != 3 * x - 4 || bounds[2 * x + 1] != 12) { fprintf(stderr, "unexpected window at %d\n", x); return 2; }
    for (int j = 0; j < 12; j++)
      for (int c = 0; c < 3; c++) {
        const int n = xl * 3 + c, w = kk[x * ksize + j];
        // balanced base-256 digits, all three signed: w = p0 + 256 p1 + 65536 p2
        int p[3];
        p[0] = ((w + 128) & 255) - 128;
        const int w1 = (w - p[0]) >> 8;
        p[1] = ((w1 + 128) & 255) - 128;
        p[2] = (w1 - p[1]) >> 8;
        if (p[2] < -128 || p[2] > 127 || p[0] + 256 * p[1] + 65536 * p[2] != w) { fprintf(stderr, "weight out of range\n"); return 2; }
        for (int odd = 0; odd < 2; odd++) {      // even tiles: window from byte -32 of the tile; odd tiles: from byte -48
          const int k = 9 * xl + 3 * j + c - 12 + (odd ? 48 : 32);
          for (int pl = 0; pl < 3; pl++) wplanes[odd * kSetBytes + (k >> 4) * (kMmaN * 16) + (pl * kTileN + n) * 16 + (k & 15)] = (uint8_t)p[pl];
        }
      }
  }
  for (int x = 2; x < kDstW - 2; x++)            // every interior pixel has the interior tile's taps
    for (int j = 0; j < 12; j++)
      if (bounds[2 * x] != 3 * x - 4 || bounds[2 * x + 1] != 12 || kk[x * ksize + j] != kk[(kTilePixels + (x % kTilePixels)) * ksize + j]) {
        fprintf(stderr, "pixel %d is not shift-invariant\n", x);
        return 2;
      }

  const size_t src_bytes = (size_t)n_rows * kRowBytes, dst_bytes = (size_t)n_rows * kOutRowBytes;
  uint8_t* h_src = (uint8_t*)malloc(src_bytes);
  uint32_t s = 2024u;
  for (size_t i = 0; i < src_bytes; i += 4) { s = s * 1664525u + 1013904223u; uint32_t v = s ^ (s >> 15); memcpy(h_src + i, &v, 4); }
  // a few saturating stretches: alternating 0 / 255 columns overshoot on both sides
  for (int x = 0; x < kSrcW; x++) { const uint8_t v = (x / 3) & 1 ? 255 : 0; for (int c = 0; c < 3; c++) h_src[(size_t)5 * kRowBytes + x * 3 + c] = v; }

  EdgeTaps edge;
  memset(&edge, 0, sizeof(edge));
  for (int w = 0; w < 4; w++) {
    const int x = w < 2 ? w : kDstW - 4 + w;
    edge.xmin[w] = bounds[2 * x];
    edge.count[w] = bounds[2 * x + 1];
    for (int j = 0; j < bounds[2 * x + 1]; j++) edge.k[w][j] = kk[x * ksize + j];
  }
  uint8_t *d_src_alloc, *d_dst, *d_w;
  int* d_flag;
  CHECK(cudaMalloc(&d_src_alloc, src_bytes + 2 * kSlack));
  CHECK(cudaMemset(d_src_alloc, 0x5a, src_bytes + 2 * kSlack));
  uint8_t* d_src = d_src_alloc + kSlack;
  CHECK(cudaMalloc(&d_dst, dst_bytes));
  CHECK(cudaMalloc(&d_w, 2 * kSetBytes));
  CHECK(cudaMalloc(&d_flag, sizeof(int)));
  CHECK(cudaMemset(d_flag, 0, sizeof(int)));
  CHECK(cudaMemcpy(d_src, h_src, src_bytes, cudaMemcpyHostToDevice));
  CHECK(cudaMemcpy(d_w, wplanes, 2 * kSetBytes, cudaMemcpyHostToDevice));
  CHECK(cudaMemset(d_dst, 0xee, dst_bytes));
  CHECK(cudaFuncSetAttribute(bicubic_h_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));

  // the frames as a 2-D tensor of bytes (row bytes, rows); a box = one 128-byte swizzle atom column of 128 rows
  typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess) { fprintf(stderr, "no cuTensorMapEncodeTiled\n"); return 2; }
  CUtensorMap src_map;
  const cuuint64_t dims[2] = {kRowBytes, (cuuint64_t)n_rows};
  const cuuint64_t strides[1] = {kRowBytes};
  const cuuint32_t box[2] = {128, kRowsPerBlock}, elem[2] = {1, 1};
  const CUresult mr = ((EncodeTiled)fn)(&src_map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_src, dims, strides, box, elem, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                        CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (mr != CUDA_SUCCESS) { printf("{\"kernel\": \"bicubic_h_umma\", \"error\": \"cuTensorMapEncodeTiled -> %d\"}\n", (int)mr); return 1; }
  const int grid = prop.multiProcessorCount;
  auto launch = [&]() { bicubic_h_umma_kernel<<<grid, kThreads, kSmemBytes>>>(src_map, d_src, d_w, d_dst, n_row_blocks, edge, d_flag, nullptr); };
  launch();
  CHECK(cudaDeviceSynchronize());
  int gave_up = 0;
  CHECK(cudaMemcpy(&gave_up, d_flag, sizeof(int), cudaMemcpyDeviceToHost));
  if (gave_up) { printf("{\"kernel\": \"bicubic_h_umma\", \"error\": \"a barrier wait was given up\"}\n"); return 1; }

  // parity: every 61st row and the first / last rows of the batch, all 1920 bytes, against Pillow's arithmetic
  uint8_t* h_dst = (uint8_t*)malloc(dst_bytes);
  CHECK(cudaMemcpy(h_dst, d_dst, dst_bytes, cudaMemcpyDeviceToHost));
  long long checked = 0, wrong = 0, first_wrong = -1;
  for (long long row = 0; row < n_rows; row += (row < 8 || row >= n_rows - 8) ? 1 : 61) {
    for (int x = 0; x < kDstW; x++)
      for (int c = 0; c < 3; c++) {
        int ss = 1 << 21;
        for (int j = 0; j < bounds[2 * x + 1]; j++) ss += (int)h_src[row * kRowBytes + (size_t)(bounds[2 * x] + j) * 3 + c] * kk[x * ksize + j];
        int v = ss >> 22;
        v = v < 0 ? 0 : v > 255 ? 255 : v;
        checked++;
        if (h_dst[row * kOutRowBytes + x * 3 + c] != (uint8_t)v) { if (first_wrong < 0) first_wrong = row * kOutRowBytes + x * 3 + c; wrong++; }
      }
  }

  // time: a second set of buffers would not change anything (199 MB in + 66 MB out per launch > L2 126 MB)
  cudaEvent_t e0, e1;
  CHECK(cudaEventCreate(&e0));
  CHECK(cudaEventCreate(&e1));
  for (int i = 0; i < 3; i++) launch();
  CHECK(cudaDeviceSynchronize());
  const int reps = 20;
  float main_ms = 0;
  CHECK(cudaEventRecord(e0));
  for (int i = 0; i < reps; i++) launch();                  // back to back: 265 MB per launch, nothing survives in L2
  CHECK(cudaEventRecord(e1));
  CHECK(cudaDeviceSynchronize());
  CHECK(cudaEventElapsedTime(&main_ms, e0, e1));
  main_ms /= reps;
  CHECK(cudaMemcpy(&gave_up, d_flag, sizeof(int), cudaMemcpyDeviceToHost));
  {   // one more launch with the clock counters on
    long long* d_stats;
    CHECK(cudaMalloc(&d_stats, sizeof(long long) * 8 * grid));
    CHECK(cudaMemset(d_stats, 0, sizeof(long long) * 8 * grid));
    bicubic_h_umma_kernel<<<grid, kThreads, kSmemBytes>>>(src_map, d_src, d_w, d_dst, n_row_blocks, edge, d_flag, d_stats);
    CHECK(cudaDeviceSynchronize());
    long long* h = (long long*)malloc(sizeof(long long) * 8 * grid);
    CHECK(cudaMemcpy(h, d_stats, sizeof(long long) * 8 * grid, cudaMemcpyDeviceToHost));
    double a[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < grid; i++) for (int j = 0; j < 6; j++) a[j] += (double)h[i * 8 + j] / grid;
    long long worst = 0;
    for (int i = 0; i < grid; i++) worst = h[i * 8] > worst ? h[i * 8] : worst;
    fprintf(stderr, "issuer clocks per CTA: total %.0f (slowest CTA %lld), waiting for data %.0f, waiting for a slot %.0f\n", a[0], worst, a[1], a[2]);
    if (a[5] > 0) fprintf(stderr, "an epilogue warp, per tile of its own: waiting for the accumulator %.0f clocks, TMEM read + arithmetic + store issue %.0f\n", a[3] / a[5], a[4] / a[5]);
    free(h);
  }
  const double bytes = (double)src_bytes + (double)dst_bytes;
  printf("{\"kernel\": \"bicubic_h_umma (tcgen05.mma kind::i8, M=128 N=144 K=32)\", \"frames\": %d, \"rows\": %lld, \"checked_bytes\": %lld, "
         "\"wrong_bytes\": %lld, \"first_wrong_offset\": %lld, \"gave_up\": %d, \"tiles_per_item\": %d, \"stages\": %d, \"kernel_us\": %.2f, "
         "\"algorithmic_mb\": %.1f, \"gb_per_s\": %.0f}\n",
         n_frames, n_rows, checked, wrong, first_wrong, gave_up, kTilesPerItem, kStages, main_ms * 1e3, bytes / 1e6, bytes / (main_ms * 1e-3) / 1e9);
  return wrong || gave_up ? 1 : 0;
}
