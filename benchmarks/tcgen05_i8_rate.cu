// Measurement behind DESIGN.md's "tcgen05 for the bicubic resize" paragraph: what one SM's 5th-generation tensor
// core delivers on 8-bit integer tiles (tcgen05.mma kind::i8, M = 64 and M = 128, accumulators in TMEM, operands
// read from shared memory through UMMA descriptors), next to what the legacy warp-level path delivers
// (mma.sync.m16n8k32, the instruction csrc/resize_bicubic_mma.cu is built from).  The bicubic weight matrices
// are banded, so the comparison that decides the kernel's shape is
//     dense rate (tcgen05) x band fill at M = 64 / 128      vs      dense rate (mma.sync) x band fill at M = 16.
// The M = 128 tile is also CHECKED against a host product, so the descriptors below are known to address what they
// claim to.  Stand-alone program, no part of the library:
//     nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o benchmarks/_build/tcgen05_i8_rate benchmarks/tcgen05_i8_rate.cu
//     benchmarks/_build/tcgen05_i8_rate
#include <cuda_runtime.h>
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

constexpr int kTileK = 128;                   // bytes of K held in shared memory: four K = 32 instructions

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// Shared-memory matrix descriptor, K-major, no swizzle ("interleave"): core matrices of 8 rows x 16 bytes stored
// contiguously (128 bytes); SBO = bytes between 8-row groups, LBO = bytes between the 16-byte K chunks.
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3ffffu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
  d |= 1ull << 46;                            // descriptor version of sm_100
  return d;                                   // base offset 0, LBO mode 0, layout type 0 = no swizzle
}

// Instruction descriptor of kind::i8: D = s32, A = u8, B = s8, both K-major, dense, no saturation.
__host__ __device__ constexpr uint32_t umma_idesc(int M, int N) {
  return (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t"
      "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate), "r"(0u)
      : "memory");
}

// One CTA per SM.  The tile's operands sit in shared memory in the canonical no-swizzle layout; one thread issues
// `iters` rounds of the four K = 32 instructions into one TMEM accumulator, commits, and waits on the mbarrier.
template <int M, int N>
__global__ void __launch_bounds__(128) umma_i8_kernel(const uint8_t* __restrict__ A, const int8_t* __restrict__ B,
                                                      int32_t* __restrict__ D, int iters, long long* __restrict__ clocks,
                                                      int* __restrict__ timed_out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;                                   // [kTileK / 16][M / 8][8][16]
  uint8_t* sb = smem + M * kTileK;                      // [kTileK / 16][N / 8][8][16]
  __shared__ __align__(8) unsigned long long bar;
  __shared__ uint32_t tmem_base_slot;

  const int tid = (int)threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < M * kTileK; i += 128) {
    const int m = i / kTileK, k = i % kTileK;
    sa[(k >> 4) * (M * 16) + (m >> 3) * 128 + (m & 7) * 16 + (k & 15)] = A[i];
  }
  for (int i = tid; i < N * kTileK; i += 128) {
    const int n = i / kTileK, k = i % kTileK;
    sb[(k >> 4) * (N * 16) + (n >> 3) * 128 + (n & 7) * 16 + (k & 15)] = (uint8_t)B[i];
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 256;" ::"r"(smem_u32(&tmem_base_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the generic-proxy stores above -> the MMA's reads
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem_d = tmem_base_slot;

  long long t0 = 0, t1 = 0;
  int gave_up = 0;
  if (tid == 0) {
    const uint32_t idesc = umma_idesc(M, N);
    const uint32_t a0 = smem_u32(sa), b0 = smem_u32(sb);
    t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
      for (int k = 0; k < kTileK / 32; k++) {
        // K advances by 32 bytes = two 16-byte chunks = 2 x LBO
        const uint64_t da = umma_desc(a0 + k * 2 * (M * 16), M * 16, 128);
        const uint64_t db = umma_desc(b0 + k * 2 * (N * 16), N * 16, 128);
        umma_i8(tmem_d, da, db, idesc, (it | k) ? 1u : 0u);
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    uint32_t done = 0;
    const long long give_up_at = t0 + 4000000000ll;                  // ~2 s: report instead of hanging the box
    while (!done) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}\n"
          : "=r"(done)
          : "r"(smem_u32(&bar)), "r"(0u)
          : "memory");
      if (!done && clock64() > give_up_at) { gave_up = 1; break; }
    }
    t1 = clock64();
    if (gave_up) atomicExch(timed_out, 1);
    if (clocks) clocks[blockIdx.x] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");

  // Accumulator -> global memory (block 0 only; M = 128: row m of D is TMEM lane m, column n is TMEM column n;
  // warp w may read lanes 32 w .. 32 w + 31).  M = 64 is timed, not read back.
  if (D != nullptr && blockIdx.x == 0 && M == 128) {
    for (int c = 0; c < N; c += 8) {
      uint32_t v[8];
      const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c;
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                   : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                   : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int j = 0; j < 8; j++) D[(size_t)(warp * 32 + lane) * N + c + j] = (int32_t)v[j];
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 256;" ::"r"(tmem_d));
}

// The issue pattern of benchmarks/bicubic_h_umma.cu without its loads and epilogue: M = 128, N = 48, seven K = 32 steps
// into each of ACCS accumulators in turn (B from a different plane each), optionally while the other three warps
// read a second TMEM region with tcgen05.ld all the time (READERS).
template <int ACCS, int READERS, int COMMIT_EACH_ROUND>
__global__ void __launch_bounds__(128) umma_pattern_kernel(int iters, long long* __restrict__ clocks, int* __restrict__ sink) {
  constexpr int M = 128, N = 48, K = 224;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sa = smem;
  uint8_t* sb = smem + M * K;                           // ACCS planes of N x K
  __shared__ __align__(8) unsigned long long bar, round_bar;
  __shared__ uint32_t tmem_base_slot;
  __shared__ volatile int stop;
  const int tid = (int)threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < (M + ACCS * N) * K; i += 128) smem[i] = (uint8_t)(i * 37 + 11);
  if (tid == 0) {
    stop = 0;
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&round_bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_base_slot)));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;");
  const uint32_t tmem_d = tmem_base_slot;
  if (warp == 0) {
    if (tid == 0) {
      const uint32_t idesc = umma_idesc(M, N);
      const uint32_t a0 = smem_u32(sa), b0 = smem_u32(sb);
      const long long t0 = clock64();
      for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int p = 0; p < ACCS; p++) {
#pragma unroll
          for (int k = 0; k < K / 32; k++) {
            const uint64_t da = umma_desc(a0 + k * 2 * (M * 16), M * 16, 128);
            const uint64_t db = umma_desc(b0 + p * N * K + k * 2 * (N * 16), N * 16, 128);
            umma_i8(tmem_d + p * 64, da, db, idesc, k ? 1u : 0u);
          }
        }
        if (COMMIT_EACH_ROUND)      // as a kernel with an epilogue does once per tile; nobody waits on this barrier
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&round_bar)) : "memory");
      }
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
      uint32_t done = 0;
      while (!done && clock64() - t0 < 4000000000ll) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}\n"
            : "=r"(done)
            : "r"(smem_u32(&bar)), "r"(0u)
            : "memory");
      }
      clocks[blockIdx.x] = clock64() - t0;
      stop = 1;
    }
  } else if (READERS) {
    uint32_t acc = 0;
    while (!stop) {
      uint32_t v[16];
      const uint32_t taddr = tmem_d + ((uint32_t)(warp * 32) << 16) + 256u;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
            "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      for (int j = 0; j < 16; j++) acc += v[j];
    }
    if (acc == 0x12345u) sink[0] = (int)acc;
  }
  asm volatile("tcgen05.fence::before_thread_sync;");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem_d));
}

template <int ACCS, int READERS, int COMMIT_EACH_ROUND>
void run_pattern(int n_sms, long long* d_clocks, int* d_sink) {
  const size_t smem = (size_t)(128 + ACCS * 48) * 224;
  CHECK(cudaFuncSetAttribute(umma_pattern_kernel<ACCS, READERS, COMMIT_EACH_ROUND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int iters = 512;
  for (int rep = 0; rep < 2; rep++) {
    umma_pattern_kernel<ACCS, READERS, COMMIT_EACH_ROUND><<<n_sms, 128, smem>>>(iters, d_clocks, d_sink);
    CHECK(cudaDeviceSynchronize());
  }
  long long* h = (long long*)malloc(sizeof(long long) * n_sms);
  CHECK(cudaMemcpy(h, d_clocks, sizeof(long long) * n_sms, cudaMemcpyDeviceToHost));
  long long worst = 0;
  for (int i = 0; i < n_sms; i++) worst = h[i] > worst ? h[i] : worst;
  free(h);
  printf("{\"kernel\": \"tcgen05.mma kind::i8 pattern\", \"M\": 128, \"N\": 48, \"accumulators_in_turn\": %d, \"tmem_readers\": %d, \"commit_after_each_round\": %d, "
         "\"clocks_per_instruction\": %.2f}\n", ACCS, READERS, COMMIT_EACH_ROUND, (double)worst / (iters * ACCS * 7));
}

// The legacy path: every warp keeps ACC independent m16n8k32 accumulators going.
template <int ACC>
__global__ void __launch_bounds__(256) imma_rate_kernel(int iters, int* __restrict__ sink) {
  int acc[ACC][4];
#pragma unroll
  for (int i = 0; i < ACC; i++) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
  const uint32_t a0 = threadIdx.x * 0x01010101u, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3;
  const uint32_t b0 = blockIdx.x * 0x01000193u, b1 = b0 ^ 0x5bd1e995u;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ACC; i++)
      asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                   : "+r"(acc[i][0]), "+r"(acc[i][1]), "+r"(acc[i][2]), "+r"(acc[i][3])
                   : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  int s = 0;
#pragma unroll
  for (int i = 0; i < ACC; i++) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
  if (s == 0x7fffffff) sink[0] = s;
}

float time_ms(cudaEvent_t a, cudaEvent_t b) {
  float ms = 0;
  CHECK(cudaEventElapsedTime(&ms, a, b));
  return ms;
}

template <int M, int N>
int run_umma(int n_sms, double sm_hz, const uint8_t* dA, const int8_t* dB, int32_t* dD, long long* d_clocks, int* d_timed_out,
             const uint8_t* hA, const int8_t* hB) {
  const size_t smem = (size_t)(M + N) * kTileK;
  CHECK(cudaFuncSetAttribute(umma_i8_kernel<M, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cudaEvent_t e0, e1;
  CHECK(cudaEventCreate(&e0));
  CHECK(cudaEventCreate(&e1));
  int timed_out = 0;
  // correctness first (one round over K = 128), on one CTA
  CHECK(cudaMemset(dD, 0xff, (size_t)128 * 256 * 4));
  umma_i8_kernel<M, N><<<1, 128, smem>>>(dA, dB, dD, 1, nullptr, d_timed_out);
  CHECK(cudaDeviceSynchronize());
  CHECK(cudaMemcpy(&timed_out, d_timed_out, 4, cudaMemcpyDeviceToHost));
  if (timed_out) { printf("{\"kernel\": \"tcgen05.mma kind::i8\", \"M\": %d, \"N\": %d, \"error\": \"mbarrier never completed\"}\n", M, N); return 1; }
  long long wrong = -1;
  if (M == 128) {
    int32_t* hD = (int32_t*)malloc((size_t)M * N * 4);
    CHECK(cudaMemcpy(hD, dD, (size_t)M * N * 4, cudaMemcpyDeviceToHost));
    wrong = 0;
    for (int m = 0; m < M; m++)
      for (int n = 0; n < N; n++) {
        int32_t s = 0;
        for (int k = 0; k < kTileK; k++) s += (int32_t)hA[m * kTileK + k] * (int32_t)hB[n * kTileK + k];
        wrong += (s != hD[(size_t)m * N + n]);
      }
    free(hD);
  }
  const int iters = 4096;
  umma_i8_kernel<M, N><<<n_sms, 128, smem>>>(dA, dB, nullptr, iters, d_clocks, d_timed_out);   // warm-up
  CHECK(cudaDeviceSynchronize());
  CHECK(cudaEventRecord(e0));
  umma_i8_kernel<M, N><<<n_sms, 128, smem>>>(dA, dB, nullptr, iters, d_clocks, d_timed_out);
  CHECK(cudaEventRecord(e1));
  CHECK(cudaDeviceSynchronize());
  CHECK(cudaMemcpy(&timed_out, d_timed_out, 4, cudaMemcpyDeviceToHost));
  long long* h_clocks = (long long*)malloc(sizeof(long long) * n_sms);
  CHECK(cudaMemcpy(h_clocks, d_clocks, sizeof(long long) * n_sms, cudaMemcpyDeviceToHost));
  long long worst = 0;
  for (int i = 0; i < n_sms; i++) worst = h_clocks[i] > worst ? h_clocks[i] : worst;
  free(h_clocks);
  const double macs_per_cta = (double)iters * (kTileK / 32) * M * N * 32.0;
  const float ms = time_ms(e0, e1);
  printf("{\"kernel\": \"tcgen05.mma kind::i8\", \"M\": %d, \"N\": %d, \"K_per_instruction\": 32, \"instructions_per_cta\": %d, "
         "\"wrong_elements_vs_host\": %lld, \"timed_out\": %d, \"mac_per_clk_per_sm\": %.1f, \"clocks_per_instruction\": %.2f, "
         "\"kernel_ms\": %.4f, \"chip_tmac_per_s\": %.1f, \"chip_tmac_per_s_from_clocks\": %.1f}\n",
         M, N, iters * (kTileK / 32), wrong, timed_out, macs_per_cta / (double)worst, (double)worst / (iters * (kTileK / 32)), ms,
         macs_per_cta * n_sms / (ms * 1e-3) / 1e12, macs_per_cta / (double)worst * sm_hz * n_sms / 1e12);
  return (wrong > 0 || timed_out) ? 1 : 0;
}

template <int ACC>
void run_imma(int n_sms, int blocks_per_sm, int* d_sink) {
  cudaEvent_t e0, e1;
  CHECK(cudaEventCreate(&e0));
  CHECK(cudaEventCreate(&e1));
  const int iters = 20000;
  imma_rate_kernel<ACC><<<n_sms * blocks_per_sm, 256>>>(iters, d_sink);
  CHECK(cudaDeviceSynchronize());
  CHECK(cudaEventRecord(e0));
  imma_rate_kernel<ACC><<<n_sms * blocks_per_sm, 256>>>(iters, d_sink);
  CHECK(cudaEventRecord(e1));
  CHECK(cudaDeviceSynchronize());
  const float ms = time_ms(e0, e1);
  const double macs = (double)n_sms * blocks_per_sm * 8 * ACC * (double)iters * 16 * 8 * 32;
  printf("{\"kernel\": \"mma.sync.m16n8k32 u8 x s8\", \"warps_per_sm\": %d, \"independent_accumulators_per_warp\": %d, "
         "\"kernel_ms\": %.4f, \"chip_tmac_per_s\": %.1f}\n",
         blocks_per_sm * 8, ACC, ms, macs / (ms * 1e-3) / 1e12);
}

}  // namespace

int main() {
  cudaDeviceProp prop;
  CHECK(cudaGetDeviceProperties(&prop, 0));
  int khz = 0;
  CHECK(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0));
  const int n_sms = prop.multiProcessorCount;
  const double sm_hz = khz * 1e3;
  printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\", \"sm_clock_mhz_max\": %.0f}\n", prop.name, n_sms, prop.major, prop.minor, khz / 1e3);
  if (prop.major != 10) { fprintf(stderr, "needs sm_100\n"); return 2; }

  uint8_t* hA = (uint8_t*)malloc(128 * kTileK);
  int8_t* hB = (int8_t*)malloc(256 * kTileK);
  uint32_t s = 12345u;
  for (int i = 0; i < 128 * kTileK; i++) { s = s * 1664525u + 1013904223u; hA[i] = (uint8_t)(s >> 24); }
  for (int i = 0; i < 256 * kTileK; i++) { s = s * 1664525u + 1013904223u; hB[i] = (int8_t)(s >> 24); }
  uint8_t* dA; int8_t* dB; int32_t* dD; long long* d_clocks; int* d_flag;
  CHECK(cudaMalloc(&dA, 128 * kTileK));
  CHECK(cudaMalloc(&dB, 256 * kTileK));
  CHECK(cudaMalloc(&dD, (size_t)128 * 256 * 4));
  CHECK(cudaMalloc(&d_clocks, sizeof(long long) * n_sms));
  CHECK(cudaMalloc(&d_flag, 2 * sizeof(int)));
  CHECK(cudaMemset(d_flag, 0, 2 * sizeof(int)));
  CHECK(cudaMemcpy(dA, hA, 128 * kTileK, cudaMemcpyHostToDevice));
  CHECK(cudaMemcpy(dB, hB, 256 * kTileK, cudaMemcpyHostToDevice));

  int bad = 0;
  bad |= run_umma<128, 256>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
  if (!bad) {                                  // a wrong or hung first shape: stop, nothing else is worth the box time
    bad |= run_umma<128, 128>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 64>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 192>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 96>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 48>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 32>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<128, 16>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<64, 256>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<64, 128>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
    bad |= run_umma<64, 64>(n_sms, sm_hz, dA, dB, dD, d_clocks, d_flag, hA, hB);
  }
  run_pattern<1, 0, 0>(n_sms, d_clocks, d_flag + 1);
  run_pattern<3, 0, 0>(n_sms, d_clocks, d_flag + 1);
  run_pattern<3, 1, 0>(n_sms, d_clocks, d_flag + 1);
  run_pattern<3, 0, 1>(n_sms, d_clocks, d_flag + 1);
  run_pattern<3, 1, 1>(n_sms, d_clocks, d_flag + 1);
  run_imma<8>(n_sms, 1, d_flag + 1);
  run_imma<8>(n_sms, 2, d_flag + 1);
  run_imma<8>(n_sms, 4, d_flag + 1);
  run_imma<4>(n_sms, 8, d_flag + 1);
  return bad;
}
