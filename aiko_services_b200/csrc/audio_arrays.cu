// PE_AudioFilter and PE_AudioResampler on arrays that did NOT come from this library's fused chain:
// the reference's elements take any (amplitudes, frequencies) pair of float64 arrays
// (R/elements/media/audio_io.py:213-216,250-251), e.g. produced by another element.  One CTA per
// window; float64 throughout, so the gate decisions are the reference's own.
//
//   audio_filter_f64 (audio_io.py:216-227): keep (|f|, a) with 0.1 <= a <= 12 and 10 <= |f| <= 9000,
//     stable sort by a descending, first 100.  The 100th largest amplitude is found by bisection on
//     the bit pattern of the (positive) doubles -- 63 counting passes over the window, which stays
//     in L2 -- then the at most 100 survivors are ranked against each other in shared memory; ties
//     keep their original order, as Python's stable sort does.
//   audio_bands_f64 (audio_io.py:253-277): first half of both arrays, 8 bands of width
//     (f[-1] - f[0]) / 8 / 10, per band the sum of the amplitudes whose frequency lies in
//     [start, end) and the band centre.
#include "common.cuh"

namespace aiko {

constexpr int kTop = 100;            // AF_SAMPLES_MAXIMUM, audio_io.py:206
constexpr int kBandCount = 8;        // AR_BAND_COUNT, audio_io.py:237

struct ArrayParams {
  const double* amplitudes;   // [windows][n]
  const double* frequencies;  // [windows][n], or one row shared by all windows (freq_stride 0)
  long long freq_stride;
  int n_windows, n;
  double* out_a;              // filter: [windows][100] amplitudes; bands: [windows][8] sums
  double* out_f;              // filter: [windows][100] |frequencies|; bands: [windows][8] centres
  int* out_count;             // filter: [windows]
};

__device__ __forceinline__ long long block_sum(long long v, long long* scratch) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
  if (lane == 0) scratch[warp] = v;
  __syncthreads();
  long long s = 0;
#pragma unroll
  for (int w = 0; w < kThreads / 32; w++) s += scratch[w];
  __syncthreads();
  return s;
}

__device__ __forceinline__ bool passes_gate(double a, double f) {
  const double af = fabs(f);
  return a >= 0.1 && af >= 10.0 && a <= 12.0 && af <= 9000.0;      // audio_io.py:202-205,217-220
}

__global__ void __launch_bounds__(kThreads) audio_filter_f64_kernel(const ArrayParams p) {
  __shared__ long long scratch[kThreads / 32];
  __shared__ double sel_a[kTop], sel_f[kTop];
  __shared__ int sel_i[kTop];
  __shared__ int warp_counts[kThreads / 32];
  __shared__ int appended, equal_taken;
  for (int wdw = blockIdx.x; wdw < p.n_windows; wdw += gridDim.x) {
    const double* A = p.amplitudes + (long long)wdw * p.n;
    const double* F = p.frequencies + (long long)wdw * p.freq_stride;
    // ---- how many pass the gate; the bit pattern T of the 100th largest among them (0: fewer than 100)
    long long mine = 0;
    for (int i = threadIdx.x; i < p.n; i += kThreads) mine += passes_gate(A[i], F[i]) ? 1 : 0;
    const long long eligible = block_sum(mine, scratch);
    unsigned long long T = 0;
    if (eligible > kTop) {
      for (int bit = 62; bit >= 0; bit--) {             // amplitudes are >= 0.1: sign bit clear, bits are monotonic
        const unsigned long long cand = T | (1ull << bit);
        long long c = 0;
        for (int i = threadIdx.x; i < p.n; i += kThreads) {
          const double a = A[i];
          c += (passes_gate(a, F[i]) && (unsigned long long)__double_as_longlong(a) >= cand) ? 1 : 0;
        }
        if (block_sum(c, scratch) >= kTop) T = cand;    // at least 100 values are >= cand: the 100th is too
      }
    }
    long long above = 0;
    for (int i = threadIdx.x; i < p.n; i += kThreads) {
      const double a = A[i];
      above += (passes_gate(a, F[i]) && (unsigned long long)__double_as_longlong(a) > T) ? 1 : 0;
    }
    const int n_above = (int)block_sum(above, scratch);
    const int want_equal = (eligible > kTop) ? kTop - n_above : 0;       // ties with the 100th value, in index order
    if (threadIdx.x == 0) { appended = 0; equal_taken = 0; }
    __syncthreads();
    // ---- gather the survivors in index order (chunks of 256 indices, ballot + prefix per chunk)
    for (int base = 0; base < p.n; base += kThreads) {
      const int i = base + threadIdx.x;
      bool gt = false, eq = false;
      double a = 0, f = 0;
      if (i < p.n) {
        a = A[i]; f = F[i];
        if (passes_gate(a, f)) {
          const unsigned long long key = (unsigned long long)__double_as_longlong(a);
          gt = (eligible <= kTop) || key > T;
          eq = (eligible > kTop) && key == T;
        }
      }
      // equal-to-threshold values are admitted in index order until want_equal are in
      const unsigned eq_ballot = __ballot_sync(0xffffffffu, eq);
      const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
      if (lane == 0) warp_counts[warp] = __popc(eq_ballot);
      __syncthreads();
      int eq_before = equal_taken;
      for (int w = 0; w < warp; w++) eq_before += warp_counts[w];
      eq_before += __popc(eq_ballot & ((1u << lane) - 1));
      const bool take = gt || (eq && eq_before < want_equal);
      int eq_total = 0;
      for (int w = 0; w < kThreads / 32; w++) eq_total += warp_counts[w];
      __syncthreads();
      const unsigned take_ballot = __ballot_sync(0xffffffffu, take);
      if (lane == 0) warp_counts[warp] = __popc(take_ballot);
      __syncthreads();
      int pos = appended;
      for (int w = 0; w < warp; w++) pos += warp_counts[w];
      pos += __popc(take_ballot & ((1u << lane) - 1));
      if (take && pos < kTop) { sel_a[pos] = a; sel_f[pos] = fabs(f); sel_i[pos] = i; }
      int take_total = 0;
      for (int w = 0; w < kThreads / 32; w++) take_total += warp_counts[w];
      __syncthreads();
      if (threadIdx.x == 0) { appended += take_total; equal_taken += eq_total; }
      __syncthreads();
    }
    const int count = min(appended, kTop);
    // ---- rank the survivors: larger amplitude first, equal amplitudes in original order
    if ((int)threadIdx.x < count) {
      const double a = sel_a[threadIdx.x];
      const int idx = sel_i[threadIdx.x];
      int rank = 0;
      for (int j = 0; j < count; j++) rank += (sel_a[j] > a || (sel_a[j] == a && sel_i[j] < idx)) ? 1 : 0;
      p.out_a[(long long)wdw * kTop + rank] = a;
      p.out_f[(long long)wdw * kTop + rank] = sel_f[threadIdx.x];
    }
    if (threadIdx.x == 0) p.out_count[wdw] = count;
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kThreads) audio_bands_f64_kernel(const ArrayParams p) {
  __shared__ double warp_sums[kThreads / 32][kBandCount];
  const int half = p.n / 2;
  for (int wdw = blockIdx.x; wdw < p.n_windows; wdw += gridDim.x) {
    const double* A = p.amplitudes + (long long)wdw * p.n;
    const double* F = p.frequencies + (long long)wdw * p.freq_stride;
    const double frequency_range = F[half - 1] - F[0];                   // audio_io.py:256
    const double band_width = frequency_range / kBandCount / 10;         // audio_io.py:257
    double sums[kBandCount];
#pragma unroll
    for (int b = 0; b < kBandCount; b++) sums[b] = 0.0;
    for (int i = threadIdx.x; i < half; i += kThreads) {
      const double f = F[i], a = A[i];
#pragma unroll
      for (int b = 0; b < kBandCount; b++) {
        const double band_start = b * band_width, band_end = band_start + band_width;   // audio_io.py:262-263
        if (f >= band_start && f < band_end) sums[b] += a;
      }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int b = 0; b < kBandCount; b++) {
      double v = sums[b];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
      if (lane == 0) warp_sums[warp][b] = v;
    }
    __syncthreads();
    if (threadIdx.x < kBandCount) {
      double v = 0.0;
      for (int w = 0; w < kThreads / 32; w++) v += warp_sums[w][threadIdx.x];
      const double band_start = threadIdx.x * band_width, band_end = band_start + band_width;
      p.out_a[(long long)wdw * kBandCount + threadIdx.x] = v;
      p.out_f[(long long)wdw * kBandCount + threadIdx.x] = (band_start + band_end) / 2;   // audio_io.py:273
    }
    __syncthreads();
  }
}

static int check_arrays(const char* what, const double* a, const double* f, int n_windows, int n, long long stride) {
  AIKO_CHECK_ARG(n_windows >= 0 && n >= 0, "%s: negative size", what);
  AIKO_CHECK_ARG(n_windows == 0 || (a && f), "%s: null array", what);
  AIKO_CHECK_ARG(stride == 0 || stride >= n, "%s: frequency stride %lld smaller than a window", what, stride);
  return AIKO_B200_OK;
}

}  // namespace aiko

using namespace aiko;

extern "C" int aiko_b200_audio_filter_f64(const double* amplitudes, const double* frequencies, int n_windows, int n,
                                          int64_t frequency_stride, double* top_frequencies, double* top_amplitudes,
                                          int* top_count, void* stream) {
  const int rc = check_arrays("audio_filter", amplitudes, frequencies, n_windows, n, frequency_stride);
  if (rc != AIKO_B200_OK) return rc;
  if (n_windows == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(top_frequencies && top_amplitudes && top_count, "audio_filter: null output");
  ArrayParams p{amplitudes, frequencies, frequency_stride, n_windows, n, top_amplitudes, top_frequencies, top_count};
  long long grid = (long long)sm_count() * 4;
  if (grid > n_windows) grid = n_windows;
  audio_filter_f64_kernel<<<(unsigned)grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return cuda_status(cudaGetLastError(), "audio_filter launch");
}

extern "C" int aiko_b200_audio_bands_f64(const double* amplitudes, const double* frequencies, int n_windows, int n,
                                         int64_t frequency_stride, double* band_frequencies, double* band_amplitudes,
                                         void* stream) {
  const int rc = check_arrays("audio_bands", amplitudes, frequencies, n_windows, n, frequency_stride);
  if (rc != AIKO_B200_OK) return rc;
  if (n_windows == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(n >= 2, "audio_bands: window of %d values has no first half", n);
  AIKO_CHECK_ARG(band_frequencies && band_amplitudes, "audio_bands: null output");
  ArrayParams p{amplitudes, frequencies, frequency_stride, n_windows, n, band_amplitudes, band_frequencies, nullptr};
  long long grid = (long long)sm_count() * 4;
  if (grid > n_windows) grid = n_windows;
  audio_bands_f64_kernel<<<(unsigned)grid, kThreads, 0, static_cast<cudaStream_t>(stream)>>>(p);
  return cuda_status(cudaGetLastError(), "audio_bands launch");
}
