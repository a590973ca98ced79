// The reference's audio chain, batched over windows: PE_FFT -> PE_AudioResampler (8 band sums)
// -> PE_AudioFilter (spectral gate + 100 loudest bins), R/elements/media/audio_io.py:202-320
// (dead code in the reference: it sits inside a string literal; oracle = oracle/audio.py).
//
// One CTA per window.  A real window of N samples is packed into N/2 complex points
// (z[n] = x[2n] + i x[2n+1]), transformed in shared memory by a mixed-radix (4/5/3/2) Stockham
// autosort FFT that ping-pongs between two buffers, and unpacked to the N-point spectrum;
// magnitudes / 1e6 go to HBM once, and the band sums and the gate + top-100 selection run on
// the copy that is still in shared memory.  Twiddles come from a table generated in double
// precision (aiko_b200_audio_prepare); all arithmetic is fp32 (tolerance 1e-4).
// 16000-sample windows (8000 = 20^3 points) take three radix-20 passes whose butterflies stay in
// registers, and the next window's samples are fetched while the current one is unpacked and gated.
// Per window HBM sees 2N bytes in (int16) and 4N bytes out: the path is HBM-bound.
#include <stdlib.h>

#include "common.cuh"

namespace aiko {

constexpr int kAudioThreads = 1024;
constexpr int kMaxStages = 16;
constexpr int kMaxHalf = 8192;       // N/2 complex points must fit twice in shared memory
constexpr int kTopK = 100;           // AF_SAMPLES_MAXIMUM, audio_io.py:206
constexpr int kBands = 8;            // AR_BAND_COUNT, audio_io.py:237
constexpr int kHistBins = 4096;
constexpr int kDirectSort = 1024;    // up to this many gated bins are sorted directly
constexpr int kRankSort = 512;
constexpr int kDirectList = kHistBins * 4 / 8;   // u64 entries that fit the histogram's storage (2048)       // ... and up to this many by counting ranks instead of exchanging

struct AudioParams {
  const void* audio;
  int dtype;  // 0 = int16, 1 = float32
  int n_windows, n, half;
  int n_stages, radix[kMaxStages];
  const float2* tw_half;  // exp(-2 pi i j / half), j < half
  const float2* tw_full;  // exp(-2 pi i k / n),    k <= half
  float* amplitudes;      // [windows, n]
  float* band_sums;       // [windows, 8] or null
  int band_lo[kBands], band_hi[kBands];
  float* top_freq;        // [windows, 100] or null
  float* top_amp;
  int* top_count;
  int gate_lo, gate_hi;   // bins k (folded: min(i, n - i)) passing the frequency gate
  float amp_min, amp_max, hz_per_bin;
  int tw_in_smem;
};

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(__fmaf_rn(a.x, b.x, -a.y * b.y), __fmaf_rn(a.x, b.y, a.y * b.x));
}
__device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)
__device__ __forceinline__ float2 cscale(float2 a, float s) { return make_float2(a.x * s, a.y * s); }

template <int R>
__device__ __forceinline__ void dft_small(float2* a) {
  if (R == 2) {
    const float2 t = a[1];
    a[1] = csub(a[0], t);
    a[0] = cadd(a[0], t);
  } else if (R == 3) {
    const float2 t1 = cadd(a[1], a[2]);
    const float2 t2 = make_float2(a[0].x - 0.5f * t1.x, a[0].y - 0.5f * t1.y);
    const float2 t3 = mul_mi(cscale(csub(a[1], a[2]), 0.86602540378443864676f));
    a[0] = cadd(a[0], t1);
    a[1] = cadd(t2, t3);
    a[2] = csub(t2, t3);
  } else if (R == 4) {
    const float2 t0 = cadd(a[0], a[2]), t1 = csub(a[0], a[2]);
    const float2 t2 = cadd(a[1], a[3]), t3 = mul_mi(csub(a[1], a[3]));
    a[0] = cadd(t0, t2);
    a[1] = cadd(t1, t3);
    a[2] = csub(t0, t2);
    a[3] = csub(t1, t3);
  } else {  // 5
    const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
    const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
    const float2 t1 = cadd(a[1], a[4]), t2 = cadd(a[2], a[3]);
    const float2 t3 = csub(a[1], a[4]), t4 = csub(a[2], a[3]);
    const float2 m1 = make_float2(a[0].x + c1 * t1.x + c2 * t2.x, a[0].y + c1 * t1.y + c2 * t2.y);
    const float2 m2 = make_float2(a[0].x + c2 * t1.x + c1 * t2.x, a[0].y + c2 * t1.y + c1 * t2.y);
    const float2 n1 = mul_mi(make_float2(s1 * t3.x + s2 * t4.x, s1 * t3.y + s2 * t4.y));
    const float2 n2 = mul_mi(make_float2(s2 * t3.x - s1 * t4.x, s2 * t3.y - s1 * t4.y));
    a[0] = cadd(a[0], cadd(t1, t2));
    a[1] = cadd(m1, n1);
    a[4] = csub(m1, n1);
    a[2] = cadd(m2, n2);
    a[3] = csub(m2, n2);
  }
}

// One Stockham pass: length-n sub-transforms interleaved with stride s; reads x, writes y.
// b = p * s + q enumerates the n / R * s butterflies; p = floor(b / s) comes from a float divide
// that is exact here ((b + 0.5) / s is never within rounding of an integer for b, s < 2^20).
// Each thread runs two butterflies at a time so their shared-memory latencies overlap.
// Complex points live in shared memory with one pad element after every 16, so the strided
// writes of the early Stockham passes (stride 4 and 16 elements) spread over all banks.
__device__ __forceinline__ int pidx(int i) { return i + (i >> 4); }
__host__ __device__ inline size_t padded_points(size_t n) { return n + (n >> 4) + 1; }

template <int R>
__device__ __forceinline__ void butterfly_load(const float2* __restrict__ x, int b, int m, int s, float inv_s,
                                               float2* a, int* p_out, int* q_out) {
  const int p = __float2int_rd(((float)b + 0.5f) * inv_s), q = b - p * s;
#pragma unroll
  for (int k = 0; k < R; k++) a[k] = x[pidx(q + s * (p + m * k))];
  *p_out = p;
  *q_out = q;
}
template <int R>
__device__ __forceinline__ void butterfly_store(float2* __restrict__ y, float2* a, int p, int q, int s,
                                                const float2* __restrict__ tw, int tw_step) {
  dft_small<R>(a);
  y[pidx(q + s * (R * p))] = a[0];
#pragma unroll
  for (int j = 1; j < R; j++) y[pidx(q + s * (R * p + j))] = cmul(a[j], tw[p * j * tw_step]);
}
template <int R, int NT>
__device__ __forceinline__ void stockham_pass(const float2* __restrict__ x, float2* __restrict__ y, int n, int s,
                                              const float2* __restrict__ tw, int half) {
  const int m = n / R, total = m * s, tw_step = half / n;
  const float inv_s = 1.0f / (float)s;
  int b = threadIdx.x;
  for (; b + NT < total; b += 2 * NT) {
    float2 a0[R], a1[R];
    int p0, q0, p1, q1;
    butterfly_load<R>(x, b, m, s, inv_s, a0, &p0, &q0);
    butterfly_load<R>(x, b + NT, m, s, inv_s, a1, &p1, &q1);
    butterfly_store<R>(y, a0, p0, q0, s, tw, tw_step);
    butterfly_store<R>(y, a1, p1, q1, s, tw, tw_step);
  }
  if (b < total) {
    float2 a0[R];
    int p0, q0;
    butterfly_load<R>(x, b, m, s, inv_s, a0, &p0, &q0);
    butterfly_store<R>(y, a0, p0, q0, s, tw, tw_step);
  }
}

// ---- half = 8000 = 20^3 (the reference's 16 kHz x 1 s window): three radix-20 passes ----------
// A radix-20 butterfly runs in registers as 5 DFTs of 4 points, 12 twiddles, 4 DFTs of 5 points
// (n = 5 n1 + n2, k = k1 + 4 k2), so the window crosses shared memory three times instead of six.
// Points are padded by one element after every 20: the stride-20 stores of the first pass then
// step through shared memory with an odd stride and are conflict-free.
__device__ __forceinline__ int pad20(int i) { return i + i / 20; }

__device__ __forceinline__ void dft20(float2* a) {   // in: a[n]; out: X[k1 + 4 k2] in a[5 k1 + k2]
  // exp(-2 pi i e / 20), e = 1 .. 9, 12
  const float2 w1 = make_float2(0.95105651629515357f, -0.30901699437494742f);
  const float2 w2 = make_float2(0.80901699437494742f, -0.58778525229247313f);
  const float2 w3 = make_float2(0.58778525229247313f, -0.80901699437494742f);
  const float2 w4 = make_float2(0.30901699437494742f, -0.95105651629515357f);
  const float2 w6 = make_float2(-0.30901699437494742f, -0.95105651629515357f);
  const float2 w8 = make_float2(-0.80901699437494742f, -0.58778525229247313f);
  const float2 w9 = make_float2(-0.95105651629515357f, -0.30901699437494742f);
  const float2 w12 = make_float2(-0.80901699437494742f, 0.58778525229247313f);
  float2 y[4][5];                       // y[k1][n2]
#pragma unroll
  for (int n2 = 0; n2 < 5; n2++) {
    float2 t[4] = {a[n2], a[5 + n2], a[10 + n2], a[15 + n2]};
    dft_small<4>(t);
#pragma unroll
    for (int k1 = 0; k1 < 4; k1++) y[k1][n2] = t[k1];
  }
  y[1][1] = cmul(y[1][1], w1); y[1][2] = cmul(y[1][2], w2); y[1][3] = cmul(y[1][3], w3); y[1][4] = cmul(y[1][4], w4);
  y[2][1] = cmul(y[2][1], w2); y[2][2] = cmul(y[2][2], w4); y[2][3] = cmul(y[2][3], w6); y[2][4] = cmul(y[2][4], w8);
  y[3][1] = cmul(y[3][1], w3); y[3][2] = cmul(y[3][2], w6); y[3][3] = cmul(y[3][3], w9); y[3][4] = cmul(y[3][4], w12);
#pragma unroll
  for (int k1 = 0; k1 < 4; k1++) {
    dft_small<5>(y[k1]);
#pragma unroll
    for (int k2 = 0; k2 < 5; k2++) a[5 * k1 + k2] = y[k1][k2];
  }
}

// One radix-20 Stockham pass over 8000 points by threads 0 .. 399: sub-transform stride S
// (1, 20, 400), reads x, writes y; twiddles exp(-2 pi i p j S / 8000) from the table.
template <int S>
__device__ __forceinline__ void radix20_pass(const float2* __restrict__ x, float2* __restrict__ y,
                                             const float2* __restrict__ tw) {
  constexpr int M = 400 / S;            // butterflies per sub-transform = stride between the 20 inputs / S
  const int b = threadIdx.x;
  if (b >= 400) return;
  const int p = b / S, q = b - p * S;
  float2 a[20];
#pragma unroll
  for (int k = 0; k < 20; k++) a[k] = x[pad20(q + S * (p + M * k))];
  dft20(a);
#pragma unroll
  for (int j = 0; j < 20; j++) {
    float2 v = a[5 * (j & 3) + (j >> 2)];
    if (S < 400 && j > 0) v = cmul(v, tw[p * j * S]);
    y[pad20(q + S * (20 * p + j))] = v;
  }
}

template <bool FAST>
__device__ __forceinline__ int point_index(int i) { return FAST ? pad20(i) : pidx(i); }

template <int NT, bool FAST>
__global__ void __launch_bounds__(NT, 1) audio_chain_kernel(const AudioParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  float2* buf0 = reinterpret_cast<float2*>(smem);
  float2* buf1 = buf0 + padded_points(p.half);
  const float2* tw = p.tw_half;
  if (p.tw_in_smem) {           // the pass twiddles next to the data: one copy per CTA, reused for every window
    float2* tws = buf1 + padded_points(p.half);
    for (int i = threadIdx.x; i < p.half; i += NT) tws[i] = __ldg(p.tw_half + i);
    tw = tws;
    __syncthreads();
  }
  __shared__ int s_count, s_threshold_bin, s_kept;
  __shared__ float s_band[kBands][4];
  __shared__ unsigned long long s_top[kTopK];
  __shared__ int hist[kHistBins];
  const int half = p.half, n = p.n;

  // int16 windows are fetched one window ahead into registers (while the previous window is
  // unpacked and gated), so the FFT never waits for HBM
  constexpr int PRE = kMaxHalf / NT;
  uint32_t pre[PRE];
  auto prefetch = [&](int w) {
    const uint32_t* in = reinterpret_cast<const uint32_t*>(static_cast<const int16_t*>(p.audio) + (size_t)w * n);
#pragma unroll
    for (int j = 0; j < PRE; j++) {
      const int i = threadIdx.x + j * NT;
      pre[j] = i < half ? __ldg(in + i) : 0u;
    }
  };
  if (p.dtype == 0 && (int)blockIdx.x < p.n_windows) prefetch(blockIdx.x);
  const float2 w_odd = __ldg(p.tw_full + 1);          // exp(-2 pi i / n): tw_full[k] = tw_half[k / 2] * w_odd^(k & 1)

  for (int w = blockIdx.x; w < p.n_windows; w += gridDim.x) {
    if (threadIdx.x == 0) s_count = 0;   // gated bins of this window (ordered before their appends by the barriers below)
    // ---- pack the real window into half complex points
    if (p.dtype == 0) {
#pragma unroll
      for (int j = 0; j < PRE; j++) {
        const int i = threadIdx.x + j * NT;
        if (i < half)
          buf0[point_index<FAST>(i)] = make_float2((float)(short)(pre[j] & 0xffffu), (float)(short)(pre[j] >> 16));
      }
    } else {
      const float2* in = reinterpret_cast<const float2*>(static_cast<const float*>(p.audio) + (size_t)w * n);
      for (int i = threadIdx.x; i < half; i += NT) buf0[point_index<FAST>(i)] = __ldg(in + i);
    }
    __syncthreads();
    // ---- half-length complex FFT
    float2 *x = buf0, *y = buf1;
    if (FAST) {
      radix20_pass<1>(buf0, buf1, tw);
      __syncthreads();
      radix20_pass<20>(buf1, buf0, tw);
      __syncthreads();
      radix20_pass<400>(buf0, buf1, tw);
      __syncthreads();
      x = buf1; y = buf0;               // spectrum in buf1, buf0 is idle
    } else {
      int len = half, stride = 1;
      for (int st = 0; st < p.n_stages; st++) {
        const int r = p.radix[st];
        if (r == 4) stockham_pass<4, NT>(x, y, len, stride, tw, half);
        else if (r == 5) stockham_pass<5, NT>(x, y, len, stride, tw, half);
        else if (r == 3) stockham_pass<3, NT>(x, y, len, stride, tw, half);
        else stockham_pass<2, NT>(x, y, len, stride, tw, half);
        len /= r;
        stride *= r;
        float2* t = x; x = y; y = t;
        __syncthreads();
      }
    }
    // ---- unpack to the spectrum of the real window, magnitudes / 1e6 (audio_io.py:299,309)
    float* amp = reinterpret_cast<float*>(y);          // half + 1 magnitudes, in the idle buffer
    float* out = p.amplitudes + (size_t)w * n;
    // PE_AudioFilter's gates are applied right here, while the magnitudes are in registers: the few
    // bins that pass (a tone and its neighbours) go to a small list in shared memory (it shares its
    // storage with the histogram of the rare many-candidates path, which re-collects everything).
    const bool gate = p.top_amp != nullptr;
    const int gate_klo = max(p.gate_lo, 0), gate_khi = min(p.gate_hi, half);
    unsigned long long* direct = reinterpret_cast<unsigned long long*>(hist);
    auto gate_bin = [&](int k, float a) {
      if (k < gate_klo || k > gate_khi || !(a >= p.amp_min && a <= p.amp_max)) return;
      const int copies = (k != 0 && k != half) ? 2 : 1;
      const int start = atomicAdd(&s_count, copies);
      const unsigned long long hi = (unsigned long long)__float_as_uint(a) << 32;
      if (start < kDirectList) direct[start] = hi | (0xffffffffu - (unsigned)k);
      if (copies == 2 && start + 1 < kDirectList) direct[start + 1] = hi | (0xffffffffu - (unsigned)(n - k));
    };
    if (p.dtype == 0 && w + (int)gridDim.x < p.n_windows) prefetch(w + gridDim.x);
    // Bins k and half - k come from the same pair (z[k], z[half - k]): X[k] = E + W^k O and
    // X[half - k] = conj(E - W^k O), so one thread makes both magnitudes from one set of loads.
    for (int k = threadIdx.x; k <= half / 2; k += NT) {
      const int kr = half - k;                            // (k == 0 pairs bin 0 with bin half)
      const float2 zk = x[point_index<FAST>(k)];
      const float2 zr = x[point_index<FAST>(k == 0 ? 0 : kr)];
      const float2 zc = make_float2(zr.x, -zr.y);
      const float2 e = cscale(cadd(zk, zc), 0.5f);
      const float2 o = cscale(mul_mi(csub(zk, zc)), 0.5f);
      float2 tk = tw[k >> 1];
      if (k & 1) tk = cmul(tk, w_odd);
      const float2 t = cmul(tk, o);
      const float2 xa = cadd(e, t), xb = csub(e, t);
      float a, b;                                         // |X| / 1e6; the approximate root is within 1e-6 relative
      asm("sqrt.approx.f32 %0, %1;" : "=f"(a) : "f"(__fmaf_rn(xa.x, xa.x, xa.y * xa.y)));
      asm("sqrt.approx.f32 %0, %1;" : "=f"(b) : "f"(__fmaf_rn(xb.x, xb.x, xb.y * xb.y)));
      a *= 1e-6f;
      b *= 1e-6f;
      amp[k] = a;
      out[k] = a;
      if (k != 0) out[n - k] = a;                         // |X[n-k]| = |X[k]| for real input
      if (gate) gate_bin(k, a);
      if (kr != k) {
        amp[kr] = b;
        out[kr] = b;
        if (k != 0) out[n - kr] = b;                      // (bin half has no mirror)
        if (gate) gate_bin(kr, b);
      }
    }
    __syncthreads();
    // ---- PE_AudioResampler: 8 band sums over the first half of the spectrum (audio_io.py:250-277)
    if (p.band_sums != nullptr) {
      // every warp sums one slice of one band; the slices of a band are then added in a fixed order
      constexpr int kParts = NT / 32 / kBands;
      const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
      const int band = warp % kBands, part = warp / kBands;
      const int lo = p.band_lo[band], len = p.band_hi[band] - lo;
      const int k0 = lo + (int)((long long)len * part / kParts), k1 = lo + (int)((long long)len * (part + 1) / kParts);
      float s0 = 0.f, s1 = 0.f;
      int k = k0 + lane;
      for (; k + 32 < k1; k += 64) { s0 += amp[k]; s1 += amp[k + 32]; }
      if (k < k1) s0 += amp[k];
      float sum = s0 + s1;
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
      if (lane == 0) s_band[band][part] = sum;
      __syncthreads();
      if (threadIdx.x < kBands) {
        float total = 0.f;
#pragma unroll
        for (int q = 0; q < kParts; q++) total += s_band[threadIdx.x][q];
        p.band_sums[(size_t)w * kBands + threadIdx.x] = total;
      }
    }
    // ---- PE_AudioFilter: gate on amplitude and |frequency|, then the 100 largest by amplitude,
    //      ties in original index order (a stable descending sort, audio_io.py:213-230)
    if (p.top_amp != nullptr) {
      unsigned long long* list = reinterpret_cast<unsigned long long*>(x);     // the spectrum is no longer needed
      const int list_cap = half;                                               // half * 8 bytes = one buffer
      const int klo = max(p.gate_lo, 0), khi = min(p.gate_hi, half);
      const int lane = threadIdx.x & 31;
      // Appends every bin that passes the gates and whose key bin is >= min_bin (each bin k stands
      // for spectrum indices k and n - k).  One atomic per warp; order in the list is irrelevant.
      auto collect = [&](unsigned min_bin) {
        for (int base = klo; base <= khi; base += NT) {
          const int k = base + (int)threadIdx.x;
          float a = 0.f;
          int copies = 0;
          if (k <= khi) {
            a = amp[k];
            if (a >= p.amp_min && a <= p.amp_max && (__float_as_uint(a) >> 20) >= min_bin)
              copies = (k != 0 && k != half) ? 2 : 1;
          }
          if (__ballot_sync(0xffffffffu, copies != 0) == 0) continue;
          // exclusive prefix of `copies` (0, 1 or 2) over the warp from two ballots
          const unsigned m1 = __ballot_sync(0xffffffffu, copies >= 1), m2 = __ballot_sync(0xffffffffu, copies == 2);
          const unsigned below = (1u << lane) - 1u;
          int start = 0;
          if (lane == 0) start = atomicAdd(&s_count, __popc(m1) + __popc(m2));
          start = __shfl_sync(0xffffffffu, start, 0) + __popc(m1 & below) + __popc(m2 & below);
          if (copies) {
            const unsigned long long hi = (unsigned long long)__float_as_uint(a) << 32;
            if (start < list_cap) list[start] = hi | (0xffffffffu - (unsigned)k);
            if (copies == 2 && start + 1 < list_cap) list[start + 1] = hi | (0xffffffffu - (unsigned)(n - k));
          }
        }
      };
      const bool direct_ok = s_count <= kDirectList;      // every gated bin is in the small list already
      if (!direct_ok) {
        __syncthreads();
        if (threadIdx.x == 0) s_count = 0;
        __syncthreads();
        collect(0u);
        __syncthreads();
      }
      if (!direct_ok && s_count > kDirectSort) {
        // Many bins pass the gates (noise-like input): keep only key bins that can reach the top 100.
        __syncthreads();
        for (int i = threadIdx.x; i < kHistBins; i += NT) hist[i] = 0;
        if (threadIdx.x == 0) { s_count = 0; s_threshold_bin = 0; }
        __syncthreads();
        for (int k = klo + threadIdx.x; k <= khi; k += NT) {
          const float a = amp[k];
          if (a >= p.amp_min && a <= p.amp_max)
            atomicAdd(&hist[__float_as_uint(a) >> 20], (k != 0 && k != half) ? 2 : 1);
        }
        __syncthreads();
        if (threadIdx.x < 32) {   // highest bin T with count(bins >= T) >= 100
          int running = 0, found = -1;
          for (int base = kHistBins - 32; base >= 0 && found < 0; base -= 32) {
            const int c = hist[base + 31 - threadIdx.x];            // lane 0 = highest bin of the chunk
            int incl = c;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
              const int t = __shfl_up_sync(0xffffffffu, incl, d);
              if ((int)threadIdx.x >= d) incl += t;
            }
            const unsigned hit = __ballot_sync(0xffffffffu, running + incl >= kTopK);
            if (hit) found = base + 31 - (__ffs(hit) - 1);
            running += __shfl_sync(0xffffffffu, incl, 31);
          }
          if (threadIdx.x == 0) s_threshold_bin = max(found, 0);
        }
        __syncthreads();
        collect((unsigned)s_threshold_bin);
        __syncthreads();
      }
      const int count = min(s_count, list_cap);
      if (direct_ok) list = direct;
      const unsigned long long* sorted = list;
      if (count <= kRankSort) {
        // Few candidates: every thread finds the rank of its own key by counting the larger ones
        // (keys are distinct: the bin index is part of them).  No exchanges.  With more than 100
        // candidates (a loud tone leaks into ~300 bins) only those that can reach the top 100 are
        // ranked: a 512-bin histogram over the amplitude bits (the gate spans 442 such bins) gives the
        // bin holding the 100th value, and the candidates at or above it are compacted first --
        // the quadratic part shrinks from count^2 to little more than 100^2 comparisons.
        const unsigned long long* cand = list;
        int ncand = count;
        if (count > kTopK && direct_ok) {
          int* fine = hist + 2048;                                      // (the list itself sits in hist[0 .. 1024))
          unsigned long long* keep = reinterpret_cast<unsigned long long*>(hist + 2560);
          const unsigned base = __float_as_uint(p.amp_min) >> 17;
          auto fine_bin = [&](unsigned long long key) { return min((int)((unsigned)(key >> 49) - base), 511); };
          if (threadIdx.x < 512) fine[threadIdx.x] = 0;
          if (threadIdx.x == 0) { s_threshold_bin = 0; s_kept = 0; }
          __syncthreads();
          for (int i = threadIdx.x; i < count; i += NT) atomicAdd(&fine[fine_bin(list[i])], 1);
          __syncthreads();
          if (threadIdx.x < 32) {                                       // highest bin T with count(bins >= T) >= 100
            int running = 0, found = -1;
            for (int top = 511; top >= 0 && found < 0; top -= 32) {
              const int c = fine[top - (int)threadIdx.x];               // lane 0 = highest bin of the chunk
              int incl = c;
#pragma unroll
              for (int d = 1; d < 32; d <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, d);
                if ((int)threadIdx.x >= d) incl += t;
              }
              const unsigned hit = __ballot_sync(0xffffffffu, running + incl >= kTopK);
              if (hit) found = top - (__ffs(hit) - 1);
              running += __shfl_sync(0xffffffffu, incl, 31);
            }
            if (threadIdx.x == 0) s_threshold_bin = max(found, 0);
          }
          __syncthreads();
          for (int i = threadIdx.x; i < count; i += NT) {
            const unsigned long long key = list[i];
            if (fine_bin(key) >= s_threshold_bin) keep[atomicAdd(&s_kept, 1)] = key;
          }
          __syncthreads();
          cand = keep;
          ncand = s_kept;
        }
        for (int i = threadIdx.x; i < ncand; i += NT) {
          const unsigned long long key = cand[i];
          int rank = 0;
#pragma unroll 4
          for (int j = 0; j < ncand; j++) rank += (cand[j] > key) ? 1 : 0;
          if (rank < kTopK) s_top[rank] = key;
        }
        sorted = s_top;
        __syncthreads();
      } else {
      int padded = 1;
      while (padded < count) padded <<= 1;
      for (int i = count + threadIdx.x; i < padded; i += NT) list[i] = 0ull;   // sorts last
      __syncthreads();
      // Bitonic sort, descending.  Only the warps that hold elements take part; exchanges with
      // a stride below 32 stay inside a warp (__syncwarp), wider ones use a named barrier over
      // the participating threads.
      const int sorters = min(max(padded, 32), NT);
      if ((int)threadIdx.x < sorters) {
        for (int size = 2; size <= padded; size <<= 1) {
          for (int strd = size >> 1; strd > 0; strd >>= 1) {
            for (int i = threadIdx.x; i < padded; i += sorters) {
              const int j = i ^ strd;
              if (j > i) {
                const bool desc = (i & size) == 0;
                const unsigned long long a = list[i], b = list[j];
                if ((a < b) == desc) { list[i] = b; list[j] = a; }
              }
            }
            // the next step crosses warps if its stride is >= 32
            const int next = (strd > 1) ? (strd >> 1) : size;
            const bool more = (strd > 1) || (size < padded);
            if (more && (strd >= 32 || next >= 32)) asm volatile("bar.sync 1, %0;" ::"r"(sorters) : "memory");
            else __syncwarp();
          }
        }
      }
      __syncthreads();
      }
      const int take = min(count, kTopK);
      for (int i = threadIdx.x; i < kTopK; i += NT) {
        float a = 0.f, f = 0.f;
        if (i < take) {
          const unsigned long long e = sorted[i];
          const unsigned idx = 0xffffffffu - (unsigned)(e & 0xffffffffu);
          a = __uint_as_float((unsigned)(e >> 32));
          f = (float)min(idx, (unsigned)n - idx) * p.hz_per_bin;
        }
        p.top_amp[(size_t)w * kTopK + i] = a;
        p.top_freq[(size_t)w * kTopK + i] = f;
      }
      if (threadIdx.x == 0) p.top_count[w] = take;
    }
    __syncthreads();   // shared buffers are reused by the next window
  }
}

__global__ void audio_twiddles_kernel(float2* tw_half, float2* tw_full, int half) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double s, c;
  if (i < half) {
    sincospi(-2.0 * (double)i / (double)half, &s, &c);
    tw_half[i] = make_float2((float)c, (float)s);
  }
  if (i <= half) {
    sincospi(-(double)i / (double)half, &s, &c);   // exp(-2 pi i k / n), n = 2 * half
    tw_full[i] = make_float2((float)c, (float)s);
  }
}

static int factorize(int m, int* radix) {
  int count = 0;
  const int order[4] = {4, 5, 3, 2};
  for (int r : order)
    while (m % r == 0 && count < kMaxStages) { radix[count++] = r; m /= r; }
  return m == 1 ? count : -1;
}

}  // namespace aiko

using namespace aiko;

extern "C" size_t aiko_b200_audio_workspace_bytes(int n_samples) {
  if (n_samples < 4 || (n_samples & 1)) return 0;
  const size_t half = (size_t)n_samples / 2;
  return (half + (half + 1) + 1) * sizeof(float2);
}

extern "C" int aiko_b200_audio_prepare(void* workspace, size_t workspace_bytes, int n_samples, void* stream) {
  const size_t need = aiko_b200_audio_workspace_bytes(n_samples);
  AIKO_CHECK_ARG(need != 0, "audio_prepare: window length %d must be even and >= 4", n_samples);
  if (!workspace || workspace_bytes < need) {
    set_last_error("audio_prepare: workspace of %zu bytes, need %zu", workspace_bytes, need);
    return AIKO_B200_ERR_WORKSPACE;
  }
  const int half = n_samples / 2;
  float2* tw_half = static_cast<float2*>(workspace);
  audio_twiddles_kernel<<<(half + 1 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(tw_half, tw_half + half,
                                                                                               half);
  return cuda_status(cudaGetLastError(), "audio_twiddles launch");
}

extern "C" int aiko_b200_audio_chain(const void* audio, int dtype, int n_windows, int n_samples, double sample_rate,
                                     float* amplitudes, float* band_sums, const int* band_ranges_host,
                                     float* top_freq, float* top_amp, int* top_count, int gate_lo, int gate_hi,
                                     const void* workspace, size_t workspace_bytes, void* stream) {
  AIKO_CHECK_ARG(n_windows >= 0, "audio_chain: n_windows < 0");
  if (n_windows == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(audio && amplitudes, "audio_chain: null pointer");
  AIKO_CHECK_ARG(dtype == 0 || dtype == 1, "audio_chain: dtype %d (0 = int16, 1 = float32)", dtype);
  const size_t need = aiko_b200_audio_workspace_bytes(n_samples);
  AIKO_CHECK_ARG(need != 0, "audio_chain: window length %d must be even and >= 4", n_samples);
  if (!workspace || workspace_bytes < need) {
    set_last_error("audio_chain: workspace of %zu bytes, need %zu (and it must have been prepared)", workspace_bytes,
                   need);
    return AIKO_B200_ERR_WORKSPACE;
  }
  AIKO_CHECK_ARG(!band_sums || band_ranges_host, "audio_chain: band_sums without band ranges");
  AIKO_CHECK_ARG((top_freq != nullptr) == (top_amp != nullptr) && (top_amp != nullptr) == (top_count != nullptr),
                 "audio_chain: top_freq, top_amp and top_count go together");
  AudioParams p{};
  p.audio = audio; p.dtype = dtype; p.n_windows = n_windows; p.n = n_samples; p.half = n_samples / 2;
  if (p.half > kMaxHalf) {
    set_last_error("audio_chain: window of %d samples exceeds the %d this kernel keeps in shared memory", n_samples,
                   2 * kMaxHalf);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  p.n_stages = factorize(p.half, p.radix);
  if (p.n_stages < 0) {
    set_last_error("audio_chain: %d / 2 has a prime factor other than 2, 3, 5", n_samples);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  p.tw_half = static_cast<const float2*>(workspace);
  p.tw_full = p.tw_half + p.half;
  p.amplitudes = amplitudes;
  p.band_sums = band_sums;
  for (int b = 0; b < kBands; b++) {
    p.band_lo[b] = band_sums ? band_ranges_host[2 * b] : 0;
    p.band_hi[b] = band_sums ? band_ranges_host[2 * b + 1] : 0;
    AIKO_CHECK_ARG(p.band_lo[b] >= 0 && p.band_hi[b] <= p.half && p.band_lo[b] <= p.band_hi[b],
                   "audio_chain: band %d range outside the first half of the spectrum", b);
  }
  p.top_freq = top_freq; p.top_amp = top_amp; p.top_count = top_count;
  p.gate_lo = gate_lo; p.gate_hi = gate_hi;
  p.amp_min = 0.1f;   // AF_AMPLITUDE_MINIMUM / MAXIMUM, audio_io.py:202-203
  p.amp_max = 12.0f;
  p.hz_per_bin = (float)(sample_rate / (double)n_samples);
  size_t smem = (size_t)2 * padded_points(p.half) * sizeof(float2);
  // static shared memory: 16 KB histogram + ~1 KB of small tables; keep the twiddles on chip when there is room
  p.tw_in_smem = (smem + (size_t)p.half * sizeof(float2) + 18 * 1024 <= (size_t)device_max_smem_optin()) ? 1 : 0;
  if (p.tw_in_smem) smem += (size_t)p.half * sizeof(float2);
  // 16000-sample windows (half = 20^3) take the radix-20 kernel: 512 threads, 128 registers each
  const bool fast = p.half == 8000 && !getenv("AIKO_B200_AUDIO_GENERIC");
  auto kernel = fast ? audio_chain_kernel<512, true> : audio_chain_kernel<kAudioThreads, false>;
  const int threads = fast ? 512 : kAudioThreads;
  AIKO_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  AIKO_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
  if (per_sm < 1) per_sm = 1;
  long long grid = (long long)sm_count() * per_sm;
  if (grid > n_windows) grid = n_windows;
  kernel<<<(unsigned)grid, threads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return cuda_status(cudaGetLastError(), "audio_chain launch");
}
