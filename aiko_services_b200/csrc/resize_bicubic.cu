// Batched antialiased bicubic resize, bit-exact with PIL.Image.resize((W, H)) for 8-bit
// images (default BICUBIC filter) as called by the reference's ImageResize on PIL inputs
// (R/elements/media/image_io.py:218).  Pillow's algorithm: per axis, a table of 22-bit
// fixed-point weights over a support of 2 * max(scale, 1) source pixels; a horizontal pass
// into a uint8 intermediate image, then a vertical pass.
//
//   bicubic_coeffs_kernel  fills the two weight tables in a caller-provided workspace using
//                          explicitly rounded IEEE double operations (same results as the C
//                          library on the host, no FMA contraction);
//   resize_bicubic_rows    one CTA per band of output rows of one frame: source rows stream
//                          through the horizontal pass into a shared-memory ring of uint8
//                          intermediate rows, and each output row is produced from that ring,
//                          so a source byte is read from HBM once per band it belongs to and the
//                          intermediate image never exists in global memory.
#include "bicubic.cuh"

namespace aiko {

__device__ __forceinline__ double bicubic_filter(double x) {
  // a = -0.5: ((a + 2) x - (a + 3)) x x + 1   |   (((x - 5) x + 8) x - 4) a
  if (x < 0.0) x = -x;
  if (x < 1.0) return __dadd_rn(__dmul_rn(__dmul_rn(__dadd_rn(__dmul_rn(1.5, x), -2.5), x), x), 1.0);
  if (x < 2.0)
    return __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(x, -5.0), x), 8.0), x), -4.0), -0.5);
  return 0.0;
}

// one thread per output index of one axis
__global__ void bicubic_coeffs_kernel(int* ws, int in_size, int out_size, int ksize, int off_bounds, int off_k) {
  const int xx = blockIdx.x * blockDim.x + threadIdx.x;
  if (xx >= out_size) return;
  int* bounds = ws + off_bounds;
  int* k = ws + off_k + (size_t)xx * ksize;
  if (in_size == out_size) {  // Pillow skips the pass: identity weights reproduce the copy exactly
    bounds[2 * xx] = xx;
    bounds[2 * xx + 1] = 1;
    k[0] = 1 << kPrecisionBits;
    for (int t = 1; t < ksize; t++) k[t] = 0;
    return;
  }
  const double scale = __ddiv_rn((double)in_size, (double)out_size);
  const double filterscale = scale < 1.0 ? 1.0 : scale;
  const double support = __dmul_rn(2.0, filterscale);
  const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), scale);
  const double ss = __ddiv_rn(1.0, filterscale);
  int xmin = __double2int_rz(__dadd_rn(__dsub_rn(center, support), 0.5));
  if (xmin < 0) xmin = 0;
  int xmax = __double2int_rz(__dadd_rn(__dadd_rn(center, support), 0.5));
  if (xmax > in_size) xmax = in_size;
  xmax -= xmin;
  double ww = 0.0;
  for (int x = 0; x < xmax; x++)
    ww = __dadd_rn(ww, bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss)));
  for (int x = 0; x < ksize; x++) {
    int v = 0;
    if (x < xmax) {
      double w = bicubic_filter(__dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss));
      if (ww != 0.0) w = __ddiv_rn(w, ww);
      const double f = __dmul_rn(w, (double)(1 << kPrecisionBits));
      v = f < 0 ? __double2int_rz(__dadd_rn(-0.5, f)) : __double2int_rz(__dadd_rn(0.5, f));
    }
    k[x] = v;
  }
  bounds[2 * xx] = xmin;
  bounds[2 * xx + 1] = xmax;
}

__device__ __forceinline__ uint8_t clip8_q22(int v) {
  v >>= kPrecisionBits;
  return (uint8_t)min(max(v, 0), 255);
}

struct BicubicParams {
  const uint8_t* src;
  uint8_t* dst;
  const int* ws;
  int n, sh, sw, dh, dw, ch;
  int ksx, ksy, off_bx, off_kx, off_by, off_ky;
  int band, bands_per_frame, ring_rows;
  int kx_in_smem;
  int chunks, chunk_w;            // output columns are split when a whole row does not fit in shared memory
  int src_stride, mid_stride;     // shared-memory pitch of a staged source row segment / an intermediate row
};

constexpr int kBicubicThreads = 512;
constexpr int kRowsPerRound = 4;   // source rows staged and filtered between two barriers

__global__ void __launch_bounds__(kBicubicThreads) resize_bicubic_rows(const BicubicParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int src_row_bytes = p.sw * p.ch, out_row_bytes = p.dw * p.ch;
  const int src_stride = p.src_stride, mid_stride = p.mid_stride;
  uint8_t* srow = smem;                               // kRowsPerRound staged source row segments
  uint8_t* ring = srow + (size_t)kRowsPerRound * src_stride;   // ring_rows intermediate rows
  int* bx_s = reinterpret_cast<int*>(ring + (size_t)p.ring_rows * mid_stride);   // xmin, count per column
  int* kx_s = bx_s + ((2 * p.chunk_w + 3) & ~3);      // weights, when they fit (16-byte aligned rows)

  const int* bx = p.ws + p.off_bx;
  const int* kx = p.ws + p.off_kx;
  const int* by = p.ws + p.off_by;
  const int* ky = p.ws + p.off_ky;
  // this CTA: (frame, band of output rows, chunk of output columns cx0 .. cx1)
  const int chunk = blockIdx.x % p.chunks, fb = blockIdx.x / p.chunks;
  const int cx0 = chunk * p.chunk_w, cx1 = min(p.dw, cx0 + p.chunk_w), cw = cx1 - cx0;
  for (int i = threadIdx.x; i < 2 * cw; i += kBicubicThreads) bx_s[i] = bx[2 * cx0 + i];
  if (p.kx_in_smem)
    for (int i = threadIdx.x; i < cw * p.ksx; i += kBicubicThreads) kx_s[i] = kx[(size_t)cx0 * p.ksx + i];
  const int* kxt = p.kx_in_smem ? kx_s : kx + (size_t)cx0 * p.ksx;
  // source bytes of a row that the chunk's columns read, from a 16-byte boundary
  const int s_lo = (bx[2 * cx0] * p.ch) & ~15;
  const int s_hi = min(src_row_bytes, (bx[2 * (cx1 - 1)] + bx[2 * (cx1 - 1) + 1]) * p.ch);
  const int seg_bytes = s_hi - s_lo, mid_row_bytes = cw * p.ch;
  __syncthreads();

  const int frame = fb / p.bands_per_frame;
  const int dy0 = (fb - frame * p.bands_per_frame) * p.band;
  const int dy1 = min(dy0 + p.band, p.dh);
  const uint8_t* fsrc = p.src + (size_t)frame * p.sh * src_row_bytes;
  uint8_t* fdst = p.dst + (size_t)frame * p.dh * out_row_bytes + (size_t)cx0 * p.ch;
  const int band_last = by[2 * (dy1 - 1)] + by[2 * (dy1 - 1) + 1] - 1;   // last source row this band needs

  int next_src = by[2 * dy0];
  for (int dy = dy0; dy < dy1; dy++) {
    const int ymin = by[2 * dy], ycnt = by[2 * dy + 1];
    // horizontal pass for the source rows this output row still needs (and up to
    // kRowsPerRound - 1 rows ahead, which the ring is sized for)
    while (next_src < ymin + ycnt) {
      const int nrows = min(kRowsPerRound, band_last - next_src + 1);
      const bool vec = ((reinterpret_cast<uintptr_t>(fsrc) | (uintptr_t)src_row_bytes) & 15) == 0;
      for (int r = 0; r < nrows; r++) {
        const uint8_t* g = fsrc + (size_t)(next_src + r) * src_row_bytes + s_lo;
        uint8_t* s = srow + (size_t)r * src_stride;
        if (vec) {                                  // (rows are a multiple of 16 bytes: rounding up stays inside the row)
          for (int i = threadIdx.x; i < ((seg_bytes + 15) >> 4); i += kBicubicThreads)
            reinterpret_cast<uint4*>(s)[i] = ldg_stream(reinterpret_cast<const uint4*>(g) + i);
        } else {
          for (int i = threadIdx.x; i < seg_bytes; i += kBicubicThreads) s[i] = g[i];
        }
      }
      __syncthreads();
      for (int r = 0; r < nrows; r++) {
        const uint8_t* srcrow = srow + (size_t)r * src_stride;
        uint8_t* mid = ring + (size_t)((next_src + r) % p.ring_rows) * mid_stride;
        for (int dx = threadIdx.x; dx < cw; dx += kBicubicThreads) {     // dx: column inside the chunk
          const int xmin = bx_s[2 * dx], cnt = bx_s[2 * dx + 1];
          const int4* k4 = reinterpret_cast<const int4*>(kxt + (size_t)dx * p.ksx);   // rows of 4-padded weights
          const uint8_t* s = srcrow + xmin * p.ch - s_lo;
          if (p.ch == 3) {
            int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
            for (int t = 0; t < cnt; t += 4) {       // padding weights are zero; staged rows have slack bytes
              const int4 w = k4[t >> 2];
              const uint8_t* q = s + 3 * t;
              a0 += q[0] * w.x + q[3] * w.y + q[6] * w.z + q[9] * w.w;
              a1 += q[1] * w.x + q[4] * w.y + q[7] * w.z + q[10] * w.w;
              a2 += q[2] * w.x + q[5] * w.y + q[8] * w.z + q[11] * w.w;
            }
            mid[3 * dx] = clip8_q22(a0); mid[3 * dx + 1] = clip8_q22(a1); mid[3 * dx + 2] = clip8_q22(a2);
          } else {
            const int* k = reinterpret_cast<const int*>(k4);
            for (int c = 0; c < p.ch; c++) {
              int a = 1 << (kPrecisionBits - 1);
              for (int t = 0; t < cnt; t++) a += s[t * p.ch + c] * k[t];
              mid[dx * p.ch + c] = clip8_q22(a);
            }
          }
        }
      }
      __syncthreads();
      next_src += nrows;
    }
    // vertical pass: one output row from ycnt intermediate rows
    const int* k = ky + (size_t)dy * p.ksy;
    uint8_t* o = fdst + (size_t)dy * out_row_bytes;
    for (int i4 = threadIdx.x * 4; i4 < mid_row_bytes; i4 += kBicubicThreads * 4) {
      int a[4];
#pragma unroll
      for (int j = 0; j < 4; j++) a[j] = 1 << (kPrecisionBits - 1);
      for (int t = 0; t < ycnt; t++) {
        const int w = __ldg(k + t);
        const uint32_t v = *reinterpret_cast<const uint32_t*>(ring + (size_t)((ymin + t) % p.ring_rows) * mid_stride + i4);
        a[0] += (int)(v & 0xff) * w; a[1] += (int)((v >> 8) & 0xff) * w;
        a[2] += (int)((v >> 16) & 0xff) * w; a[3] += (int)(v >> 24) * w;
      }
      if (i4 + 3 < mid_row_bytes && (reinterpret_cast<uintptr_t>(o) & 3) == 0) {
        *reinterpret_cast<uint32_t*>(o + i4) = (uint32_t)clip8_q22(a[0]) | ((uint32_t)clip8_q22(a[1]) << 8) |
                                               ((uint32_t)clip8_q22(a[2]) << 16) | ((uint32_t)clip8_q22(a[3]) << 24);
      } else {
        for (int j = 0; j < 4 && i4 + j < mid_row_bytes; j++) o[i4 + j] = clip8_q22(a[j]);
      }
    }
    // the next staging round starts with a barrier before any ring row is rewritten, which
    // orders these ring reads before it
  }
}

}  // namespace aiko

using namespace aiko;

extern "C" size_t aiko_b200_resize_bicubic_workspace_bytes(int src_h, int src_w, int dst_h, int dst_w) {
  if (src_h <= 0 || src_w <= 0 || dst_h <= 0 || dst_w <= 0) return 0;
  return bicubic_layout(src_h, src_w, dst_h, dst_w).ints * sizeof(int);
}

extern "C" int aiko_b200_resize_bicubic_prepare(void* workspace, size_t workspace_bytes, int src_h, int src_w,
                                                int dst_h, int dst_w, void* stream) {
  AIKO_CHECK_ARG(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "resize_bicubic_prepare: bad size");
  const BicubicLayout l = bicubic_layout_for(src_h, src_w, dst_h, dst_w, workspace_bytes);
  if (!workspace || workspace_bytes < l.ints * sizeof(int)) {
    set_last_error("resize_bicubic_prepare: workspace of %zu bytes, need %zu", workspace_bytes, l.ints * sizeof(int));
    return AIKO_B200_ERR_WORKSPACE;
  }
  AIKO_CHECK_ARG((reinterpret_cast<uintptr_t>(workspace) & 15) == 0, "resize_bicubic_prepare: workspace not 16-byte aligned");
  int* ws = static_cast<int*>(workspace);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  AIKO_CUDA(cudaMemsetAsync(ws, 0, 4 * sizeof(int), st));   // header; int 1 is the table builders' error flag
  bicubic_coeffs_kernel<<<(dst_w + 127) / 128, 128, 0, st>>>(ws, src_w, dst_w, l.x.ksize, l.x.off_bounds, l.x.off_k);
  bicubic_coeffs_kernel<<<(dst_h + 127) / 128, 128, 0, st>>>(ws, src_h, dst_h, l.y.ksize, l.y.off_bounds, l.y.off_k);
  const int rc = cuda_status(cudaGetLastError(), "bicubic_coeffs launch");
  if (rc != AIKO_B200_OK) return rc;
  return bicubic_mma_prepare(ws, l, src_h, src_w, dst_h, dst_w, st);
}

extern "C" int aiko_b200_resize_bicubic_aa_u8(const uint8_t* src, int n_frames, int src_h, int src_w, int channels,
                                              uint8_t* dst, int dst_h, int dst_w, const void* workspace,
                                              size_t workspace_bytes, void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0, "resize_bicubic: n_frames < 0");
  if (n_frames == 0) return AIKO_B200_OK;
  AIKO_CHECK_ARG(src && dst, "resize_bicubic: null frame pointer");
  AIKO_CHECK_ARG(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "resize_bicubic: bad size");
  AIKO_CHECK_ARG(channels == 1 || channels == 3 || channels == 4, "resize_bicubic: channels %d not in {1,3,4}",
                 channels);
  const BicubicLayout l = bicubic_layout_for(src_h, src_w, dst_h, dst_w, workspace_bytes);
  if (!workspace || workspace_bytes < l.ints * sizeof(int)) {
    set_last_error("resize_bicubic: workspace of %zu bytes, need %zu (and it must have been prepared)", workspace_bytes,
                   l.ints * sizeof(int));
    return AIKO_B200_ERR_WORKSPACE;
  }
  {
    bool done = false;
    const int rc = bicubic_mma_launch(src, n_frames, src_h, src_w, channels, dst, dst_h, dst_w,
                                      static_cast<const int*>(workspace), l, static_cast<cudaStream_t>(stream), &done);
    if (rc != AIKO_B200_OK || done) return rc;
  }
  BicubicParams p{};
  p.src = src; p.dst = dst; p.ws = static_cast<const int*>(workspace);
  p.n = n_frames; p.sh = src_h; p.sw = src_w; p.dh = dst_h; p.dw = dst_w; p.ch = channels;
  p.ksx = l.x.ksize; p.ksy = l.y.ksize;
  p.off_bx = l.x.off_bounds; p.off_kx = l.x.off_k; p.off_by = l.y.off_bounds; p.off_ky = l.y.off_k;
  p.ring_rows = l.y.ksize + kRowsPerRound;
  // bands: long enough that the vertical halo (ksize source rows re-filtered per band) stays small,
  // and a CTA count that fills whole waves of the machine (2 CTAs per SM fit by shared memory)
  const double scale_y = (double)src_h / dst_h;
  const double rows_per_out = scale_y < 1.0 ? 1.0 : scale_y;
  const long long slots = (long long)sm_count() * 2;
  int band = 1;
  double best = -1.0;
  for (int cand = 4; cand <= 64 && cand <= dst_h; cand++) {
    const long long ctas = (long long)n_frames * ((dst_h + cand - 1) / cand);
    const double wave_eff = (double)ctas / (double)(((ctas + slots - 1) / slots) * slots);
    const double halo_eff = cand * rows_per_out / (cand * rows_per_out + l.y.ksize);
    if (wave_eff * halo_eff > best) { best = wave_eff * halo_eff; band = cand; }
  }
  if (dst_h < 4) band = dst_h;
  p.band = band;
  p.bands_per_frame = (dst_h + band - 1) / band;
  // Shared memory per CTA: staged source segments, the ring of intermediate rows, the column tables.
  // When a whole output row does not fit (wide rows under a large vertical reduction), the columns
  // are split into chunks, each with its own CTAs.
  const size_t max_smem = (size_t)device_max_smem_optin();
  const double scale_x = (double)src_w / dst_w, support_x = 2.0 * (scale_x < 1.0 ? 1.0 : scale_x);
  size_t smem = 0;
  for (int chunks = 1;; chunks *= 2) {
    const int chunk_w = (dst_w + chunks - 1) / chunks;
    p.chunk_w = chunk_w;
    p.chunks = (dst_w + chunk_w - 1) / chunk_w;
    long long seg = (long long)((chunk_w - 1) * scale_x + 2.0 * support_x + 3.0) * channels + 16;   // bytes a chunk reads per row
    if (seg > (long long)src_w * channels) seg = (long long)src_w * channels;
    const size_t src_stride = ((size_t)seg + 31) & ~(size_t)15;
    const size_t mid_stride = ((size_t)chunk_w * channels + 15) & ~(size_t)15;
    p.src_stride = (int)src_stride;
    p.mid_stride = (int)mid_stride;
    smem = (size_t)kRowsPerRound * src_stride + 16 + (size_t)p.ring_rows * mid_stride +
           (size_t)((2 * chunk_w + 3) & ~3) * sizeof(int);
    const size_t kx_bytes = (size_t)chunk_w * l.x.ksize * sizeof(int);
    p.kx_in_smem = (smem + kx_bytes <= 100 * 1024) ? 1 : 0;
    if (p.kx_in_smem) smem += kx_bytes;
    if (smem <= max_smem) break;
    if (chunk_w <= 16) {
      set_last_error("resize_bicubic: %dx%d -> %dx%d needs %zu bytes of shared memory for 16 columns (max %zu)", src_w,
                     src_h, dst_w, dst_h, smem, max_smem);
      return AIKO_B200_ERR_UNSUPPORTED;
    }
  }
  AIKO_CUDA(cudaFuncSetAttribute(resize_bicubic_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long grid = (long long)n_frames * p.bands_per_frame * p.chunks;
  AIKO_CHECK_ARG(grid < ((long long)1 << 31), "resize_bicubic: too many bands");
  resize_bicubic_rows<<<(unsigned)grid, kBicubicThreads, smem, static_cast<cudaStream_t>(stream)>>>(p);
  return cuda_status(cudaGetLastError(), "resize_bicubic launch");
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant.
//
// Some users of this library cannot give the bicubic resize a workspace because they cannot
// allocate device memory at the point where they call it: an element whose process_frame runs
// on the pipeline's single event-loop thread (a cudaMalloc there synchronises the device and
// holds up every other stream's frames), a service that handed all device memory to its own
// pool at start-up, or a thin foreign-language binding that can only pass the frame buffers it
// was given.  For them the weight tables live in the fixed array below instead.
//
// The array is part of the loaded library image: every GPU that loads libaiko_b200.so holds
// these kStaticScratchBytes until the process exits, whether or not this entry point is used.
constexpr size_t kStaticScratchBytes = 256 * 1024;
static __device__ __align__(256) int g_bicubic_static_scratch[kStaticScratchBytes / sizeof(int)];

// In host code `g_bicubic_static_scratch` is only a handle for the symbol, not an address a
// kernel could use; the device address is looked up once per device (with that device
// current) and remembered.
static int* g_static_scratch_ptr[64];

static int static_scratch_for_current_device(int** out) {
  int dev = 0;
  AIKO_CUDA(cudaGetDevice(&dev));
  AIKO_CHECK_ARG(dev >= 0 && dev < 64, "resize_bicubic_noalloc: device index %d out of range", dev);
  int* p = __atomic_load_n(&g_static_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    void* addr = nullptr;
    AIKO_CUDA(cudaGetSymbolAddress(&addr, g_bicubic_static_scratch));
    p = static_cast<int*>(addr);
    __atomic_store_n(&g_static_scratch_ptr[dev], p, __ATOMIC_RELEASE);  // same value from any thread
  }
  *out = p;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_resize_bicubic_aa_u8_noalloc(const uint8_t* src, int n_frames, int src_h, int src_w,
                                                      int channels, uint8_t* dst, int dst_h, int dst_w,
                                                      void* stream) {
  AIKO_CHECK_ARG(n_frames >= 0, "resize_bicubic_noalloc: n_frames < 0");
  AIKO_CHECK_ARG(src_h > 0 && src_w > 0 && dst_h > 0 && dst_w > 0, "resize_bicubic_noalloc: bad size");
  // the coefficient tables must fit; the tensor-core tables are used when they fit as well
  const size_t need = bicubic_layout(src_h, src_w, dst_h, dst_w, false).ints * sizeof(int);
  if (need > kStaticScratchBytes) {
    set_last_error(
        "resize_bicubic_noalloc: %dx%d -> %dx%d needs %zu bytes of weight tables but the built-in scratch holds %zu; "
        "use aiko_b200_resize_bicubic_aa_u8 with a workspace of your own",
        src_w, src_h, dst_w, dst_h, need, kStaticScratchBytes);
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  if (n_frames == 0) return AIKO_B200_OK;
  int* scratch = nullptr;
  const int rc = static_scratch_for_current_device(&scratch);
  if (rc != AIKO_B200_OK) return rc;
  // Both tables are rewritten in full (every bound, every weight, padding weights as zero)
  // before the resize reads them, on the same stream: nothing left by an earlier call is ever
  // read, so there is nothing that has to be cleared first.
  const int prc = aiko_b200_resize_bicubic_prepare(scratch, kStaticScratchBytes, src_h, src_w, dst_h, dst_w, stream);
  if (prc != AIKO_B200_OK) return prc;
  return aiko_b200_resize_bicubic_aa_u8(src, n_frames, src_h, src_w, channels, dst, dst_h, dst_w, scratch,
                                        kStaticScratchBytes, stream);
}
