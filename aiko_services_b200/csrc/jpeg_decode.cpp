// libaiko_b200_jpeg.so: batched JPEG decode into device-resident RGB frames through nvJPEG
// (include/aiko_b200_jpeg.h).  Host code only; the decode kernels / engine work are nvJPEG's.
#include <cuda_runtime.h>
#include <nvjpeg.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../../include/aiko_b200.h"
#include "../../include/aiko_b200_jpeg.h"

namespace {

thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

const char* status_name(nvjpegStatus_t s) {
  switch (s) {
    case NVJPEG_STATUS_SUCCESS: return "success";
    case NVJPEG_STATUS_NOT_INITIALIZED: return "not initialized";
    case NVJPEG_STATUS_INVALID_PARAMETER: return "invalid parameter";
    case NVJPEG_STATUS_BAD_JPEG: return "bad jpeg";
    case NVJPEG_STATUS_JPEG_NOT_SUPPORTED: return "jpeg not supported";
    case NVJPEG_STATUS_ALLOCATOR_FAILURE: return "allocator failure";
    case NVJPEG_STATUS_EXECUTION_FAILED: return "execution failed";
    case NVJPEG_STATUS_ARCH_MISMATCH: return "arch mismatch";
    case NVJPEG_STATUS_INTERNAL_ERROR: return "internal error";
    case NVJPEG_STATUS_IMPLEMENTATION_NOT_SUPPORTED: return "implementation not supported";
    default: return "unknown status";
  }
}

#define NVJ(call, code)                                                  \
  do {                                                                   \
    nvjpegStatus_t _s = (call);                                          \
    if (_s != NVJPEG_STATUS_SUCCESS) {                                   \
      set_error("%s: %s (%d)", #call, status_name(_s), (int)_s);         \
      return (code);                                                     \
    }                                                                    \
  } while (0)

}  // namespace

// Without the NVJPG engines the decode is split the way nvJPEG's hybrid decoder splits it: the
// entropy (Huffman) decode of each image, serial by nature, runs on a host thread -- one image per
// thread, all host cores busy -- into a page-locked buffer; inverse DCT, chroma upsampling and
// colour conversion run on the GPU.  Each image of a batch has its own nvJPEG state and buffers.
struct JpegSlot {
  nvjpegJpegState_t state = nullptr;
  nvjpegBufferPinned_t pinned = nullptr;
  nvjpegBufferDevice_t device = nullptr;
  nvjpegJpegStream_t stream = nullptr;
};

struct aiko_b200_jpeg_decoder {
  nvjpegHandle_t handle = nullptr;
  nvjpegJpegState_t state = nullptr;      // hardware path: one batched state
  int hardware = 0;
  int max_batch = 0;
  int initialized_batch = 0;       // batch size nvjpegDecodeBatchedInitialize was last called with
  std::vector<nvjpegImage_t> dst;
  // hybrid path
  nvjpegJpegDecoder_t hybrid = nullptr;
  nvjpegDecodeParams_t params = nullptr;
  std::vector<JpegSlot> slots;
  cudaEvent_t drained = nullptr;   // the transfers of the previous batch have left the pinned buffers
  bool in_flight = false;
  int threads = 1;
};

static void destroy_slots(aiko_b200_jpeg_decoder* d) {
  for (JpegSlot& s : d->slots) {
    if (s.state) nvjpegJpegStateDestroy(s.state);
    if (s.pinned) nvjpegBufferPinnedDestroy(s.pinned);
    if (s.device) nvjpegBufferDeviceDestroy(s.device);
    if (s.stream) nvjpegJpegStreamDestroy(s.stream);
  }
  d->slots.clear();
  if (d->params) { nvjpegDecodeParamsDestroy(d->params); d->params = nullptr; }
  if (d->hybrid) { nvjpegDecoderDestroy(d->hybrid); d->hybrid = nullptr; }
  if (d->drained) { cudaEventDestroy(d->drained); d->drained = nullptr; }
}

static int create_hybrid(aiko_b200_jpeg_decoder* d) {
  // chroma is upsampled with interpolation, as libjpeg's "fancy upsampling" (Pillow's decoder) does
  unsigned flags = getenv("AIKO_B200_JPEG_PLAIN_UPSAMPLING") ? NVJPEG_FLAGS_DEFAULT : NVJPEG_FLAGS_UPSAMPLING_WITH_INTERPOLATION;
  nvjpegStatus_t s = nvjpegCreateEx(NVJPEG_BACKEND_DEFAULT, nullptr, nullptr, flags, &d->handle);
  if (s != NVJPEG_STATUS_SUCCESS && flags != NVJPEG_FLAGS_DEFAULT)
    s = nvjpegCreateEx(NVJPEG_BACKEND_DEFAULT, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->handle);
  if (s != NVJPEG_STATUS_SUCCESS) { set_error("nvjpegCreateEx: %s", status_name(s)); d->handle = nullptr; return -1; }
  bool ok = nvjpegDecoderCreate(d->handle, NVJPEG_BACKEND_HYBRID, &d->hybrid) == NVJPEG_STATUS_SUCCESS &&
            nvjpegDecodeParamsCreate(d->handle, &d->params) == NVJPEG_STATUS_SUCCESS &&
            nvjpegDecodeParamsSetOutputFormat(d->params, NVJPEG_OUTPUT_RGBI) == NVJPEG_STATUS_SUCCESS &&
            cudaEventCreateWithFlags(&d->drained, cudaEventDisableTiming) == cudaSuccess;
  d->slots.resize(d->max_batch);
  for (JpegSlot& sl : d->slots) {
    if (!ok) break;
    ok = nvjpegDecoderStateCreate(d->handle, d->hybrid, &sl.state) == NVJPEG_STATUS_SUCCESS &&
         nvjpegBufferPinnedCreate(d->handle, nullptr, &sl.pinned) == NVJPEG_STATUS_SUCCESS &&
         nvjpegBufferDeviceCreate(d->handle, nullptr, &sl.device) == NVJPEG_STATUS_SUCCESS &&
         nvjpegJpegStreamCreate(d->handle, &sl.stream) == NVJPEG_STATUS_SUCCESS &&
         nvjpegStateAttachPinnedBuffer(sl.state, sl.pinned) == NVJPEG_STATUS_SUCCESS &&
         nvjpegStateAttachDeviceBuffer(sl.state, sl.device) == NVJPEG_STATUS_SUCCESS;
  }
  if (!ok) {
    set_error("nvJPEG hybrid decoder: could not create its states / buffers");
    destroy_slots(d);
    nvjpegDestroy(d->handle);
    d->handle = nullptr;
    return -1;
  }
  unsigned hc = std::thread::hardware_concurrency();
  int t = hc ? (int)hc : 4;
  if (const char* e = getenv("AIKO_B200_JPEG_THREADS")) t = atoi(e);
  d->threads = t < 1 ? 1 : (t > d->max_batch ? d->max_batch : t);
  return 0;
}

static int try_create(aiko_b200_jpeg_decoder* d, nvjpegBackend_t backend) {
  nvjpegStatus_t s = nvjpegCreateEx(backend, nullptr, nullptr, NVJPEG_FLAGS_DEFAULT, &d->handle);
  if (s != NVJPEG_STATUS_SUCCESS) {
    set_error("nvjpegCreateEx(backend %d): %s", (int)backend, status_name(s));
    d->handle = nullptr;
    return -1;
  }
  s = nvjpegJpegStateCreate(d->handle, &d->state);
  if (s != NVJPEG_STATUS_SUCCESS) {
    set_error("nvjpegJpegStateCreate: %s", status_name(s));
    nvjpegDestroy(d->handle);
    d->handle = nullptr;
    return -1;
  }
  return 0;
}

extern "C" const char* aiko_b200_jpeg_last_error(void) { return g_error; }

extern "C" int aiko_b200_jpeg_decoder_create(aiko_b200_jpeg_decoder** out, int backend, int max_batch) {
  if (!out || max_batch < 1 || backend < 0 || backend > 2) {
    set_error("jpeg_decoder_create: bad argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  aiko_b200_jpeg_decoder* d = new aiko_b200_jpeg_decoder();
  d->max_batch = max_batch;
  int rc = -1;
  if (backend == 0 || backend == 2) {
    rc = try_create(d, NVJPEG_BACKEND_HARDWARE);
    if (rc == 0) d->hardware = 1;
  }
  if (rc != 0 && backend != 2) rc = create_hybrid(d);
  if (rc != 0) {
    delete d;
    return AIKO_B200_ERR_UNSUPPORTED;
  }
  d->dst.resize(max_batch);
  *out = d;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_jpeg_decoder_destroy(aiko_b200_jpeg_decoder* d) {
  if (!d) return AIKO_B200_OK;
  if (d->in_flight && d->drained) cudaEventSynchronize(d->drained);
  destroy_slots(d);
  if (d->state) nvjpegJpegStateDestroy(d->state);
  if (d->handle) nvjpegDestroy(d->handle);
  delete d;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_jpeg_decoder_is_hardware(const aiko_b200_jpeg_decoder* d) { return d ? d->hardware : 0; }

extern "C" int aiko_b200_jpeg_info(aiko_b200_jpeg_decoder* d, const uint8_t* jpeg, size_t length, int* width,
                                   int* height, int* components, int* subsampling) {
  if (!d || !jpeg || !width || !height) {
    set_error("jpeg_info: null argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  int comps = 0, widths[NVJPEG_MAX_COMPONENT], heights[NVJPEG_MAX_COMPONENT];
  nvjpegChromaSubsampling_t ss;
  NVJ(nvjpegGetImageInfo(d->handle, jpeg, length, &comps, &ss, widths, heights), AIKO_B200_ERR_INVALID_ARGUMENT);
  *width = widths[0];
  *height = heights[0];
  if (components) *components = comps;
  if (subsampling) *subsampling = (int)ss;
  return AIKO_B200_OK;
}

extern "C" int aiko_b200_jpeg_decode_rgb_u8(aiko_b200_jpeg_decoder* d, const uint8_t* const* jpegs,
                                            const size_t* lengths, int n, uint8_t* dst, int height, int width,
                                            int64_t frame_stride, void* stream) {
  if (!d || n < 0 || (n > 0 && (!jpegs || !lengths || !dst)) || height <= 0 || width <= 0 ||
      frame_stride < (int64_t)height * width * 3) {
    set_error("jpeg_decode: bad argument");
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  if (n == 0) return AIKO_B200_OK;
  if (n > d->max_batch) {
    set_error("jpeg_decode: batch of %d exceeds the decoder's max_batch %d", n, d->max_batch);
    return AIKO_B200_ERR_INVALID_ARGUMENT;
  }
  for (int i = 0; i < n; i++) {
    int comps = 0, widths[NVJPEG_MAX_COMPONENT], heights[NVJPEG_MAX_COMPONENT];
    nvjpegChromaSubsampling_t ss;
    NVJ(nvjpegGetImageInfo(d->handle, jpegs[i], lengths[i], &comps, &ss, widths, heights), AIKO_B200_ERR_INVALID_ARGUMENT);
    if (widths[0] != width || heights[0] != height) {
      set_error("jpeg_decode: image %d is %dx%d, the batch is %dx%d", i, widths[0], heights[0], width, height);
      return AIKO_B200_ERR_INVALID_ARGUMENT;
    }
    nvjpegImage_t& im = d->dst[i];
    for (int c = 0; c < NVJPEG_MAX_COMPONENT; c++) { im.channel[c] = nullptr; im.pitch[c] = 0; }
    im.channel[0] = dst + (size_t)i * frame_stride;
    im.pitch[0] = (size_t)width * 3;
  }
  if (!d->hardware) {
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (d->in_flight) {                      // the pinned buffers are about to be rewritten
      cudaEventSynchronize(d->drained);
      d->in_flight = false;
    }
    // phase 1, host threads: parse + entropy decode, one image per thread at a time
    std::atomic<int> next(0), failed(-1);
    std::vector<std::string> errors(n);
    auto work = [&]() {
      for (;;) {
        const int i = next.fetch_add(1);
        if (i >= n) return;
        JpegSlot& sl = d->slots[i];
        nvjpegStatus_t s = nvjpegJpegStreamParse(d->handle, jpegs[i], lengths[i], 0, 0, sl.stream);
        if (s == NVJPEG_STATUS_SUCCESS) s = nvjpegDecodeJpegHost(d->handle, d->hybrid, sl.state, d->params, sl.stream);
        if (s != NVJPEG_STATUS_SUCCESS) {
          errors[i] = status_name(s);
          int expected = -1;
          failed.compare_exchange_strong(expected, i);
        }
      }
    };
    const int nthreads = d->threads < n ? d->threads : n;
    std::vector<std::thread> pool;
    for (int t = 1; t < nthreads; t++) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();
    if (failed.load() >= 0) {
      set_error("jpeg_decode: image %d: %s", failed.load(), errors[failed.load()].c_str());
      return AIKO_B200_ERR_UNSUPPORTED;
    }
    // phase 2, the caller's stream: coefficients to the device, IDCT + upsampling + colour conversion there
    for (int i = 0; i < n; i++) {
      JpegSlot& sl = d->slots[i];
      NVJ(nvjpegDecodeJpegTransferToDevice(d->handle, d->hybrid, sl.state, sl.stream, st), AIKO_B200_ERR_CUDA);
      NVJ(nvjpegDecodeJpegDevice(d->handle, d->hybrid, sl.state, &d->dst[i], st), AIKO_B200_ERR_CUDA);
    }
    cudaEventRecord(d->drained, st);
    d->in_flight = true;
    return AIKO_B200_OK;
  }
  if (d->initialized_batch != n) {
    NVJ(nvjpegDecodeBatchedInitialize(d->handle, d->state, n, 1, NVJPEG_OUTPUT_RGBI), AIKO_B200_ERR_CUDA);
    d->initialized_batch = n;
  }
  nvjpegStatus_t s = nvjpegDecodeBatched(d->handle, d->state, jpegs, lengths, d->dst.data(), static_cast<cudaStream_t>(stream));
  if (s != NVJPEG_STATUS_SUCCESS) {
    d->initialized_batch = 0;          // a failed batch must be re-initialised
    set_error("nvjpegDecodeBatched: %s (%d)", status_name(s), (int)s);
    return s == NVJPEG_STATUS_JPEG_NOT_SUPPORTED || s == NVJPEG_STATUS_BAD_JPEG ? AIKO_B200_ERR_UNSUPPORTED : AIKO_B200_ERR_CUDA;
  }
  return AIKO_B200_OK;
}
