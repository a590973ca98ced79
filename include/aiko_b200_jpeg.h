/*
 * aiko_b200_jpeg.h -- C ABI of libaiko_b200_jpeg.so: JPEG frames decoded on the GPU so that they are
 * BORN device-resident.  Replaces the decode half of
 *   PIL.Image.open(path) ... image.resize(...)   R/elements/media/image_io.py:184,218
 * (Pillow decodes lazily, inside the resize call) for baseline JPEG files / buffers.
 *
 * The decoder is NVIDIA's nvJPEG (CUDA toolkit library; the NVJPG hardware engines when the driver
 * offers that backend, else nvJPEG's CUDA decoder) -- library code, like cuBLAS, not a kernel of this
 * repository.  It lives in its own shared object so that libaiko_b200.so keeps no dependency on
 * libnvjpeg.  Unlike libaiko_b200.so the decoder owns memory: nvJPEG's internal device / pinned
 * buffers belong to the handle (allocated on first use per batch shape, freed by _destroy).
 *
 * Decoded pixels are not bit-identical to Pillow's (libjpeg-turbo): IDCT rounding and, for
 * subsampled chroma, the upsampling filter differ between decoders -- as they do between any two
 * JPEG decoders.  Tolerance used by the tests: mean |diff| <= 1.0 and max |diff| <= 16 on 4:2:0
 * content (tests/test_gpu_jpeg.py); everything downstream of the decoded pixels stays bit-exact.
 */
#ifndef AIKO_B200_JPEG_H_
#define AIKO_B200_JPEG_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct aiko_b200_jpeg_decoder aiko_b200_jpeg_decoder;

/* backend: 0 = best available (hardware engines, else the hybrid decoder), 1 = hybrid decoder
 * (entropy decode on host threads, one image each; IDCT, upsampling and colour conversion on the
 * GPU), 2 = hardware engines only.  max_batch: largest batch a decode call will pass.
 * Returns 0 or a negative AIKO_B200_ERR_* code (aiko_b200_jpeg_last_error() has the text). */
int aiko_b200_jpeg_decoder_create(aiko_b200_jpeg_decoder** decoder_host, int backend, int max_batch);
int aiko_b200_jpeg_decoder_destroy(aiko_b200_jpeg_decoder* decoder);
/* 1 if the decoder runs on the NVJPG hardware engines, 0 if it is the hybrid (host threads + GPU) decoder. */
int aiko_b200_jpeg_decoder_is_hardware(const aiko_b200_jpeg_decoder* decoder);
const char* aiko_b200_jpeg_last_error(void);

/* Width, height, components (1 gray / 3 colour) and chroma subsampling (nvjpegChromaSubsampling_t
 * value) from the header of one JPEG buffer in HOST memory. */
int aiko_b200_jpeg_info(aiko_b200_jpeg_decoder* decoder, const uint8_t* jpeg_host, size_t length, int* width_host,
                        int* height_host, int* components_host, int* subsampling_host);

/* Decodes n JPEG buffers (HOST memory, page-locked or not) into RGB frames on the device:
 * dst[i] = uint8 [height][width][3] at dst + i * frame_stride, interleaved R, G, B -- the layout
 * np.asarray(PIL.Image.open(...).convert("RGB")) has.  Every image must be width x height.
 * Work is enqueued on `stream` (cudaStream_t as void*); the host buffers may be reused when the call
 * returns. */
int aiko_b200_jpeg_decode_rgb_u8(aiko_b200_jpeg_decoder* decoder, const uint8_t* const* jpegs_host,
                                 const size_t* lengths_host, int n, uint8_t* dst, int height, int width,
                                 int64_t frame_stride, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AIKO_B200_JPEG_H_ */
