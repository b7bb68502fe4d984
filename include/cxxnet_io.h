/*
 * C ABI of the native data pipeline (cxxnet_b200/lib/libcxxnet_io.so, sources csrc/runtime/).
 * Host code only: no CUDA, no Python.  Counterparts in the reference: utils::BinaryPage / ThreadBuffer
 * (src/utils/io.h:254-326, thread_buffer.h), tools/im2bin.cpp, utils::JpegDecoder (src/utils/decoder.h:21-124),
 * AugmentIterator / ImageAugmenter / BatchAdaptIterator (src/io/iter_augment_proc-inl.hpp, image_augmenter-inl.hpp,
 * iter_batch_proc-inl.hpp).
 */
#ifndef CXXNET_IO_H_
#define CXXNET_IO_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- imgbin pages: blobs of one or more .bin files, pages read ahead by a worker thread */
void* cxnio_reader_open(const char** paths, int npaths, int nbuffers);
/* 1 = *blob / *size describe the next record (valid until the next call), 0 = end of data */
int cxnio_reader_next(void* reader, const void** blob, unsigned long long* size);
/* the next whole page (64 MiB, layout of csrc/runtime/binary_page.h) copied into caller memory; 1, or 0 at the end.
 * Use either this or cxnio_reader_next within one pass. */
int cxnio_reader_next_page(void* reader, void* dst);
void cxnio_reader_rewind(void* reader);
void cxnio_reader_close(void* reader);
/* pack the files of an image list (index <tab> label.. <tab> path, relative to root) into pages; returns the
 * number of images or -1 */
long long cxnio_pack(const char* list_path, const char* root, const char* out_path, int silent);

/* ---- JPEG (baseline / progressive Huffman, 8 bit, grey or YCbCr / RGB, sampling factors 1 or 2).
 * 0 = success; -1 truncated, -2 not a JPEG, -3 unsupported variant, -4 corrupt, -5 size mismatch */
int cxnio_jpeg_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
int cxnio_jpeg_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height);
int cxnio_jpeg_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height);
/* Bytes that differ between the AVX2 inverse DCT and the portable one over nblocks pseudo-random coefficient blocks
 * (must be 0: the vector code performs the same operations in the same order); -1 on a CPU without AVX2.  For tests. */
int cxnio_jpeg_idct_selftest(int nblocks, unsigned seed);
/* `count` images on `nthreads` worker threads; status[i] per image; returns the number of failures */
int cxnio_jpeg_decode_many(int count, const uint8_t* const* bufs, const size_t* sizes, float* const* outs_chw,
                           const int* widths, const int* heights, int* status, int nthreads);

/* ---- PNG (non-interlaced; grey / RGB / palette, bit depths 1-16, alpha dropped); codes as for JPEG */
int cxnio_png_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
int cxnio_png_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height);

/* ---- pre-decoded records ("CXNR", LE uint16 width, height, uint8 channels = 3, 3 zero bytes, then interleaved RGB
 * bytes) next to JPEG and PNG: the functions below take any of them, chosen by the first bytes.  No reference counterpart. */
int cxnio_image_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
int cxnio_image_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height);
/* interleaved RGB bytes -> raw record (out: 12 + width*height*3 bytes); returns the record size, 0 on bad sizes */
size_t cxnio_raw_encode(const uint8_t* rgb_hwc, int width, int height, uint8_t* out);
/* cxnio_pack that decodes every (JPEG) file and stores raw records (im2bin ... raw=1) */
long long cxnio_pack_raw(const char* list_path, const char* root, const char* out_path, int silent);

/* ---- one batch: decode + (affine warp) + crop + mean + contrast / illumination + mirror + scale */
typedef struct CxnioAugItem {
  const uint8_t* buf;    /* encoded image, or NULL when `pixels` is given */
  uint64_t size;
  const float* pixels;   /* decoded planar float image (channels x src_h x src_w), or NULL */
  int32_t src_h, src_w;  /* size of the decoded image (checked against the file) */
  int32_t y0, x0;        /* crop origin */
  int32_t flip;          /* mirror horizontally */
  float contrast, illum; /* v -> (v - mean) * contrast + illum */
  int32_t affine;        /* 1: take the crop from the warped image described by the fields below (then y0 = x0 = 0) */
  int32_t warp_h, warp_w, warp_y, warp_x;
  float fill;            /* value outside the decoded image */
  double ai[6];          /* inverse affine map, two rows: source x = ai[0] x + ai[1] y + ai[2], source y = ai[3..5] */
  int32_t interp;        /* resampling kernel of the warp: 0 bilinear, 1 bicubic (a = -0.75, as cv::INTER_CUBIC) */
  int32_t pad_;
} CxnioAugItem;
/* mean_mode 0 none | 1 mean[c] per channel | 2 mean image of the decoded size (subtracted before crop / mirror) |
 * 3 mean image of the crop size (after).  dst: [n][C][H][W], fp32, (dst_bf16 = 1) bf16 or (dst_bf16 = 2) uint8 - the
 * last for mean_mode 0 / scale 1 only: cropped and mirrored raw pixels.  status[i] = 0 when
 * instance i was written.  Returns the number of instances that were not. */
int cxnio_decode_augment_batch(const CxnioAugItem* items, int n, int C, int H, int W, int mean_mode, const float* mean,
                               int mean_h, int mean_w, float scale, void* dst, int dst_bf16, int* status, int nthreads);
int cxnio_aug_item_bytes(void);

#ifdef __cplusplus
}
#endif
#endif /* CXXNET_IO_H_ */
