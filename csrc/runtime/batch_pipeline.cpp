// Decode + augment + batch assembly on native worker threads: the per-instance work of the reference's
// imgbinx decode thread (src/io/iter_thread_imbin_x-inl.hpp:18-395), AugmentIterator::SetData
// (src/io/iter_augment_proc-inl.hpp:98-162) and BatchAdaptIterator (src/io/iter_batch_proc-inl.hpp:16-133),
// done for a whole batch in one call.  The caller (cxxnet_b200/io/iterators.py) draws the random decisions
// in instance order - so results do not depend on the number of threads - and hands over encoded images;
// each worker decodes one image (jpeg_decoder.cpp) and writes its crop, mean-subtracted, jittered, mirrored
// and scaled, straight into that instance's slot of the (pinned) batch buffer, as fp32, bf16 (dst_bf16 = 1) or -
// cropped and mirrored only, normalisation left to the device - 8-bit pixels (dst_bf16 = 2).
//
//   mean_mode 0: out = v * scale                                   (no mean configured)
//             1: out = ((v - mean[c]) * contrast + illum) * scale  (mean_value: one number per channel)
//             2: mean image of the decoded size, subtracted before the crop / mirror
//             3: mean image of the crop size, subtracted after the crop / mirror
// The arithmetic is single precision in this order, the same as the NumPy path it replaces.
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

// JPEG or pre-decoded raw record (raw_image.cpp)
extern "C" int cxnio_image_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
extern "C" int cxnio_image_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height);

extern "C" {

struct CxnioAugItem {
  const uint8_t* buf;     // encoded image, or NULL when `pixels` is given
  uint64_t size;
  const float* pixels;    // already decoded planar float image (channels x src_h x src_w), or NULL
  int32_t src_h, src_w;   // size of the decoded image (checked against the file)
  int32_t y0, x0;         // crop origin
  int32_t flip;           // mirror horizontally
  float contrast, illum;
  // optional affine stage (reference src/io/image_augmenter-inl.hpp: rotation / shear / scale / aspect): the crop is
  // taken from a virtual warp_h x warp_w image V(y, x) = bilinear(decoded image at (ai[0..2] . (x, y, 1),
  // ai[3..5] . (x, y, 1))), positions outside the decoded image reading `fill`; (y0, x0) then refer to the
  // input_shape window of V that starts at (warp_y, warp_x)
  int32_t affine;
  int32_t warp_h, warp_w, warp_y, warp_x;
  float fill;
  double ai[6];
  int32_t interp;         // 0 bilinear, 1 bicubic (Keys kernel, a = -0.75: cv::INTER_CUBIC of the reference)
  int32_t pad_;
};

static inline uint16_t F32ToBf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  // (branch-free, so that the row loops vectorise: both results are computed and one is selected)
  const uint32_t nan = (u >> 16) | 0x40u;                                       // NaN stays NaN
  const uint32_t rne = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16;                  // round to nearest even
  return uint16_t((u & 0x7fffffffu) > 0x7f800000u ? nan : rne);
}

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(CXNIO_NO_CLONES)
#define CXNIO_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define CXNIO_CLONES
#endif

extern "C++" {
// One row of one channel: W outputs.  MODE = mean_mode, DST = 0 fp32 / 1 bf16 / 2 8-bit, FLIP = mirrored read.  All
// per-instance decisions are template parameters, so the loop body is straight-line code the compiler vectorises
// (it was one loop with the branches inside: 0.42 ms per 3 x 227 x 227 instance, 0.8 ms with bf16 output).
// The arithmetic and its order are those of the NumPy path: ((v - mean) * contrast + illum) * scale, single precision,
// no contraction (-ffp-contract=off).
template <int MODE, int DST, bool FLIP>
CXNIO_CLONES static void AugmentRow(const float* srow, const float* mrow, float meanc, float contrast, float illum, float scale,
                                    int W, float* df, uint16_t* dh, uint8_t* db) {
  // bf16 output: the values of a stretch first, then their conversion in a loop of its own (integer operations on a
  // contiguous array: that one vectorises, the combined loop did not)
  constexpr int kStretch = 256;
  float buf[kStretch];
  for (int x0 = 0; x0 < W; x0 += kStretch) {
    const int n = W - x0 < kStretch ? W - x0 : kStretch;
    float* o = DST == 0 ? df + x0 : buf;
    for (int i = 0; i < n; ++i) {
      const int x = x0 + i;
      const int xs = FLIP ? W - 1 - x : x;
      float v = srow[xs];
      if (MODE == 1) v = (v - meanc) * contrast + illum;
      else if (MODE == 2) v = (v - mrow[xs]) * contrast + illum;
      else if (MODE == 3) v = (v - mrow[x]) * contrast + illum;
      o[i] = v * scale;
    }
    if (DST == 1) {
      for (int i = 0; i < n; ++i) dh[x0 + i] = F32ToBf16(buf[i]);
    } else if (DST == 2) {
      for (int i = 0; i < n; ++i) {
        const float v = buf[i];
        db[x0 + i] = uint8_t(v < 0.0f ? 0.0f : (v > 255.0f ? 255.0f : v + 0.5f));   // raw pixels
      }
    }
  }
}
typedef void (*AugmentRowFn)(const float*, const float*, float, float, float, float, int, float*, uint16_t*, uint8_t*);
template <int MODE, int DST>
static AugmentRowFn PickFlip(bool flip) { return flip ? AugmentRow<MODE, DST, true> : AugmentRow<MODE, DST, false>; }
template <int MODE>
static AugmentRowFn PickDst(int dst, bool flip) {
  return dst == 1 ? PickFlip<MODE, 1>(flip) : dst == 2 ? PickFlip<MODE, 2>(flip) : PickFlip<MODE, 0>(flip);
}
static AugmentRowFn PickRow(int mode, int dst, bool flip) {
  switch (mode) {
    case 1: return PickDst<1>(dst, flip);
    case 2: return PickDst<2>(dst, flip);
    case 3: return PickDst<3>(dst, flip);
    default: return PickDst<0>(dst, flip);
  }
}
}  // extern "C++"

// one instance: src (C x sh x sw planar float) -> dst slot
static int AugmentOne(const float* src, int C, int sh, int sw, const CxnioAugItem& it, int H, int W, int mean_mode,
                      const float* mean, int mean_h, int mean_w, float scale, void* dst, int dst_bf16) {
  if (it.y0 < 0 || it.x0 < 0 || it.y0 + H > sh || it.x0 + W > sw) return -10;
  if (mean_mode == 2 && (mean_h != sh || mean_w != sw)) return -11;
  if (mean_mode == 3 && (mean_h != H || mean_w != W)) return -11;
  float* df = static_cast<float*>(dst);
  uint16_t* dh = static_cast<uint16_t*>(dst);
  uint8_t* db = static_cast<uint8_t*>(dst);
  const AugmentRowFn row = PickRow(mean_mode, dst_bf16, it.flip != 0);
  for (int c = 0; c < C; ++c) {
    const float meanc = mean_mode == 1 ? mean[c] : 0.0f;
    for (int y = 0; y < H; ++y) {
      const float* srow = src + (size_t(c) * sh + it.y0 + y) * sw + it.x0;
      const float* mrow = nullptr;
      if (mean_mode == 2) mrow = mean + (size_t(c) * sh + it.y0 + y) * sw + it.x0;
      if (mean_mode == 3) mrow = mean + (size_t(c) * H + y) * W;
      const size_t o = (size_t(c) * H + y) * W;
      row(srow, mrow, meanc, it.contrast, it.illum, scale, W, df + o, dh + o, db + o);
    }
  }
  return 0;
}

// the four Keys (a = -0.75) weights of taps -1, 0, 1, 2 for fraction t; the expressions of ImageAugmenter.cubic_weights
static inline void CubicWeights(float t, float* c) {
  const float A = -0.75f;
  const float t1 = t + 1.0f;
  c[0] = ((A * t1 - 5.0f * A) * t1 + 8.0f * A) * t1 - 4.0f * A;
  c[1] = ((A + 2.0f) * t - (A + 3.0f)) * t * t + 1.0f;
  const float u = 1.0f - t;
  c[2] = ((A + 2.0f) * u - (A + 3.0f)) * u * u + 1.0f;
  c[3] = 1.0f - c[0] - c[1] - c[2];
}

// the H x W window at (wy, wx) of the affine-warped image, channel by channel, into `out` (C x H x W)
CXNIO_CLONES static void WarpWindow(const float* src, int C, int sh, int sw, const CxnioAugItem& it, int H, int W,
                                    float* out) {
  const double a00 = it.ai[0], a01 = it.ai[1], a02 = it.ai[2], a10 = it.ai[3], a11 = it.ai[4], a12 = it.ai[5];
  const float fill = it.fill;
  for (int y = 0; y < H; ++y) {
    const int yy = it.warp_y + y;
    for (int x = 0; x < W; ++x) {
      const int xx = it.warp_x + x;
      const double sx = a00 * xx + a01 * yy + a02, sy = a10 * xx + a11 * yy + a12;
      const double fxd = std::floor(sx), fyd = std::floor(sy);
      const long x0 = long(fxd), y0 = long(fyd);
      const float fx = float(sx - fxd), fy = float(sy - fyd);
      const bool okx0 = x0 >= 0 && x0 < sw, okx1 = x0 + 1 >= 0 && x0 + 1 < sw;
      const bool oky0 = y0 >= 0 && y0 < sh, oky1 = y0 + 1 >= 0 && y0 + 1 < sh;
      if (it.interp == 1) {
        // bicubic: 4 x 4 taps around (y0, x0), weights and summation order as ImageAugmenter.warp (io/iterators.py)
        float cx[4], cy[4];
        CubicWeights(fx, cx);
        CubicWeights(fy, cy);
        for (int c = 0; c < C; ++c) {
          const float* pl = src + size_t(c) * sh * sw;
          float v = 0.0f;
          for (int j = 0; j < 4; ++j) {
            const long yr = y0 - 1 + j;
            const bool oky = yr >= 0 && yr < sh;
            float t[4];
            for (int i = 0; i < 4; ++i) {
              const long xc = x0 - 1 + i;
              t[i] = (oky && xc >= 0 && xc < sw) ? pl[yr * sw + xc] : fill;
            }
            const float r = ((t[0] * cx[0] + t[1] * cx[1]) + t[2] * cx[2]) + t[3] * cx[3];
            v = j == 0 ? r * cy[0] : v + r * cy[j];
          }
          out[(size_t(c) * H + y) * W + x] = v;
        }
        continue;
      }
      const float w00 = (1 - fx), w01 = fx;
      for (int c = 0; c < C; ++c) {
        const float* pl = src + size_t(c) * sh * sw;
        const float t00 = (oky0 && okx0) ? pl[y0 * sw + x0] : fill;
        const float t01 = (oky0 && okx1) ? pl[y0 * sw + x0 + 1] : fill;
        const float t10 = (oky1 && okx0) ? pl[(y0 + 1) * sw + x0] : fill;
        const float t11 = (oky1 && okx1) ? pl[(y0 + 1) * sw + x0 + 1] : fill;
        // the NumPy statement's order: ((t00 (1-fx) (1-fy) + t01 fx (1-fy)) + t10 (1-fx) fy) + t11 fx fy
        const float v = ((t00 * w00 * (1 - fy) + t01 * w01 * (1 - fy)) + t10 * w00 * fy) + t11 * w01 * fy;
        out[(size_t(c) * H + y) * W + x] = v;
      }
    }
  }
}

// crop / mean / jitter / mirror / scale of one decoded image, through the affine stage when the item asks for it
static int ProcessOne(const float* src, int C, int sh, int sw, const CxnioAugItem& it, int H, int W, int mean_mode,
                      const float* mean, int mean_h, int mean_w, float scale, void* dst, int dst_bf16,
                      std::vector<float>& win) {
  if (!it.affine)
    return AugmentOne(src, C, sh, sw, it, H, W, mean_mode, mean, mean_h, mean_w, scale, dst, dst_bf16);
  if (it.warp_y < 0 || it.warp_x < 0 || it.warp_y + H > it.warp_h || it.warp_x + W > it.warp_w || it.y0 || it.x0)
    return -13;
  win.resize(size_t(C) * H * W);
  WarpWindow(src, C, sh, sw, it, H, W, win.data());
  return AugmentOne(win.data(), C, H, W, it, H, W, mean_mode, mean, mean_h, mean_w, scale, dst, dst_bf16);
}

// status[i]: 0 ok, < 0 this instance was not written (undecodable here, size mismatch, crop outside the image)
int cxnio_decode_augment_batch(const CxnioAugItem* items, int n, int C, int H, int W, int mean_mode, const float* mean,
                               int mean_h, int mean_w, float scale, void* dst, int dst_bf16, int* status,
                               int nthreads) {
  if (n <= 0) return 0;
  if (nthreads < 1) nthreads = 1;
  if (nthreads > n) nthreads = n;
  const size_t inst = size_t(C) * H * W * (dst_bf16 == 1 ? 2 : dst_bf16 == 2 ? 1 : 4);
  std::atomic<int> next{0};
  auto work = [&]() {
    std::vector<float> tmp, win;
    for (;;) {
      const int i = next.fetch_add(1);
      if (i >= n) return;
      const CxnioAugItem& it = items[i];
      void* slot = static_cast<char*>(dst) + inst * i;
      if (it.pixels) {
        status[i] = ProcessOne(it.pixels, C, it.src_h, it.src_w, it, H, W, mean_mode, mean, mean_h, mean_w, scale,
                               slot, dst_bf16, win);
        continue;
      }
      int w = 0, h = 0, comps = 0;
      int r = cxnio_image_info(it.buf, it.size, &w, &h, &comps);
      if (r == 0 && (w != it.src_w || h != it.src_h || (C != 3 && C != 1))) r = -12;
      if (r == 0) {
        tmp.resize(size_t(3) * h * w);
        r = cxnio_image_decode_planar_f32(it.buf, it.size, tmp.data(), w, h);
      }
      if (r == 0 && C == 1) {
        // a single-channel net: the mean of the three decoded planes (for a grey-scale file, the image)
        const size_t plane = size_t(h) * w;
        float* t = tmp.data();
        for (size_t i = 0; i < plane; ++i) t[i] = ((t[i] + t[plane + i]) + t[2 * plane + i]) / 3.0f;
      }
      if (r == 0) r = ProcessOne(tmp.data(), C, h, w, it, H, W, mean_mode, mean, mean_h, mean_w, scale, slot, dst_bf16, win);
      status[i] = r;
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nthreads; ++t) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  int bad = 0;
  for (int i = 0; i < n; ++i) bad += status[i] != 0;
  return bad;
}

int cxnio_aug_item_bytes() { return int(sizeof(CxnioAugItem)); }

}  // extern "C"
