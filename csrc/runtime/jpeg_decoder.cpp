// In-memory JPEG decoder for the image iterators (the role of the reference's utils::JpegDecoder,
// src/utils/decoder.h:21-124, which wraps libjpeg; libjpeg's headers are not available here, so this is a
// self-contained implementation).  Sequential (baseline / extended, 8 bit) and progressive Huffman JPEG,
// 1 or 3 components, sampling factors 1 or 2, restart intervals.  The output stages follow the usual
// libjpeg defaults so that decoded pixels agree with it to within rounding: triangle-filter ("fancy")
// chroma up-sampling for 2x1 and 2x2, fixed-point YCbCr -> RGB.  Anything else (arithmetic coding, 12 bit,
// CMYK, lossless) returns an error code and the caller falls back to another decoder.
//
// C ABI (ctypes, cxxnet_b200/io/native.py):
//   int cxnio_jpeg_info(const uint8_t* buf, size_t n, int* width, int* height, int* components)
//   int cxnio_jpeg_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height)
//   int cxnio_jpeg_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height)
//   int cxnio_jpeg_decode_many(int count, const uint8_t* const* bufs, const size_t* sizes,
//                              float* const* outs_chw, const int* widths, const int* heights,
//                              int* status, int nthreads)        // worker threads, one image per task
// return 0 on success, a negative code otherwise.
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>
#if defined(__x86_64__) && defined(__GNUC__) && !defined(CXNIO_NO_CLONES)
#include <immintrin.h>
#endif

namespace cxnio {
namespace jpeg {

enum Err { kOk = 0, kTruncated = -1, kNotJpeg = -2, kUnsupported = -3, kCorrupt = -4, kSizeMismatch = -5 };

static const uint8_t kZigzag[64] = {0,  1,  8,  16, 9,  2,  3,  10, 17, 24, 32, 25, 18, 11, 4,  5,
                                    12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6,  7,  14, 21, 28,
                                    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51,
                                    58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63};

struct HuffTable {
  bool present = false;
  uint8_t fast_len[512];   // 9-bit prefix -> code length (0 = longer than 9 bits)
  uint8_t fast_sym[512];
  int32_t maxcode[18];     // largest code of each length, left-aligned to 16 bits, +1
  int32_t valoff[17];      // symbol index = code + valoff[len]
  uint8_t syms[256];
  // AC tables only: 9-bit prefix -> (value << 8) | (run << 4) | (code bits + magnitude bits) when a whole
  // run/size code and its magnitude bits fit into the prefix, else 0
  int16_t fast_ac[512];
  void BuildFastAc() {
    for (int i = 0; i < 512; ++i) {
      fast_ac[i] = 0;
      const int len = fast_len[i];
      if (!len) continue;
      const int rs = fast_sym[i], run = rs >> 4, mag = rs & 15;
      if (mag == 0 || len + mag > 9) continue;
      int v = ((i << len) & 511) >> (9 - mag);
      if (v < (1 << (mag - 1))) v += 1 - (1 << mag);
      if (v >= -128 && v <= 127) fast_ac[i] = int16_t(v * 256 + run * 16 + len + mag);
    }
  }
  bool Build(const uint8_t counts[16], const uint8_t* symbols, int nsym) {
    memcpy(syms, symbols, nsym);
    memset(fast_len, 0, sizeof(fast_len));
    int code = 0, k = 0;
    for (int len = 1; len <= 16; ++len) {
      valoff[len] = k - code;
      if (code + counts[len - 1] > (1 << len) || k + counts[len - 1] > nsym) return false;   // over-subscribed
      for (int i = 0; i < counts[len - 1]; ++i, ++k, ++code) {
        if (len <= 9) {
          const int first = code << (9 - len);
          for (int j = 0; j < (1 << (9 - len)); ++j) { fast_len[first + j] = uint8_t(len); fast_sym[first + j] = syms[k]; }
        }
      }
      maxcode[len] = code << (16 - len);
      code <<= 1;
    }
    maxcode[17] = 0x7fffffff;
    present = true;
    return k == nsym;
  }
};

struct BitReader {
  const uint8_t* p;
  const uint8_t* end;
  uint32_t acc = 0;
  int nbits = 0;
  int marker = 0;          // a marker met inside entropy data (0xD0.. etc.), 0 if none
  void Reset(const uint8_t* b, const uint8_t* e) { p = b; end = e; acc = 0; nbits = 0; marker = 0; }
  inline void Fill() {
    while (nbits <= 24) {
      uint32_t byte = 0;
      if (!marker && p < end) {
        byte = *p++;
        if (byte == 0xFF) {
          uint8_t nx = p < end ? *p : 0xD9;
          while (nx == 0xFF && p + 1 < end) nx = *++p;       // fill bytes
          if (nx == 0) { ++p; }
          else { marker = nx; byte = 0; ++p; }
        }
      }
      acc |= byte << (24 - nbits);
      nbits += 8;
    }
  }
  inline int Peek(int n) { return int(acc >> (32 - n)); }
  inline void Skip(int n) { acc <<= n; nbits -= n; }
  inline int Bits(int n) { if (n == 0) return 0; if (nbits < n) Fill(); int v = Peek(n); Skip(n); return v; }
  inline int Bit() { if (nbits < 1) Fill(); int v = int(acc >> 31); Skip(1); return v; }
  // value of an n-bit magnitude category ("receive and extend")
  inline int Extend(int n) { if (n == 0) return 0; int v = Bits(n); return v < (1 << (n - 1)) ? v - (1 << n) + 1 : v; }
  inline int Decode(const HuffTable& h) {
    if (nbits < 16) Fill();
    const int look = Peek(9);
    const int l = h.fast_len[look];
    if (l) { Skip(l); return h.fast_sym[look]; }
    const int32_t code16 = int32_t(acc >> 16);
    int len = 10;
    while (code16 >= h.maxcode[len]) ++len;
    if (len > 16) return -1;
    const int idx = (code16 >> (16 - len)) + h.valoff[len];
    Skip(len);
    return (idx >= 0 && idx < 256) ? h.syms[idx] : -1;
  }
};

struct Component {
  int id = 0, h = 1, v = 1, tq = 0, td = 0, ta = 0;
  int bw = 0, bh = 0;          // blocks per row / column, padded to whole MCUs
  int w = 0, hgt = 0;          // real down-sampled size in samples
  int pred = 0;
  std::vector<int16_t> coef;   // progressive: bw*bh*64 coefficients
  std::vector<uint8_t> plane;  // bw*8 x bh*8 samples
};

// scale factors of the AAN factorisation; folded into the de-quantisation multipliers together with the 1/8
static const float kAanScale[8] = {1.0f, 1.387039845f, 1.306562965f, 1.175875602f, 1.0f, 0.785694958f, 0.541196100f,
                                   0.275899379f};

// Hot loops are compiled twice (AVX2 and baseline x86-64) and chosen at load time.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(CXNIO_NO_CLONES)
#define CXNIO_CLONES __attribute__((target_clones("avx2", "default")))
#else
#define CXNIO_CLONES
#endif

// Eight 8-point inverse DCTs side by side (Arai-Agui-Nakajima, 5 multiplications each): element k of
// transform `lane` is in[k * 8 + lane], results go to out[k * 8 + lane]; the lane loop vectorises.
static inline void Idct8Lanes(const float* in, float* out) {
  for (int l = 0; l < 8; ++l) {
    float t0 = in[l], t1 = in[16 + l], t2 = in[32 + l], t3 = in[48 + l];
    float t10 = t0 + t2, t11 = t0 - t2;
    float t13 = t1 + t3, t12 = (t1 - t3) * 1.414213562f - t13;
    t0 = t10 + t13; t3 = t10 - t13; t1 = t11 + t12; t2 = t11 - t12;
    float t4 = in[8 + l], t5 = in[24 + l], t6 = in[40 + l], t7 = in[56 + l];
    const float z13 = t6 + t5, z10 = t6 - t5, z11 = t4 + t7, z12 = t4 - t7;
    t7 = z11 + z13;
    t11 = (z11 - z13) * 1.414213562f;
    const float z5 = (z10 + z12) * 1.847759065f;
    t10 = 1.082392200f * z12 - z5;
    t12 = -2.613125930f * z10 + z5;
    t6 = t12 - t7; t5 = t11 - t6; t4 = t10 + t5;
    out[l] = t0 + t7; out[56 + l] = t0 - t7; out[8 + l] = t1 + t6; out[48 + l] = t1 - t6;
    out[16 + l] = t2 + t5; out[40 + l] = t2 - t5; out[32 + l] = t3 + t4; out[24 + l] = t3 - t4;
  }
}

#if defined(__x86_64__) && defined(__GNUC__) && !defined(CXNIO_NO_CLONES)
#define CXNIO_HAVE_AVX2_IDCT 1
// The same 8x8 inverse DCT with explicit AVX2 vectors: a row of the block is one register, both 1-D passes are the
// operations of Idct8Lanes IN THE SAME ORDER (no fused multiply-adds: the result is bit-identical to the portable
// code below, cxnio_jpeg_idct_selftest checks that), the two transposes are register shuffles instead of 64 scalar
// moves each, and the level shift / clamp / byte conversion handles a row per instruction.  This was the largest
// single cost of a decode (about 330 ns of the ~600 ns a block takes).
#define CXNIO_AVX2 __attribute__((target("avx2")))
CXNIO_AVX2 static inline void Idct8Rows(__m256 (&r)[8]) {
  const __m256 k1414 = _mm256_set1_ps(1.414213562f), k1847 = _mm256_set1_ps(1.847759065f);
  const __m256 k1082 = _mm256_set1_ps(1.082392200f), km2613 = _mm256_set1_ps(-2.613125930f);
  __m256 t0 = r[0], t1 = r[2], t2 = r[4], t3 = r[6];
  __m256 t10 = _mm256_add_ps(t0, t2), t11 = _mm256_sub_ps(t0, t2);
  __m256 t13 = _mm256_add_ps(t1, t3);
  __m256 t12 = _mm256_sub_ps(_mm256_mul_ps(_mm256_sub_ps(t1, t3), k1414), t13);
  t0 = _mm256_add_ps(t10, t13); t3 = _mm256_sub_ps(t10, t13); t1 = _mm256_add_ps(t11, t12); t2 = _mm256_sub_ps(t11, t12);
  __m256 t4 = r[1], t5 = r[3], t6 = r[5], t7 = r[7];
  const __m256 z13 = _mm256_add_ps(t6, t5), z10 = _mm256_sub_ps(t6, t5), z11 = _mm256_add_ps(t4, t7), z12 = _mm256_sub_ps(t4, t7);
  t7 = _mm256_add_ps(z11, z13);
  t11 = _mm256_mul_ps(_mm256_sub_ps(z11, z13), k1414);
  const __m256 z5 = _mm256_mul_ps(_mm256_add_ps(z10, z12), k1847);
  t10 = _mm256_sub_ps(_mm256_mul_ps(k1082, z12), z5);
  t12 = _mm256_add_ps(_mm256_mul_ps(km2613, z10), z5);
  t6 = _mm256_sub_ps(t12, t7); t5 = _mm256_sub_ps(t11, t6); t4 = _mm256_add_ps(t10, t5);
  r[0] = _mm256_add_ps(t0, t7); r[7] = _mm256_sub_ps(t0, t7); r[1] = _mm256_add_ps(t1, t6); r[6] = _mm256_sub_ps(t1, t6);
  r[2] = _mm256_add_ps(t2, t5); r[5] = _mm256_sub_ps(t2, t5); r[4] = _mm256_add_ps(t3, t4); r[3] = _mm256_sub_ps(t3, t4);
}
CXNIO_AVX2 static inline void Transpose8(__m256 (&r)[8]) {
  const __m256 a0 = _mm256_unpacklo_ps(r[0], r[1]), a1 = _mm256_unpackhi_ps(r[0], r[1]);
  const __m256 a2 = _mm256_unpacklo_ps(r[2], r[3]), a3 = _mm256_unpackhi_ps(r[2], r[3]);
  const __m256 a4 = _mm256_unpacklo_ps(r[4], r[5]), a5 = _mm256_unpackhi_ps(r[4], r[5]);
  const __m256 a6 = _mm256_unpacklo_ps(r[6], r[7]), a7 = _mm256_unpackhi_ps(r[6], r[7]);
  const __m256 b0 = _mm256_shuffle_ps(a0, a2, 0x44), b1 = _mm256_shuffle_ps(a0, a2, 0xEE);
  const __m256 b2 = _mm256_shuffle_ps(a1, a3, 0x44), b3 = _mm256_shuffle_ps(a1, a3, 0xEE);
  const __m256 b4 = _mm256_shuffle_ps(a4, a6, 0x44), b5 = _mm256_shuffle_ps(a4, a6, 0xEE);
  const __m256 b6 = _mm256_shuffle_ps(a5, a7, 0x44), b7 = _mm256_shuffle_ps(a5, a7, 0xEE);
  r[0] = _mm256_permute2f128_ps(b0, b4, 0x20); r[1] = _mm256_permute2f128_ps(b1, b5, 0x20);
  r[2] = _mm256_permute2f128_ps(b2, b6, 0x20); r[3] = _mm256_permute2f128_ps(b3, b7, 0x20);
  r[4] = _mm256_permute2f128_ps(b0, b4, 0x31); r[5] = _mm256_permute2f128_ps(b1, b5, 0x31);
  r[6] = _mm256_permute2f128_ps(b2, b6, 0x31); r[7] = _mm256_permute2f128_ps(b3, b7, 0x31);
}
CXNIO_AVX2 static void Idct8x8Avx2(const float* in, uint8_t* out, int stride) {
  __m256 r[8];
  for (int i = 0; i < 8; ++i) r[i] = _mm256_loadu_ps(in + 8 * i);   // row v of the coefficients, lanes = u
  Idct8Rows(r);        // along v for every column u: r[y] lane u
  Transpose8(r);       // r[u] lane y
  Idct8Rows(r);        // along u for every row y: r[x] lane y
  Transpose8(r);       // r[y] lane x
  const __m256 k128 = _mm256_set1_ps(128.5f), k1024 = _mm256_set1_ps(1024.0f);
  const __m256i i1024 = _mm256_set1_epi32(1024), i255 = _mm256_set1_epi32(255), zero = _mm256_setzero_si256();
  for (int y = 0; y < 8; ++y) {
    // int(a + 128.5f + 1024.0f) - 1024, clamped to [0, 255]
    __m256i v = _mm256_cvttps_epi32(_mm256_add_ps(_mm256_add_ps(r[y], k128), k1024));
    v = _mm256_min_epi32(_mm256_max_epi32(_mm256_sub_epi32(v, i1024), zero), i255);
    const __m128i lo = _mm256_castsi256_si128(v), hi = _mm256_extracti128_si256(v, 1);
    const __m128i w = _mm_packus_epi32(lo, hi);            // 8 x u16
    const __m128i b = _mm_packus_epi16(w, w);              // 8 x u8 (twice)
    _mm_storel_epi64(reinterpret_cast<__m128i*>(out + y * stride), b);
  }
}
static const bool g_has_avx2 = __builtin_cpu_supports("avx2");
#endif

static void Idct8x8Portable(const float* in, uint8_t* out, int stride);

// 8x8 inverse DCT of coefficients already multiplied by their (AAN-scaled) quantisation steps, natural order
// -> samples, level shift +128, clamp.  `dc_only`: every AC coefficient is zero.
static inline void Idct8x8(const float* in, uint8_t* out, int stride, bool dc_only) {
  if (dc_only) {
    int v = int(in[0] + 128.5f + 1024.0f) - 1024;
    v = v < 0 ? 0 : (v > 255 ? 255 : v);
    for (int y = 0; y < 8; ++y) memset(out + y * stride, v, 8);
    return;
  }
#ifdef CXNIO_HAVE_AVX2_IDCT
  if (g_has_avx2) { Idct8x8Avx2(in, out, stride); return; }
#endif
  Idct8x8Portable(in, out, stride);
}

CXNIO_CLONES static void Idct8x8Portable(const float* in, uint8_t* out, int stride) {
  float a[64], b[64];
  Idct8Lanes(in, a);                                   // along v for every column u: a[y * 8 + u]
  for (int y = 0; y < 8; ++y)
    for (int u = 0; u < 8; ++u) b[u * 8 + y] = a[y * 8 + u];
  Idct8Lanes(b, a);                                    // along u for every row y: a[x * 8 + y]
  for (int y = 0; y < 8; ++y) {
    uint8_t* o = out + y * stride;
    for (int x = 0; x < 8; ++x) {
      int v = int(a[x * 8 + y] + 128.5f + 1024.0f) - 1024;      // floor for the value range of interest
      o[x] = uint8_t(v < 0 ? 0 : (v > 255 ? 255 : v));
    }
  }
}

// one row of YCbCr samples -> interleaved RGB bytes (libjpeg's fixed-point constants)
CXNIO_CLONES static void YccRowToRgb(const uint8_t* yy, const uint8_t* cb, const uint8_t* cr, uint8_t* px, int width) {
  for (int x = 0; x < width; ++x) {
    const int Y = yy[x], u = cb[x] - 128, v = cr[x] - 128;
    int r = Y + ((91881 * v + 32768) >> 16);
    int g = Y + ((-22554 * u - 46802 * v + 32768) >> 16);
    int b = Y + ((116130 * u + 32768) >> 16);
    r = r < 0 ? 0 : (r > 255 ? 255 : r);
    g = g < 0 ? 0 : (g > 255 ? 255 : g);
    b = b < 0 ? 0 : (b > 255 ? 255 : b);
    px[3 * x] = uint8_t(r); px[3 * x + 1] = uint8_t(g); px[3 * x + 2] = uint8_t(b);
  }
}

// one row of YCbCr samples -> three planar float rows
CXNIO_CLONES static void YccRowToPlanar(const uint8_t* yy, const uint8_t* cb, const uint8_t* cr, float* o0, float* o1,
                                        float* o2, int width) {
  for (int x = 0; x < width; ++x) {
    const int Y = yy[x], u = cb[x] - 128, v = cr[x] - 128;
    int r = Y + ((91881 * v + 32768) >> 16);
    int g = Y + ((-22554 * u - 46802 * v + 32768) >> 16);
    int b = Y + ((116130 * u + 32768) >> 16);
    o0[x] = float(r < 0 ? 0 : (r > 255 ? 255 : r));
    o1[x] = float(g < 0 ? 0 : (g > 255 ? 255 : g));
    o2[x] = float(b < 0 ? 0 : (b > 255 ? 255 : b));
  }
}

class Decoder {
 public:
  int width = 0, height = 0, ncomp = 0;

  int ParseHeader(const uint8_t* buf, size_t n) {
    b_ = buf; e_ = buf + n;
    if (n < 4 || buf[0] != 0xFF || buf[1] != 0xD8) return kNotJpeg;
    pos_ = buf + 2;
    for (;;) {
      int m = NextMarker();
      if (m < 0) return kTruncated;
      if (m == 0xD9) return kCorrupt;
      if (m >= 0xC0 && m <= 0xCF && m != 0xC4 && m != 0xC8 && m != 0xCC) {
        if (m != 0xC0 && m != 0xC1 && m != 0xC2) return kUnsupported;
        progressive_ = (m == 0xC2);
        return ParseSOF();
      }
      int r = SkipOrParseSegment(m);
      if (r != kOk) return r;
    }
  }

  // decodes everything after the frame header; rgb: interleaved uint8 (may be null), planar: float CHW (may be null)
  int Decode(uint8_t* rgb, float* planar) {
    for (;;) {
      int m = NextMarker();
      if (m < 0) break;                    // truncated file: use what was decoded
      if (m == 0xD9) break;
      if (m == 0xDA) {
        int r = ParseScan();
        if (r != kOk) return r;
        continue;
      }
      int r = SkipOrParseSegment(m);
      if (r != kOk) return r;
    }
    if (!seen_scan_) return kCorrupt;
    if (progressive_) FinishProgressive();
    Output(rgb, planar);
    return kOk;
  }

 private:
  const uint8_t *b_ = nullptr, *e_ = nullptr, *pos_ = nullptr;
  bool progressive_ = false, seen_scan_ = false;
  int restart_interval_ = 0, hmax_ = 1, vmax_ = 1, mcux_ = 0, mcuy_ = 0;
  uint16_t qt_[4][64];
  float qf_[4][64];        // quantisation step x AAN scale / 8, natural order
  bool qt_present_[4] = {false, false, false, false};
  HuffTable dc_[4], ac_[4];
  Component comp_[3];
  int eobrun_ = 0;
  int adobe_transform_ = -1;

  int NextMarker() {
    while (pos_ + 1 < e_) {
      if (pos_[0] != 0xFF) { ++pos_; continue; }
      uint8_t m = pos_[1];
      if (m == 0xFF) { ++pos_; continue; }
      pos_ += 2;
      if (m == 0x00) continue;
      return m;
    }
    return -1;
  }
  int SegLen() {
    if (pos_ + 2 > e_) return -1;
    int len = (pos_[0] << 8) | pos_[1];
    if (len < 2 || pos_ + len > e_) return -1;
    return len;
  }
  int SkipOrParseSegment(int m) {
    if ((m >= 0xD0 && m <= 0xD8) || m == 0x01) return kOk;      // markers without a length field
    const int len = SegLen();
    if (len < 0) return kTruncated;
    const uint8_t* p = pos_ + 2;
    const uint8_t* end = pos_ + len;
    pos_ = end;
    if (m == 0xDB) {            // quantisation tables
      while (p < end) {
        const int pq = p[0] >> 4, tq = p[0] & 15;
        ++p;
        if (tq > 3) return kCorrupt;
        if (p + (pq ? 128 : 64) > end) return kCorrupt;
        for (int i = 0; i < 64; ++i) {
          qt_[tq][kZigzag[i]] = pq ? uint16_t((p[0] << 8) | p[1]) : p[0];
          p += pq ? 2 : 1;
        }
        for (int k = 0; k < 64; ++k) qf_[tq][k] = float(qt_[tq][k]) * kAanScale[k >> 3] * kAanScale[k & 7] * 0.125f;
        qt_present_[tq] = true;
      }
    } else if (m == 0xC4) {     // Huffman tables
      while (p < end) {
        const int tc = p[0] >> 4, th = p[0] & 15;
        if (tc > 1 || th > 3 || p + 17 > end) return kCorrupt;
        int nsym = 0;
        for (int i = 0; i < 16; ++i) nsym += p[1 + i];
        if (nsym > 256 || p + 17 + nsym > end) return kCorrupt;
        HuffTable& h = tc ? ac_[th] : dc_[th];
        if (!h.Build(p + 1, p + 17, nsym)) return kCorrupt;
        if (tc) h.BuildFastAc();
        p += 17 + nsym;
      }
    } else if (m == 0xDD) {     // restart interval
      if (len < 4) return kCorrupt;
      restart_interval_ = (p[0] << 8) | p[1];
    } else if (m == 0xCC) {
      return kUnsupported;      // arithmetic conditioning
    } else if (m == 0xEE && len >= 14 && memcmp(p, "Adobe", 5) == 0) {
      adobe_transform_ = p[11];  // 0: components are RGB, 1: YCbCr
    }
    return kOk;
  }
  int ParseSOF() {
    const int len = SegLen();
    if (len < 8) return kTruncated;
    const uint8_t* p = pos_ + 2;
    pos_ += len;
    if (p[0] != 8) return kUnsupported;
    height = (p[1] << 8) | p[2];
    width = (p[3] << 8) | p[4];
    ncomp = p[5];
    if (width <= 0 || height <= 0) return kUnsupported;       // height 0 = DNL, not handled
    if (int64_t(width) * height > (int64_t(1) << 26)) return kUnsupported;   // keeps a damaged header from asking for gigabytes
    if (ncomp != 1 && ncomp != 3) return kUnsupported;
    if (len < 8 + 3 * ncomp) return kCorrupt;
    for (int i = 0; i < ncomp; ++i) {
      Component& c = comp_[i];
      c.id = p[6 + 3 * i];
      c.h = p[7 + 3 * i] >> 4;
      c.v = p[7 + 3 * i] & 15;
      c.tq = p[8 + 3 * i];
      if (c.h < 1 || c.h > 2 || c.v < 1 || c.v > 2 || c.tq > 3) return kUnsupported;
      if (c.h > hmax_) hmax_ = c.h;
      if (c.v > vmax_) vmax_ = c.v;
    }
    if (ncomp == 1) { comp_[0].h = comp_[0].v = 1; hmax_ = vmax_ = 1; }
    mcux_ = (width + 8 * hmax_ - 1) / (8 * hmax_);
    mcuy_ = (height + 8 * vmax_ - 1) / (8 * vmax_);
    for (int i = 0; i < ncomp; ++i) {
      Component& c = comp_[i];
      if (hmax_ % c.h || vmax_ % c.v) return kUnsupported;
      c.bw = mcux_ * c.h;
      c.bh = mcuy_ * c.v;
      c.w = (width * c.h + hmax_ - 1) / hmax_;
      c.hgt = (height * c.v + vmax_ - 1) / vmax_;
    }
    return kOk;
  }
  void Allocate() {
    for (int i = 0; i < ncomp; ++i) {
      Component& c = comp_[i];
      if (c.plane.empty()) c.plane.assign(size_t(c.bw) * 8 * c.bh * 8, 128);
      if (progressive_ && c.coef.empty()) c.coef.assign(size_t(c.bw) * c.bh * 64, 0);
    }
  }

  // ---------------------------------------------------------------- sequential block
  int DecodeBlockSeq(BitReader& br, Component& c, float* deq, bool* dc_only) {
    const HuffTable& hd = dc_[c.td];
    const HuffTable& ha = ac_[c.ta];
    const float* q = qf_[c.tq];
    memset(deq, 0, 64 * sizeof(float));
    int t = br.Decode(hd);
    if (t < 0 || t > 15) return kCorrupt;
    c.pred += br.Extend(t);
    deq[0] = float(c.pred) * q[0];
    *dc_only = true;
    for (int k = 1; k < 64;) {
      if (br.nbits < 16) br.Fill();
      const int fast = ha.fast_ac[br.Peek(9)];
      if (fast) {               // code and magnitude bits together in the 9-bit window
        k += (fast >> 4) & 15;
        if (k > 63) return kCorrupt;
        br.Skip(fast & 15);
        const int z = kZigzag[k++];
        deq[z] = float(fast >> 8) * q[z];
        *dc_only = false;
        continue;
      }
      int rs = br.Decode(ha);
      if (rs < 0) return kCorrupt;
      const int r = rs >> 4, s = rs & 15;
      if (s == 0) {
        if (r != 15) break;     // end of block
        k += 16;
        continue;
      }
      k += r;
      if (k > 63) return kCorrupt;
      const int z = kZigzag[k++];
      deq[z] = float(br.Extend(s)) * q[z];
      *dc_only = false;
    }
    return kOk;
  }

  // ---------------------------------------------------------------- progressive blocks (coefficients stay quantised)
  int DecodeBlockProgDC(BitReader& br, Component& c, int16_t* blk, int ah, int al) {
    if (ah == 0) {
      int t = br.Decode(dc_[c.td]);
      if (t < 0 || t > 15) return kCorrupt;
      c.pred += br.Extend(t);
      blk[0] = int16_t(c.pred * (1 << al));
    } else if (br.Bit()) {
      blk[0] = int16_t(blk[0] | (1 << al));
    }
    return kOk;
  }
  int DecodeBlockProgAC(BitReader& br, Component& c, int16_t* blk, int ss, int se, int ah, int al) {
    const HuffTable& ha = ac_[c.ta];
    if (ah == 0) {
      if (eobrun_ > 0) { --eobrun_; return kOk; }
      for (int k = ss; k <= se;) {
        int rs = br.Decode(ha);
        if (rs < 0) return kCorrupt;
        const int r = rs >> 4, s = rs & 15;
        if (s == 0) {
          if (r < 15) {
            eobrun_ = (1 << r) - 1;
            if (r) eobrun_ += br.Bits(r);
            break;
          }
          k += 16;
        } else {
          k += r;
          if (k > 63) return kCorrupt;
          blk[kZigzag[k++]] = int16_t(br.Extend(s) * (1 << al));
        }
      }
      return kOk;
    }
    // refinement scan
    const int p1 = 1 << al, m1 = -(1 << al);
    int k = ss;
    if (eobrun_ == 0) {
      while (k <= se) {
        int rs = br.Decode(ha);
        if (rs < 0) return kCorrupt;
        int r = rs >> 4;
        const int s = rs & 15;
        int newval = 0;
        if (s == 0) {
          if (r < 15) {
            eobrun_ = (1 << r);
            if (r) eobrun_ += br.Bits(r);
            break;
          }
        } else {
          if (s != 1) return kCorrupt;
          newval = br.Bit() ? p1 : m1;
        }
        while (k <= se) {
          int16_t& v = blk[kZigzag[k++]];
          if (v != 0) {
            if (br.Bit() && (v & p1) == 0) v = int16_t(v >= 0 ? v + p1 : v + m1);
          } else {
            if (r == 0) { if (newval) v = int16_t(newval); break; }
            --r;
          }
        }
      }
    }
    if (eobrun_ > 0) {
      while (k <= se) {
        int16_t& v = blk[kZigzag[k++]];
        if (v != 0 && br.Bit() && (v & p1) == 0) v = int16_t(v >= 0 ? v + p1 : v + m1);
      }
      --eobrun_;
    }
    return kOk;
  }

  int ParseScan() {
    const int len = SegLen();
    if (len < 6) return kTruncated;
    const uint8_t* p = pos_ + 2;
    const int ns = p[0];
    if (ns < 1 || ns > ncomp || len < 6 + 2 * ns) return kCorrupt;
    int order[3];
    for (int i = 0; i < ns; ++i) {
      int ci = -1;
      for (int j = 0; j < ncomp; ++j) if (comp_[j].id == p[1 + 2 * i]) ci = j;
      if (ci < 0) return kCorrupt;
      comp_[ci].td = p[2 + 2 * i] >> 4;
      comp_[ci].ta = p[2 + 2 * i] & 15;
      if (comp_[ci].td > 3 || comp_[ci].ta > 3) return kCorrupt;
      order[i] = ci;
    }
    const int ss = p[1 + 2 * ns], se = p[2 + 2 * ns], ah = p[3 + 2 * ns] >> 4, al = p[3 + 2 * ns] & 15;
    pos_ += len;
    if (progressive_) {
      if (ss > se || se > 63 || (ss == 0 && se != 0) || (ss > 0 && ns != 1) || al > 13) return kCorrupt;
    }
    for (int i = 0; i < ns; ++i) {
      const Component& c = comp_[order[i]];
      if (!qt_present_[c.tq] && !progressive_) return kCorrupt;
      if ((!progressive_ || ss == 0) && !(progressive_ && ah) && !dc_[c.td].present) return kCorrupt;
      if ((!progressive_ || ss > 0) && !ac_[c.ta].present) return kCorrupt;
    }
    Allocate();
    seen_scan_ = true;
    BitReader br;
    br.Reset(pos_, e_);
    for (int j = 0; j < ncomp; ++j) comp_[j].pred = 0;
    eobrun_ = 0;
    int todo = restart_interval_ ? restart_interval_ : 0x7fffffff;
    float deq[64];
    bool dc_only = false;
    int rc = kOk;
    auto restart_check = [&]() -> bool {      // returns false when the scan's data ended
      if (--todo > 0) return true;
      // expect RSTn: drop the remaining bits, consume the marker
      if (br.nbits < 32) br.Fill();
      if (br.marker >= 0xD0 && br.marker <= 0xD7) {
        const uint8_t* at = br.p;
        br.Reset(at, e_);
        for (int j = 0; j < ncomp; ++j) comp_[j].pred = 0;
        eobrun_ = 0;
        todo = restart_interval_;
        return true;
      }
      return br.marker == 0;     // no marker yet although the interval is over: damaged, carry on
    };
    if (ns == 1) {
      Component& c = comp_[order[0]];
      // non-interleaved: only the blocks that cover the component's real size
      const int nbx = (c.w + 7) / 8, nby = (c.hgt + 7) / 8;
      for (int by = 0; by < nby && rc == kOk; ++by)
        for (int bx = 0; bx < nbx; ++bx) {
          if (progressive_) {
            int16_t* blk = c.coef.data() + (size_t(by) * c.bw + bx) * 64;
            rc = ss == 0 ? DecodeBlockProgDC(br, c, blk, ah, al) : DecodeBlockProgAC(br, c, blk, ss, se, ah, al);
          } else {
            rc = DecodeBlockSeq(br, c, deq, &dc_only);
            if (rc == kOk) Idct8x8(deq, c.plane.data() + (size_t(by) * 8 * c.bw + bx) * 8, c.bw * 8, dc_only);
          }
          if (rc != kOk) break;
          if (!restart_check()) { by = nby; break; }
        }
    } else {
      for (int my = 0; my < mcuy_ && rc == kOk; ++my)
        for (int mx = 0; mx < mcux_; ++mx) {
          for (int i = 0; i < ns && rc == kOk; ++i) {
            Component& c = comp_[order[i]];
            for (int y = 0; y < c.v && rc == kOk; ++y)
              for (int x = 0; x < c.h; ++x) {
                const int bx = mx * c.h + x, by = my * c.v + y;
                if (progressive_) {
                  rc = DecodeBlockProgDC(br, c, c.coef.data() + (size_t(by) * c.bw + bx) * 64, ah, al);
                } else {
                  rc = DecodeBlockSeq(br, c, deq, &dc_only);
                  if (rc == kOk) Idct8x8(deq, c.plane.data() + (size_t(by) * 8 * c.bw + bx) * 8, c.bw * 8, dc_only);
                }
                if (rc != kOk) break;
              }
          }
          if (rc != kOk) break;
          if (!restart_check()) { my = mcuy_; break; }
        }
    }
    // continue marker parsing where the entropy data stopped
    if (br.marker) {
      pos_ = br.p - 2;
    } else {
      pos_ = br.p;
    }
    if (pos_ < b_ + 2) pos_ = b_ + 2;
    return rc;
  }

  void FinishProgressive() {
    float deq[64];
    for (int i = 0; i < ncomp; ++i) {
      Component& c = comp_[i];
      const float* q = qf_[c.tq];
      for (int by = 0; by < c.bh; ++by)
        for (int bx = 0; bx < c.bw; ++bx) {
          const int16_t* blk = c.coef.data() + (size_t(by) * c.bw + bx) * 64;
          bool dc_only = true;
          for (int k = 0; k < 64; ++k) {
            deq[k] = float(blk[k]) * q[k];
            if (k && blk[k]) dc_only = false;
          }
          Idct8x8(deq, c.plane.data() + (size_t(by) * 8 * c.bw + bx) * 8, c.bw * 8, dc_only);
        }
    }
  }

  // ---------------------------------------------------------------- up-sampling + colour
  // one full-resolution row `y` of component c into dst[width]
  mutable std::vector<uint8_t> tmp_;
  void UpsampledRow(const Component& c, int y, uint8_t* dst, std::vector<int>& colsum) const {
    const int stride = c.bw * 8;
    const int hs = hmax_ / c.h, vs = vmax_ / c.v;
    const uint8_t* base = c.plane.data();
    if (hs == 1 && vs == 1) { memcpy(dst, base + size_t(y) * stride, width); return; }
    const int cw = c.w;
    if (vs == 1) {                 // h2v1 triangle filter
      const uint8_t* in = base + size_t(y) * stride;
      if (cw == 1) { dst[0] = in[0]; if (width > 1) dst[1] = in[0]; return; }
      tmp_.resize(size_t(cw) * 2);
      uint8_t* o = tmp_.data();
      o[0] = in[0];
      o[1] = uint8_t((3 * in[0] + in[1] + 2) >> 2);
      for (int i = 1; i + 1 < cw; ++i) {
        o[2 * i] = uint8_t((3 * in[i] + in[i - 1] + 1) >> 2);
        o[2 * i + 1] = uint8_t((3 * in[i] + in[i + 1] + 2) >> 2);
      }
      o[2 * cw - 2] = uint8_t((3 * in[cw - 1] + in[cw - 2] + 1) >> 2);
      o[2 * cw - 1] = in[cw - 1];
      memcpy(dst, o, width);
      return;
    }
    const int sy = y >> 1;
    int ny = (y & 1) ? sy + 1 : sy - 1;        // the farther of the two source rows
    if (ny < 0) ny = 0;
    if (ny >= c.hgt) ny = c.hgt - 1;
    const uint8_t* in0 = base + size_t(sy < c.hgt ? sy : c.hgt - 1) * stride;
    const uint8_t* in1 = base + size_t(ny) * stride;
    if (hs == 1) {                 // h1v2: vertical triangle filter
      for (int x = 0; x < width; ++x) dst[x] = uint8_t((3 * in0[x] + in1[x] + ((y & 1) ? 2 : 1)) >> 2);
      return;
    }
    // h2v2: 3/4 near + 1/4 far vertically, then the same horizontally on the column sums
    colsum.resize(cw);
    for (int i = 0; i < cw; ++i) colsum[i] = 3 * in0[i] + in1[i];
    if (cw == 1) { dst[0] = uint8_t((colsum[0] * 4 + 8) >> 4); if (width > 1) dst[1] = uint8_t((colsum[0] * 4 + 7) >> 4); return; }
    tmp_.resize(size_t(cw) * 2);
    uint8_t* o = tmp_.data();
    const int* cs = colsum.data();
    o[0] = uint8_t((cs[0] * 4 + 8) >> 4);
    o[1] = uint8_t((3 * cs[0] + cs[1] + 7) >> 4);
    for (int i = 1; i + 1 < cw; ++i) {
      o[2 * i] = uint8_t((3 * cs[i] + cs[i - 1] + 8) >> 4);
      o[2 * i + 1] = uint8_t((3 * cs[i] + cs[i + 1] + 7) >> 4);
    }
    o[2 * cw - 2] = uint8_t((3 * cs[cw - 1] + cs[cw - 2] + 8) >> 4);
    o[2 * cw - 1] = uint8_t((cs[cw - 1] * 4 + 7) >> 4);
    memcpy(dst, o, width);
  }

  static inline uint8_t Clamp(int v) { return uint8_t(v < 0 ? 0 : (v > 255 ? 255 : v)); }

  void Output(uint8_t* rgb, float* planar) const {
    std::vector<uint8_t> rows(size_t(width) * 3), line(size_t(width) * 3);
    std::vector<int> colsum;
    const size_t plane = size_t(width) * height;
    const bool is_rgb = ncomp == 3 && (adobe_transform_ == 0 || (adobe_transform_ < 0 && comp_[0].id == 'R' &&
                                                                comp_[1].id == 'G' && comp_[2].id == 'B'));
    for (int y = 0; y < height; ++y) {
      uint8_t* yy = rows.data();
      uint8_t* cb = yy + width;
      uint8_t* cr = cb + width;
      UpsampledRow(comp_[0], y, yy, colsum);
      if (ncomp == 3) {
        UpsampledRow(comp_[1], y, cb, colsum);
        UpsampledRow(comp_[2], y, cr, colsum);
      }
      uint8_t* px = line.data();
      float* o0 = planar ? planar + size_t(y) * width : nullptr;
      if (ncomp == 3 && !is_rgb) {
        if (rgb) YccRowToRgb(yy, cb, cr, rgb + size_t(y) * width * 3, width);
        if (planar) YccRowToPlanar(yy, cb, cr, o0, o0 + plane, o0 + 2 * plane, width);
        continue;
      }
      if (ncomp == 1) {
        for (int x = 0; x < width; ++x) px[3 * x] = px[3 * x + 1] = px[3 * x + 2] = yy[x];
      } else {
        for (int x = 0; x < width; ++x) { px[3 * x] = yy[x]; px[3 * x + 1] = cb[x]; px[3 * x + 2] = cr[x]; }
      }
      if (rgb) memcpy(rgb + size_t(y) * width * 3, px, size_t(width) * 3);
      if (planar)
        for (int x = 0; x < width; ++x) { o0[x] = px[3 * x]; o0[plane + x] = px[3 * x + 1]; o0[2 * plane + x] = px[3 * x + 2]; }
    }
  }
};

static int DecodeOne(const uint8_t* buf, size_t n, uint8_t* rgb, float* planar, int w, int h) {
  Decoder d;
  int r = d.ParseHeader(buf, n);
  if (r != kOk) return r;
  if (d.width != w || d.height != h) return kSizeMismatch;
  return d.Decode(rgb, planar);
}

}  // namespace jpeg
}  // namespace cxnio

extern "C" {
// Differences (in bytes) between the AVX2 inverse DCT and the portable one over `nblocks` pseudo-random coefficient
// blocks (a mix of sparse, dense and out-of-range ones); -1 when this CPU has no AVX2.  For tests.
int cxnio_jpeg_idct_selftest(int nblocks, unsigned seed) {
#ifdef CXNIO_HAVE_AVX2_IDCT
  if (!cxnio::jpeg::g_has_avx2) return -1;
  uint32_t st = seed * 2654435761u + 12345u;
  auto rnd = [&]() { st = st * 1664525u + 1013904223u; return st >> 8; };
  int bad = 0;
  for (int b = 0; b < nblocks; ++b) {
    float in[64];
    const int kind = b % 4;
    for (int k = 0; k < 64; ++k) {
      const float u = float(int(rnd() % 2001) - 1000) / 1000.0f;        // [-1, 1]
      const bool keep = kind == 0 ? (k < 6) : kind == 1 ? (rnd() % 4 == 0) : true;
      const float amp = kind == 3 ? 4000.0f : 300.0f / float(1 + (k >> 3) + (k & 7));
      in[k] = keep ? u * amp : 0.0f;
    }
    uint8_t o1[64], o2[64];
    cxnio::jpeg::Idct8x8Avx2(in, o1, 8);
    cxnio::jpeg::Idct8x8Portable(in, o2, 8);
    for (int k = 0; k < 64; ++k) bad += o1[k] != o2[k];
  }
  return bad;
#else
  (void)nblocks; (void)seed;
  return -1;
#endif
}

int cxnio_jpeg_info(const uint8_t* buf, size_t n, int* width, int* height, int* components) {
  cxnio::jpeg::Decoder d;
  int r = d.ParseHeader(buf, n);
  if (r != 0) return r;
  *width = d.width; *height = d.height; *components = d.ncomp;
  return 0;
}
int cxnio_jpeg_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height) {
  return cxnio::jpeg::DecodeOne(buf, n, out_hwc, nullptr, width, height);
}
int cxnio_jpeg_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height) {
  return cxnio::jpeg::DecodeOne(buf, n, nullptr, out_chw, width, height);
}
int cxnio_jpeg_decode_many(int count, const uint8_t* const* bufs, const size_t* sizes, float* const* outs_chw,
                           const int* widths, const int* heights, int* status, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  if (nthreads > count) nthreads = count;
  std::atomic<int> next{0};
  auto work = [&]() {
    for (;;) {
      const int i = next.fetch_add(1);
      if (i >= count) return;
      status[i] = cxnio::jpeg::DecodeOne(bufs[i], sizes[i], nullptr, outs_chw[i], widths[i], heights[i]);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nthreads; ++t) pool.emplace_back(work);
  work();
  for (auto& t : pool) t.join();
  int bad = 0;
  for (int i = 0; i < count; ++i) bad += status[i] != 0;
  return bad;
}
}  // extern "C"
