// In-memory PNG decoder for the image iterators (the reference decodes everything through OpenCV's imdecode,
// src/utils/decoder.h:80-124; this covers the second common format natively, zlib does the inflation).
// Non-interlaced PNG, bit depths 1-16, grey / RGB / palette, with or without alpha (alpha is dropped, as
// imdecode's colour mode does; 16-bit samples keep their high byte).  Interlaced files and anything malformed
// return an error code and the caller falls back to another decoder.
//
//   int cxnio_png_info(const uint8_t* buf, size_t n, int* width, int* height, int* components)
//   int cxnio_png_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height)
// return 0 on success; -1 truncated, -2 not a PNG, -3 unsupported variant, -4 corrupt, -5 size mismatch.
#include <zlib.h>

#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace cxnio {
namespace png {

static const uint8_t kSig[8] = {0x89, 'P', 'N', 'G', 0x0D, 0x0A, 0x1A, 0x0A};

static inline uint32_t Be32(const uint8_t* p) { return (uint32_t(p[0]) << 24) | (uint32_t(p[1]) << 16) | (uint32_t(p[2]) << 8) | p[3]; }

struct Header {
  int w = 0, h = 0, depth = 0, color = 0, interlace = 0;
  int channels() const { return color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4; }
};

static int ParseHeader(const uint8_t* buf, size_t n, Header* hd) {
  if (n < 8 || memcmp(buf, kSig, 8) != 0) return -2;
  if (n < 8 + 25) return -1;
  if (Be32(buf + 8) != 13 || memcmp(buf + 12, "IHDR", 4) != 0) return -4;
  const uint8_t* p = buf + 16;
  const uint32_t w = Be32(p), h = Be32(p + 4);
  hd->depth = p[8]; hd->color = p[9]; hd->interlace = p[12];
  if (w == 0 || h == 0 || w > 65535 || h > 65535 || uint64_t(w) * h > (uint64_t(1) << 26)) return -3;
  if (p[10] != 0 || p[11] != 0) return -4;
  hd->w = int(w); hd->h = int(h);
  const int d = hd->depth, c = hd->color;
  const bool ok = (c == 0 && (d == 1 || d == 2 || d == 4 || d == 8 || d == 16)) || (c == 3 && (d == 1 || d == 2 || d == 4 || d == 8)) ||
                  ((c == 2 || c == 4 || c == 6) && (d == 8 || d == 16));
  if (!ok) return -4;
  if (hd->interlace != 0) return -3;
  return 0;
}

static inline int Paeth(int a, int b, int c) {
  const int p = a + b - c, pa = abs(p - a), pb = abs(p - b), pc = abs(p - c);
  return (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c);
}

static int Decode(const uint8_t* buf, size_t n, uint8_t* out, int width, int height) {
  Header hd;
  int r = ParseHeader(buf, n, &hd);
  if (r) return r;
  if (hd.w != width || hd.h != height) return -5;
  // walk the chunks: palette + concatenated IDAT payload
  std::vector<uint8_t> idat;
  uint8_t pal[256][3];
  int npal = 0;
  size_t pos = 8;
  bool end = false;
  while (!end && pos + 12 <= n) {
    const uint32_t len = Be32(buf + pos);
    if (len > n - pos - 12) return -1;
    const uint8_t* type = buf + pos + 4;
    const uint8_t* data = buf + pos + 8;
    if (memcmp(type, "PLTE", 4) == 0) {
      if (len % 3 || len > 768) return -4;
      npal = int(len / 3);
      memcpy(pal, data, len);
    } else if (memcmp(type, "IDAT", 4) == 0) {
      idat.insert(idat.end(), data, data + len);
    } else if (memcmp(type, "IEND", 4) == 0) {
      end = true;
    }
    pos += 12 + size_t(len);
  }
  if (idat.empty()) return -1;
  if (hd.color == 3 && npal == 0) return -4;
  const int ch = hd.channels();
  const size_t bpp_bits = size_t(ch) * hd.depth;
  const size_t stride = (size_t(hd.w) * bpp_bits + 7) / 8;       // bytes per scanline without the filter byte
  const size_t bpp = bpp_bits >= 8 ? bpp_bits / 8 : 1;           // filter distance in bytes
  std::vector<uint8_t> raw((stride + 1) * hd.h);
  uLongf got = uLongf(raw.size());
  const int zr = uncompress(raw.data(), &got, idat.data(), uLong(idat.size()));
  if (zr != Z_OK || got != raw.size()) return -4;
  std::vector<uint8_t> prev(stride, 0), cur(stride);
  for (int y = 0; y < hd.h; ++y) {
    const uint8_t* line = raw.data() + (stride + 1) * y;
    const int ft = line[0];
    if (ft > 4) return -4;
    for (size_t i = 0; i < stride; ++i) {
      const int a = i >= bpp ? cur[i - bpp] : 0, b = prev[i], c = i >= bpp ? prev[i - bpp] : 0;
      int v = line[1 + i];
      switch (ft) {
        case 1: v += a; break;
        case 2: v += b; break;
        case 3: v += (a + b) >> 1; break;
        case 4: v += Paeth(a, b, c); break;
        default: break;
      }
      cur[i] = uint8_t(v);
    }
    uint8_t* o = out + size_t(y) * hd.w * 3;
    const int step = hd.depth == 16 ? 2 : 1;                     // 16 bit: big endian, keep the high byte
    for (int x = 0; x < hd.w; ++x) {
      uint8_t rr, gg, bb;
      if (hd.depth < 8) {                                        // packed grey or palette indices
        const int per = 8 / hd.depth, idx = x / per, sh = 8 - hd.depth * (x % per + 1);
        const int s = (cur[idx] >> sh) & ((1 << hd.depth) - 1);
        if (hd.color == 3) {
          if (s >= npal) return -4;
          rr = pal[s][0]; gg = pal[s][1]; bb = pal[s][2];
        } else {
          rr = gg = bb = uint8_t(s * 255 / ((1 << hd.depth) - 1));
        }
      } else {
        const uint8_t* px = cur.data() + size_t(x) * ch * step;
        if (hd.color == 3) {
          if (px[0] >= npal) return -4;
          rr = pal[px[0]][0]; gg = pal[px[0]][1]; bb = pal[px[0]][2];
        } else if (hd.color == 0 || hd.color == 4) {
          rr = gg = bb = px[0];
        } else {
          rr = px[0]; gg = px[step]; bb = px[2 * step];
        }
      }
      o[3 * x] = rr; o[3 * x + 1] = gg; o[3 * x + 2] = bb;
    }
    prev.swap(cur);
  }
  return 0;
}

}  // namespace png
}  // namespace cxnio

extern "C" {
int cxnio_png_info(const uint8_t* buf, size_t n, int* width, int* height, int* components) {
  cxnio::png::Header hd;
  const int r = cxnio::png::ParseHeader(buf, n, &hd);
  if (r) return r;
  *width = hd.w; *height = hd.h; *components = 3;
  return 0;
}
int cxnio_png_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height) {
  return cxnio::png::Decode(buf, n, out_hwc, width, height);
}
}  // extern "C"
