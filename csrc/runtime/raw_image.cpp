// Pre-decoded images inside imgbin pages, and the format dispatch of the native pipeline.
//
// A JPEG costs ~1.4 ms of one core to decode; a B200 trains AlexNet on ~100 images per millisecond.  For data
// sets that are read many times it pays to decode once, at packing time: `im2bin list root out.bin raw=1` stores
// every image as a raw record
//     bytes 0-3  "CXNR"      4-5  width (LE uint16)    6-7  height    8  channels (3)    9-11  zero
//     12...      height x width x channels bytes, interleaved RGB
// which the batch stage only has to crop, normalise and convert (memory speed, no entropy decoding).  No
// counterpart in the reference, whose pages always hold encoded files (tools/im2bin.cpp:6-67).
//
//   cxnio_image_info / cxnio_image_decode_planar_f32   JPEG, PNG or raw record, chosen by the first bytes
//   cxnio_raw_encode                                   planar-or-interleaved pixels -> raw record
//   cxnio_pack_raw                                     im2bin with decoding (JPEG and PNG sources)
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "binary_page.h"

extern "C" int cxnio_jpeg_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
extern "C" int cxnio_jpeg_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height);
extern "C" int cxnio_jpeg_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height);
extern "C" int cxnio_png_info(const uint8_t* buf, size_t n, int* width, int* height, int* components);
extern "C" int cxnio_png_decode_rgb(const uint8_t* buf, size_t n, uint8_t* out_hwc, int width, int height);

namespace cxnio {
constexpr size_t kRawHeader = 12;
static bool IsRaw(const uint8_t* b, size_t n) { return n >= kRawHeader && memcmp(b, "CXNR", 4) == 0; }
static bool IsPng(const uint8_t* b, size_t n) { return n >= 8 && memcmp(b, "\x89PNG", 4) == 0; }
static void HwcToPlanar(const uint8_t* px, size_t plane, float* out_chw) {
  for (size_t i = 0; i < plane; ++i) {
    out_chw[i] = float(px[3 * i]);
    out_chw[plane + i] = float(px[3 * i + 1]);
    out_chw[2 * plane + i] = float(px[3 * i + 2]);
  }
}
}  // namespace cxnio

extern "C" {

// 0 = ok; negative as the JPEG functions (-2: neither format, -4: inconsistent raw record)
int cxnio_image_info(const uint8_t* buf, size_t n, int* width, int* height, int* components) {
  if (cxnio::IsRaw(buf, n)) {
    const int w = buf[4] | (buf[5] << 8), h = buf[6] | (buf[7] << 8), c = buf[8];
    if (w <= 0 || h <= 0 || c != 3 || n != cxnio::kRawHeader + size_t(w) * h * c) return -4;
    *width = w; *height = h; *components = c;
    return 0;
  }
  if (cxnio::IsPng(buf, n)) return cxnio_png_info(buf, n, width, height, components);
  return cxnio_jpeg_info(buf, n, width, height, components);
}

int cxnio_image_decode_planar_f32(const uint8_t* buf, size_t n, float* out_chw, int width, int height) {
  if (cxnio::IsPng(buf, n)) {
    std::vector<uint8_t> rgb(size_t(width) * height * 3);
    const int rp = cxnio_png_decode_rgb(buf, n, rgb.data(), width, height);
    if (rp == 0) cxnio::HwcToPlanar(rgb.data(), size_t(width) * height, out_chw);
    return rp;
  }
  if (!cxnio::IsRaw(buf, n)) return cxnio_jpeg_decode_planar_f32(buf, n, out_chw, width, height);
  int w, h, c;
  const int r = cxnio_image_info(buf, n, &w, &h, &c);
  if (r) return r;
  if (w != width || h != height) return -5;
  cxnio::HwcToPlanar(buf + cxnio::kRawHeader, size_t(w) * h, out_chw);
  return 0;
}

// interleaved RGB bytes -> raw record; out must hold 12 + width*height*3 bytes; returns the record size or 0
size_t cxnio_raw_encode(const uint8_t* rgb_hwc, int width, int height, uint8_t* out) {
  if (width <= 0 || height <= 0 || width > 65535 || height > 65535) return 0;
  memcpy(out, "CXNR", 4);
  out[4] = uint8_t(width & 255); out[5] = uint8_t(width >> 8);
  out[6] = uint8_t(height & 255); out[7] = uint8_t(height >> 8);
  out[8] = 3; out[9] = out[10] = out[11] = 0;
  memcpy(out + cxnio::kRawHeader, rgb_hwc, size_t(width) * height * 3);
  return cxnio::kRawHeader + size_t(width) * height * 3;
}

// like cxnio_pack, but every (JPEG) file is decoded and stored as a raw record; returns the image count or -1
long long cxnio_pack_raw(const char* list_path, const char* root, const char* out_path, int silent) {
  FILE* fl = fopen(list_path, "r");
  FILE* fo = fopen(out_path, "wb");
  if (!fl || !fo) { if (fl) fclose(fl); if (fo) fclose(fo); return -1; }
  cxnio::BinaryPage pg;
  std::vector<uint8_t> file, rgb, rec;
  char line[4096];
  long long n = 0;
  auto fail = [&](const char* what, const std::string& path) {
    fprintf(stderr, "im2bin raw: %s: %s\n", what, path.c_str());
    fclose(fl); fclose(fo);
    return -1LL;
  };
  while (fgets(line, sizeof(line), fl)) {
    std::string s(line);
    while (!s.empty() && (s.back() == '\n' || s.back() == '\r' || s.back() == ' ')) s.pop_back();
    const size_t p = s.find_last_of("\t ");
    if (p == std::string::npos) continue;
    const std::string path = std::string(root) + s.substr(p + 1);
    FILE* fi = fopen(path.c_str(), "rb");
    if (!fi) return fail("cannot open", path);
    fseek(fi, 0, SEEK_END);
    const long sz = ftell(fi);
    fseek(fi, 0, SEEK_SET);
    file.resize(sz);
    const bool ok = fread(file.data(), 1, sz, fi) == size_t(sz);
    fclose(fi);
    if (!ok) return fail("cannot read", path);
    int w, h, c;
    const bool is_png = cxnio::IsPng(file.data(), file.size());
    if ((is_png ? cxnio_png_info(file.data(), file.size(), &w, &h, &c) : cxnio_jpeg_info(file.data(), file.size(), &w, &h, &c)) != 0)
      return fail("neither a JPEG nor a PNG this decoder reads", path);
    rgb.resize(size_t(w) * h * 3);
    if ((is_png ? cxnio_png_decode_rgb(file.data(), file.size(), rgb.data(), w, h)
                : cxnio_jpeg_decode_rgb(file.data(), file.size(), rgb.data(), w, h)) != 0)
      return fail("damaged image", path);
    rec.resize(cxnio::kRawHeader + rgb.size());
    const size_t rsz = cxnio_raw_encode(rgb.data(), w, h, rec.data());
    if (!rsz) return fail("image too large for a raw record", path);
    if (!pg.Push(rec.data(), rsz)) {
      pg.Save(fo);
      pg.Clear();
      if (!pg.Push(rec.data(), rsz)) return fail("image larger than a page", path);
    }
    ++n;
    if (!silent && n % 1000 == 0) { printf("\r%lld images decoded and packed", n); fflush(stdout); }
  }
  if (pg.Size() != 0) pg.Save(fo);
  fclose(fl);
  fclose(fo);
  if (!silent) printf("\rall %lld images decoded and packed into %s\n", n, out_path);
  return n;
}

}  // extern "C"
