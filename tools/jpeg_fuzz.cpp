// Mutation fuzzer for the native image decoders (csrc/runtime/jpeg_decoder.cpp, png_decoder.cpp).
//   g++ -O1 -g -std=c++17 -fsanitize=address,undefined -fno-sanitize-recover=undefined -pthread \
//       -o jpeg_fuzz tools/jpeg_fuzz.cpp csrc/runtime/jpeg_decoder.cpp csrc/runtime/png_decoder.cpp -lz
//   ./jpeg_fuzz <iterations per file> a.jpg b.png ...
// Every iteration flips / overwrites a few bytes of a seed file (and sometimes truncates it), copies the result
// into an exact-size heap block so that the sanitizer sees any over-read, and decodes it.  The decoder may
// reject the input; it must not crash, read or write out of bounds, or overflow.  Prints the two counts.
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

extern "C" int cxnio_jpeg_info(const uint8_t*, size_t, int*, int*, int*);
extern "C" int cxnio_jpeg_decode_rgb(const uint8_t*, size_t, uint8_t*, int, int);
extern "C" int cxnio_png_info(const uint8_t*, size_t, int*, int*, int*);
extern "C" int cxnio_png_decode_rgb(const uint8_t*, size_t, uint8_t*, int, int);

int main(int argc, char** argv) {
  if (argc < 3) {
    printf("usage: jpeg_fuzz <iterations> file.jpg ...\n");
    return 2;
  }
  const int iters = atoi(argv[1]);
  long decoded = 0, rejected = 0;
  for (int f = 2; f < argc; ++f) {
    FILE* fi = fopen(argv[f], "rb");
    if (!fi) { fprintf(stderr, "cannot open %s\n", argv[f]); return 2; }
    std::vector<uint8_t> seed;
    for (int c; (c = fgetc(fi)) != EOF;) seed.push_back(uint8_t(c));
    fclose(fi);
    srand(f);
    for (int it = 0; it < iters; ++it) {
      std::vector<uint8_t> m = seed;
      const int edits = 1 + rand() % 6;
      for (int k = 0; k < edits; ++k) {
        const size_t p = rand() % m.size();
        switch (rand() % 3) {
          case 0: m[p] = uint8_t(rand()); break;
          case 1: m[p] ^= uint8_t(1 << (rand() % 8)); break;
          default: m[p] = 0xFF;
        }
      }
      if (rand() % 10 == 0) m.resize(rand() % m.size() + 1);
      std::vector<uint8_t> exact(m);
      int w, h, n;
      const bool png = seed.size() > 4 && seed[1] == 'P' && seed[2] == 'N' && seed[3] == 'G';
      if ((png ? cxnio_png_info(exact.data(), exact.size(), &w, &h, &n) : cxnio_jpeg_info(exact.data(), exact.size(), &w, &h, &n)) != 0 ||
          long(w) * h > 4000000) { ++rejected; continue; }
      std::vector<uint8_t> out(size_t(w) * h * 3);
      if ((png ? cxnio_png_decode_rgb(exact.data(), exact.size(), out.data(), w, h)
               : cxnio_jpeg_decode_rgb(exact.data(), exact.size(), out.data(), w, h)) == 0) ++decoded; else ++rejected;
    }
  }
  printf("decoded %ld rejected %ld\n", decoded, rejected);
  return 0;
}
