// im2bin: pack an image list into the imgbin BinaryPage format.
//   usage: im2bin <image.lst> <image_root_dir> <output.bin> [raw=1]
// Same command line and file format as the reference tool (tools/im2bin.cpp:6-67).  With raw=1 every file (JPEG or PNG) is
// decoded now and stored as a raw RGB record (csrc/runtime/raw_image.cpp): about 8x the bytes, no decoding at
// training time.
#include <cstdio>
#include <cstring>
#include "../csrc/runtime/page_reader.cpp"
#include "../csrc/runtime/jpeg_decoder.cpp"
#include "../csrc/runtime/png_decoder.cpp"
#include "../csrc/runtime/raw_image.cpp"

int main(int argc, char* argv[]) {
  if (argc < 4) {
    printf("Usage: <image.lst> <image_root_dir> <output_file> [raw=1]\n");
    return 0;
  }
  const bool raw = argc > 4 && strcmp(argv[4], "raw=1") == 0;
  const long long n = raw ? cxnio_pack_raw(argv[1], argv[2], argv[3], 0) : cxnio_pack(argv[1], argv[2], argv[3], 0);
  return n < 0 ? 1 : 0;
}
