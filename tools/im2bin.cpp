// im2bin: pack an image list into the imgbin BinaryPage format.
//   usage: im2bin <image.lst> <image_root_dir> <output.bin>
// Same command line and file format as the reference tool (tools/im2bin.cpp:6-67).
#include <cstdio>
#include "../csrc/runtime/page_reader.cpp"

int main(int argc, char* argv[]) {
  if (argc < 4) {
    printf("Usage: <image.lst> <image_root_dir> <output_file>\n");
    return 0;
  }
  const long long n = cxnio_pack(argv[1], argv[2], argv[3], 0);
  return n < 0 ? 1 : 0;
}
