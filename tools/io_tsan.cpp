// Thread-sanitizer exercise of the native data pipeline's concurrency (page reader thread, worker-thread batch calls).
//   g++ -O1 -g -std=c++17 -fsanitize=thread -DCXNIO_NO_CLONES -pthread -o io_tsan tools/io_tsan.cpp csrc/runtime/page_reader.cpp \
//       csrc/runtime/jpeg_decoder.cpp csrc/runtime/png_decoder.cpp csrc/runtime/raw_image.cpp \
//       csrc/runtime/batch_pipeline.cpp -lz
//   ./io_tsan /tmp/scratch.bin
// Writes three pages of raw records, then: reads them record by record with a rewind in the middle of a page, reads
// them page by page, and runs batch calls from two caller threads at once (each with its own worker threads) while a
// third reader streams pages.  Any data race makes the sanitizer fail the run; prints "ok" otherwise.
// (-DCXNIO_NO_CLONES: the AVX2 / baseline function clones are dispatched by ifunc resolvers, which run before the
// sanitizer's runtime is up.)
#include <atomic>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include "../csrc/runtime/binary_page.h"
#include "../include/cxxnet_io.h"

static std::vector<uint8_t> Record(int w, int h, int seed) {
  std::vector<uint8_t> px(size_t(w) * h * 3), rec(12 + px.size());
  for (size_t i = 0; i < px.size(); ++i) px[i] = uint8_t((i * 7 + seed * 13) & 255);
  cxnio_raw_encode(px.data(), w, h, rec.data());
  return rec;
}

int main(int argc, char** argv) {
  if (argc < 2) { printf("usage: io_tsan scratch.bin\n"); return 2; }
  const int kPerPage = 40, kPages = 3, W = 64, H = 48, C = 3, OH = 40, OW = 56;
  {
    FILE* fo = fopen(argv[1], "wb");
    if (!fo) return 2;
    cxnio::BinaryPage pg;
    for (int p = 0; p < kPages; ++p) {
      pg.Clear();
      for (int i = 0; i < kPerPage; ++i) {
        std::vector<uint8_t> r = Record(W, H, p * kPerPage + i);
        if (!pg.Push(r.data(), r.size())) return 3;
      }
      pg.Save(fo);
    }
    fclose(fo);
  }
  const char* paths[1] = {argv[1]};
  // 1. record API with a rewind in the middle of the second page
  void* rd = cxnio_reader_open(paths, 1, 2);
  const void* blob; unsigned long long size; long n = 0;
  for (int i = 0; i < kPerPage + 7; ++i) if (!cxnio_reader_next(rd, &blob, &size)) return 4;
  cxnio_reader_rewind(rd);
  while (cxnio_reader_next(rd, &blob, &size)) ++n;
  if (n != kPerPage * kPages) { printf("record count %ld\n", n); return 5; }
  cxnio_reader_close(rd);
  // 2. page API feeding two concurrent batch callers, a third thread streaming pages at the same time
  std::vector<uint8_t> page(cxnio::BinaryPage::kPageBytes);
  rd = cxnio_reader_open(paths, 1, 3);
  if (!cxnio_reader_next_page(rd, page.data())) return 6;
  const int32_t* head = reinterpret_cast<const int32_t*>(page.data());
  std::vector<CxnioAugItem> items(kPerPage);
  for (int r = 0; r < kPerPage; ++r) {
    memset(&items[r], 0, sizeof(CxnioAugItem));
    items[r].buf = page.data() + (cxnio::BinaryPage::kPageBytes - head[r + 2]);
    items[r].size = uint64_t(head[r + 2] - head[r + 1]);
    items[r].src_h = H; items[r].src_w = W; items[r].y0 = r % 8; items[r].x0 = r % 8; items[r].flip = r & 1;
    items[r].contrast = 1.0f;
  }
  const float mean[3] = {100.f, 110.f, 120.f};
  std::atomic<int> bad{0};
  auto caller = [&](int kind) {
    std::vector<uint8_t> dst(size_t(kPerPage) * C * OH * OW * 4);
    std::vector<int> status(kPerPage);
    for (int it = 0; it < 6; ++it)
      bad += cxnio_decode_augment_batch(items.data(), kPerPage, C, OH, OW, kind == 2 ? 0 : 1, kind == 2 ? nullptr : mean, 0, 0,
                                        kind == 2 ? 1.0f : 0.5f, dst.data(), kind, status.data(), 4);
  };
  std::thread streamer([&] {
    std::vector<uint8_t> other(cxnio::BinaryPage::kPageBytes);
    while (cxnio_reader_next_page(rd, other.data())) {}
  });
  std::thread a(caller, 0), b(caller, 2);
  a.join(); b.join(); streamer.join();
  cxnio_reader_close(rd);
  remove(argv[1]);
  if (bad.load()) { printf("%d instances failed\n", bad.load()); return 7; }
  printf("ok\n");
  return 0;
}
