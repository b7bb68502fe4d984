// Native page reader for imgbin / imgbinx: a producer thread streams 64 MiB BinaryPages from one
// or more .bin files into a ring of buffers while the consumer walks the blobs of the current
// page - the role of the reference's utils::ThreadBuffer + ThreadImagePageIterator::Factory
// (src/utils/thread_buffer.h:22-202, src/io/iter_thread_imbin-inl.hpp:222-275), written with
// std::thread / condition variables instead of the reference's semaphore ping-pong.
//
// C ABI (used by cxxnet_b200/io/native.py through ctypes):
//   void* cxnio_reader_open(const char** paths, int npaths, int nbuffers)
//   int   cxnio_reader_next(void* h, const void** blob, unsigned long long* size)  1 = blob, 0 = end of data
//   int   cxnio_reader_next_page(void* h, void* dst)   next whole 64 MiB page into caller memory, 0 = end of data
//   void  cxnio_reader_rewind(void* h)          restart from the first file
//   void  cxnio_reader_close(void* h)
//   int   cxnio_pack(...)                        im2bin: see tools/im2bin.cpp
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "binary_page.h"

namespace cxnio {

class PageReader {
 public:
  PageReader(std::vector<std::string> paths, int nbuf) : paths_(std::move(paths)), nbuf_(nbuf < 2 ? 2 : nbuf) {
    for (int i = 0; i < nbuf_; ++i) free_.push_back(std::make_unique<BinaryPage>());
    Start();
  }
  ~PageReader() { Stop(); }

  bool Next(const void** blob, unsigned long long* size) {
    for (;;) {
      if (cur_ && top_ < cur_->Size()) {
        size_t sz;
        *blob = cur_->Get(top_++, &sz);
        *size = sz;
        return true;
      }
      std::unique_lock<std::mutex> lk(mu_);
      if (cur_) {
        free_.push_back(std::move(cur_));
        cv_free_.notify_one();
      }
      cv_full_.wait(lk, [&] { return !full_.empty() || done_; });
      if (full_.empty()) return false;
      cur_ = std::move(full_.front());
      full_.pop_front();
      top_ = 0;
    }
  }
  // the next whole page copied into caller-owned memory (kPageBytes): records can then be handed out as views of
  // that buffer with no per-record copy and no lifetime tie to this reader.  Not to be mixed with Next() in a pass.
  bool NextPage(void* dst) {
    std::unique_lock<std::mutex> lk(mu_);
    if (cur_) {
      free_.push_back(std::move(cur_));
      cv_free_.notify_one();
    }
    cv_full_.wait(lk, [&] { return !full_.empty() || done_; });
    if (full_.empty()) return false;
    std::unique_ptr<BinaryPage> pg = std::move(full_.front());
    full_.pop_front();
    lk.unlock();
    memcpy(dst, pg->raw(), BinaryPage::kPageBytes);
    lk.lock();
    free_.push_back(std::move(pg));
    cv_free_.notify_one();
    return true;
  }
  void Rewind() {
    Stop();
    {
      std::lock_guard<std::mutex> lk(mu_);
      if (cur_) free_.push_back(std::move(cur_));
      while (!full_.empty()) { free_.push_back(std::move(full_.front())); full_.pop_front(); }
      done_ = false;
      stop_ = false;
      top_ = 0;
    }
    Start();
  }

 private:
  void Start() { thr_ = std::thread([this] { Run(); }); }
  void Stop() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_free_.notify_all();
    if (thr_.joinable()) thr_.join();
  }
  void Run() {
    for (const std::string& p : paths_) {
      FILE* fi = fopen(p.c_str(), "rb");
      if (!fi) { fprintf(stderr, "cxnio: cannot open %s\n", p.c_str()); break; }
      for (;;) {
        std::unique_ptr<BinaryPage> pg;
        {
          std::unique_lock<std::mutex> lk(mu_);
          cv_free_.wait(lk, [&] { return !free_.empty() || stop_; });
          if (stop_) { fclose(fi); return; }
          pg = std::move(free_.back());
          free_.pop_back();
        }
        const bool ok = pg->Load(fi);
        std::lock_guard<std::mutex> lk(mu_);
        if (!ok) { free_.push_back(std::move(pg)); break; }
        full_.push_back(std::move(pg));
        cv_full_.notify_one();
      }
      fclose(fi);
    }
    std::lock_guard<std::mutex> lk(mu_);
    done_ = true;
    cv_full_.notify_all();
  }

  std::vector<std::string> paths_;
  int nbuf_;
  std::mutex mu_;
  std::condition_variable cv_free_, cv_full_;
  std::vector<std::unique_ptr<BinaryPage>> free_;
  std::deque<std::unique_ptr<BinaryPage>> full_;
  std::unique_ptr<BinaryPage> cur_;
  int top_ = 0;
  bool done_ = false, stop_ = false;
  std::thread thr_;
};

}  // namespace cxnio

extern "C" {
void* cxnio_reader_open(const char** paths, int npaths, int nbuffers) {
  std::vector<std::string> v(paths, paths + npaths);
  return new cxnio::PageReader(std::move(v), nbuffers);
}
int cxnio_reader_next(void* h, const void** blob, unsigned long long* size) {
  return static_cast<cxnio::PageReader*>(h)->Next(blob, size) ? 1 : 0;
}
int cxnio_reader_next_page(void* h, void* dst) { return static_cast<cxnio::PageReader*>(h)->NextPage(dst) ? 1 : 0; }
void cxnio_reader_rewind(void* h) { static_cast<cxnio::PageReader*>(h)->Rewind(); }
void cxnio_reader_close(void* h) { delete static_cast<cxnio::PageReader*>(h); }

// pack the files named in `list_path` (lines: index <tab> label.. <tab> relative path) found under
// `root` into BinaryPages appended to `out_path`; returns the number of images or -1.
long long cxnio_pack(const char* list_path, const char* root, const char* out_path, int silent) {
  FILE* fl = fopen(list_path, "r");
  FILE* fo = fopen(out_path, "wb");
  if (!fl || !fo) return -1;
  cxnio::BinaryPage pg;
  std::vector<char> buf;
  char line[4096];
  long long n = 0;
  while (fgets(line, sizeof(line), fl)) {
    std::string s(line);
    while (!s.empty() && (s.back() == '\n' || s.back() == '\r' || s.back() == ' ')) s.pop_back();
    const size_t p = s.find_last_of("\t ");
    if (p == std::string::npos) continue;
    const std::string path = std::string(root) + s.substr(p + 1);
    FILE* fi = fopen(path.c_str(), "rb");
    if (!fi) { fprintf(stderr, "im2bin: cannot open %s\n", path.c_str()); fclose(fl); fclose(fo); return -1; }
    fseek(fi, 0, SEEK_END);
    const long sz = ftell(fi);
    fseek(fi, 0, SEEK_SET);
    buf.resize(sz);
    if (fread(buf.data(), 1, sz, fi) != static_cast<size_t>(sz)) { fclose(fi); fclose(fl); fclose(fo); return -1; }
    fclose(fi);
    if (!pg.Push(buf.data(), sz)) {
      pg.Save(fo);
      pg.Clear();
      if (!pg.Push(buf.data(), sz)) { fprintf(stderr, "im2bin: image larger than a page: %s\n", path.c_str()); return -1; }
    }
    ++n;
    if (!silent && n % 1000 == 0) { printf("\r%lld images packed", n); fflush(stdout); }
  }
  if (pg.Size() != 0) pg.Save(fo);
  fclose(fl);
  fclose(fo);
  if (!silent) printf("\rall %lld images packed into %s\n", n, out_path);
  return n;
}
}  // extern "C"
