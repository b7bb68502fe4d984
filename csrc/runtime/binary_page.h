// 64 MiB packed-blob page used by the imgbin format.
// Layout parity with the reference's utils::BinaryPage (src/utils/io.h:254-326): int32 data[kPageInts];
// data[0] = object count, data[r+2] = end offset of object r measured back from the end of the
// page (data[1] = 0), blobs packed from the end of the page towards the offset table.
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

namespace cxnio {

class BinaryPage {
 public:
  static constexpr size_t kPageInts = size_t(64) << 18;
  static constexpr size_t kPageBytes = kPageInts * sizeof(int32_t);
  BinaryPage() : data_(kPageInts, 0) {}
  void Clear() { std::fill(data_.begin(), data_.end(), 0); }
  int Size() const { return data_[0]; }
  bool Load(FILE* fi) { return fread(data_.data(), 1, kPageBytes, fi) == kPageBytes; }
  void Save(FILE* fo) const { fwrite(data_.data(), 1, kPageBytes, fo); }
  size_t FreeBytes() const { return (kPageInts - (Size() + 2)) * sizeof(int32_t) - data_[Size() + 1]; }
  bool Push(const void* p, size_t sz) {
    if (FreeBytes() < sz + sizeof(int32_t)) return false;
    const int n = Size();
    data_[n + 2] = data_[n + 1] + static_cast<int32_t>(sz);
    memcpy(Offset(data_[n + 2]), p, sz);
    data_[0] = n + 1;
    return true;
  }
  const void* Get(int r, size_t* sz) const {
    *sz = static_cast<size_t>(data_[r + 2] - data_[r + 1]);
    return const_cast<BinaryPage*>(this)->Offset(data_[r + 2]);
  }
  int32_t* raw() { return data_.data(); }

 private:
  void* Offset(int32_t pos) { return reinterpret_cast<char*>(data_.data()) + (kPageBytes - pos); }
  std::vector<int32_t> data_;
};

}  // namespace cxnio
