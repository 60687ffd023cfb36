// oracle/gtb_io.h -- TEST INFRASTRUCTURE (not product code).
// Minimal named-array container ("GTB1") shared by oracle/ref_tool.cpp and the
// Python side (gtsam_petercdev_b200/gtbio.py).  Layout, little endian:
//   magic "GTB1", u32 n_records, then per record:
//     u32 name_len, name bytes, u8 dtype (0=f64 1=i64 2=i32 3=u64), u32 ndim,
//     u64 shape[ndim], raw data (C order).
#pragma once
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace gtbio {

struct Array {
  uint8_t dtype = 0;
  std::vector<uint64_t> shape;
  std::vector<char> bytes;
  size_t count() const {
    size_t n = 1;
    for (auto s : shape) n *= s;
    return n;
  }
  size_t rows() const { return shape.empty() ? 0 : shape[0]; }
  size_t cols() const { return shape.size() < 2 ? 1 : shape[1]; }
  const double* f64() const { check(0); return reinterpret_cast<const double*>(bytes.data()); }
  const int64_t* i64() const { check(1); return reinterpret_cast<const int64_t*>(bytes.data()); }
  const int32_t* i32() const { check(2); return reinterpret_cast<const int32_t*>(bytes.data()); }
  const uint64_t* u64() const { check(3); return reinterpret_cast<const uint64_t*>(bytes.data()); }
  void check(uint8_t want) const {
    if (dtype != want) throw std::runtime_error("gtbio: dtype mismatch");
  }
};

inline size_t dtype_size(uint8_t d) { return d == 2 ? 4 : 8; }

struct File {
  std::map<std::string, Array> arrays;
  std::vector<std::string> order;

  bool has(const std::string& n) const { return arrays.count(n) != 0; }
  const Array& at(const std::string& n) const {
    auto it = arrays.find(n);
    if (it == arrays.end()) throw std::runtime_error("gtbio: missing array " + n);
    return it->second;
  }

  template <class T>
  void put(const std::string& name, uint8_t dtype, const std::vector<uint64_t>& shape, const T* data) {
    Array a;
    a.dtype = dtype;
    a.shape = shape;
    size_t nb = a.count() * dtype_size(dtype);
    a.bytes.resize(nb);
    if (nb) std::memcpy(a.bytes.data(), data, nb);
    if (!arrays.count(name)) order.push_back(name);
    arrays[name] = std::move(a);
  }
  void put_f64(const std::string& n, const std::vector<double>& v, std::vector<uint64_t> shape = {}) {
    if (shape.empty()) shape = {v.size()};
    put(n, 0, shape, v.data());
  }
  void put_i64(const std::string& n, const std::vector<int64_t>& v, std::vector<uint64_t> shape = {}) {
    if (shape.empty()) shape = {v.size()};
    put(n, 1, shape, v.data());
  }
  void put_u64(const std::string& n, const std::vector<uint64_t>& v, std::vector<uint64_t> shape = {}) {
    if (shape.empty()) shape = {v.size()};
    put(n, 3, shape, v.data());
  }

  void load(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("gtbio: cannot open " + path);
    char magic[4];
    uint32_t n = 0;
    rd(f, magic, 4);
    if (std::memcmp(magic, "GTB1", 4)) throw std::runtime_error("gtbio: bad magic");
    rd(f, &n, 4);
    for (uint32_t r = 0; r < n; ++r) {
      uint32_t len;
      rd(f, &len, 4);
      std::string name(len, 0);
      rd(f, name.data(), len);
      Array a;
      rd(f, &a.dtype, 1);
      uint32_t nd;
      rd(f, &nd, 4);
      a.shape.resize(nd);
      rd(f, a.shape.data(), 8 * nd);
      a.bytes.resize(a.count() * dtype_size(a.dtype));
      rd(f, a.bytes.data(), a.bytes.size());
      order.push_back(name);
      arrays[name] = std::move(a);
    }
    std::fclose(f);
  }

  void save(const std::string& path) const {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("gtbio: cannot write " + path);
    uint32_t n = order.size();
    std::fwrite("GTB1", 1, 4, f);
    std::fwrite(&n, 4, 1, f);
    for (auto& name : order) {
      const Array& a = arrays.at(name);
      uint32_t len = name.size(), nd = a.shape.size();
      std::fwrite(&len, 4, 1, f);
      std::fwrite(name.data(), 1, len, f);
      std::fwrite(&a.dtype, 1, 1, f);
      std::fwrite(&nd, 4, 1, f);
      std::fwrite(a.shape.data(), 8, nd, f);
      if (!a.bytes.empty()) std::fwrite(a.bytes.data(), 1, a.bytes.size(), f);
    }
    std::fclose(f);
  }

 private:
  static void rd(FILE* f, void* p, size_t n) {
    if (n && std::fread(p, 1, n, f) != n) throw std::runtime_error("gtbio: short read");
  }
};

}  // namespace gtbio
