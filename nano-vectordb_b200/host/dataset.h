// dataset.h — vecbin64 / raw12 files, memory-mapped; the host-side source of bytes for the device index.
//
// Mirrors the reference's loader contract so that callers written against it keep working:
//   formats      include/nvdb/vecbin_format.h:7-29 (64-byte packed header "NVDBVEC1", version 1, dtype 1/2/3, dim,
//                count; payload row-major; INT8 followed by `count` float scales) and the legacy raw12 header
//                {u32 count, u32 reserved, u32 dim} + f32 payload (src/vector_dataset.cpp:11-15)
//   mapping      include/nvdb/mmap_file.h:21-25, src/mmap_file.cpp:39-69 (read-only private mapping, RAII, move-only)
//   dataset      include/nvdb/vector_dataset.h:14-37, src/vector_dataset.cpp:24-157 (strict file-size check, typed
//                accessors that throw std::runtime_error with the same messages)
// Written from scratch; additions over the reference: madvise(SEQUENTIAL) for the upload pass, data_ptr() also
// answers for INT8, scales_ptr().
#pragma once
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cerrno>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <utility>

namespace nvdb {

constexpr uint64_t kMagic = 0x4E56444256454331ULL;  // "NVDBVEC1"
constexpr uint32_t kVersion = 1;
enum class DType : uint32_t { Float32 = 1, Float16 = 2, Int8 = 3 };

inline size_t bytes_per_elem(uint32_t dtype) { return dtype == 1 ? 4 : dtype == 2 ? 2 : dtype == 3 ? 1 : 0; }
inline size_t bytes_for_vectors_typed(uint64_t count, uint32_t dim, uint32_t dtype) {
  return (size_t)count * dim * bytes_per_elem(dtype);
}
inline size_t bytes_for_scales(uint64_t count, uint32_t dtype) { return dtype == 3 ? (size_t)count * 4 : 0; }
inline size_t bytes_for_payload_and_aux(uint64_t count, uint32_t dim, uint32_t dtype) {
  return bytes_for_vectors_typed(count, dim, dtype) + bytes_for_scales(count, dtype);
}

class MmapFile {
 public:
  MmapFile() = default;
  ~MmapFile() { close(); }
  MmapFile(const MmapFile&) = delete;
  MmapFile& operator=(const MmapFile&) = delete;
  MmapFile(MmapFile&& o) noexcept { swap(o); }
  MmapFile& operator=(MmapFile&& o) noexcept {
    if (this != &o) {
      close();
      swap(o);
    }
    return *this;
  }

  void open_readonly(const std::string& path) {
    close();
    const int fd = ::open(path.c_str(), O_RDONLY);
    if (fd < 0) throw std::runtime_error("open: " + std::string(std::strerror(errno)));
    struct stat st;
    if (::fstat(fd, &st) != 0) {
      const int e = errno;
      ::close(fd);
      throw std::runtime_error("fstat: " + std::string(std::strerror(e)));
    }
    void* p = st.st_size > 0 ? ::mmap(nullptr, (size_t)st.st_size, PROT_READ, MAP_PRIVATE, fd, 0) : MAP_FAILED;
    if (p == MAP_FAILED) {
      const int e = st.st_size > 0 ? errno : EINVAL;
      ::close(fd);
      throw std::runtime_error("mmap: " + std::string(std::strerror(e)));
    }
    ::madvise(p, (size_t)st.st_size, MADV_SEQUENTIAL);
    fd_ = fd;
    data_ = static_cast<const uint8_t*>(p);
    size_ = (size_t)st.st_size;
  }
  void close() {
    if (data_) ::munmap(const_cast<uint8_t*>(data_), size_);
    if (fd_ >= 0) ::close(fd_);
    data_ = nullptr;
    size_ = 0;
    fd_ = -1;
  }
  const uint8_t* data() const { return data_; }
  size_t size() const { return size_; }

 private:
  void swap(MmapFile& o) {
    std::swap(fd_, o.fd_);
    std::swap(data_, o.data_);
    std::swap(size_, o.size_);
  }
  int fd_ = -1;
  const uint8_t* data_ = nullptr;
  size_t size_ = 0;
};

class VectorDataset {
 public:
  void load(const std::string& path) {
    mm_.open_readonly(path);
    payload_ = nullptr;
    scales_ = nullptr;
    count_ = 0;
    dim_ = 0;
    dtype_ = 1;
    const uint8_t* p = mm_.data();
    if (mm_.size() >= 64) {
      uint64_t magic, count;
      uint32_t version, dtype, dim;
      std::memcpy(&magic, p, 8);
      std::memcpy(&version, p + 8, 4);
      std::memcpy(&dtype, p + 12, 4);
      std::memcpy(&dim, p + 16, 4);
      std::memcpy(&count, p + 24, 8);
      if (magic == kMagic && version == kVersion && dim > 0 && count > 0 && bytes_per_elem(dtype) != 0) {
        if (mm_.size() != 64 + bytes_for_payload_and_aux(count, dim, dtype))
          throw std::runtime_error("VecbinHeader ok but file size mismatch");
        set(p + 64, dtype, count, dim);
        return;
      }
    }
    if (mm_.size() < 12) throw std::runtime_error("File too small (neither vecbin64 nor raw12)");
    uint32_t rc, rd;
    std::memcpy(&rc, p, 4);
    std::memcpy(&rd, p + 8, 4);
    if (rc == 0 || rd == 0) throw std::runtime_error("raw12 header invalid (count/dim == 0)");
    if (mm_.size() != 12 + (size_t)rc * rd * 4) throw std::runtime_error("raw12 header parsed but file size mismatch");
    set(p + 12, 1, rc, rd);
  }

  uint64_t count() const { return count_; }
  uint32_t dim() const { return dim_; }
  uint32_t dtype() const { return dtype_; }

  const float* vector_ptr_f32(uint64_t i) const { return row<float>(i, 1, "Dataset is not float32"); }
  const uint16_t* vector_ptr_f16(uint64_t i) const { return row<uint16_t>(i, 2, "Dataset is not float16"); }
  const int8_t* vector_ptr_i8(uint64_t i) const { return row<int8_t>(i, 3, "Dataset is not int8"); }
  const float* scale_ptr_i8(uint64_t i) const {
    if (dtype_ != 3 || !scales_) throw std::runtime_error("Dataset is not int8 (scales missing)");
    if (i >= count_) throw std::runtime_error("Index out of range");
    return scales_ + i;
  }
  const float* vector_ptr(uint64_t i) const { return vector_ptr_f32(i); }
  const void* data_ptr() const { return payload_; }      // all dtypes (the reference returns nullptr for INT8)
  const float* scales_ptr() const { return scales_; }    // INT8 only, else nullptr

 private:
  void set(const uint8_t* payload, uint32_t dtype, uint64_t count, uint32_t dim) {
    payload_ = payload;
    dtype_ = dtype;
    count_ = count;
    dim_ = dim;
    scales_ = dtype == 3 ? reinterpret_cast<const float*>(payload + bytes_for_vectors_typed(count, dim, dtype)) : nullptr;
  }
  template <class T>
  const T* row(uint64_t i, uint32_t want, const char* what) const {
    if (dtype_ != want || !payload_) throw std::runtime_error(what);
    if (i >= count_) throw std::runtime_error("Index out of range");
    return reinterpret_cast<const T*>(payload_) + (size_t)i * dim_;
  }
  MmapFile mm_;
  const uint8_t* payload_ = nullptr;
  const float* scales_ = nullptr;
  uint64_t count_ = 0;
  uint32_t dim_ = 0, dtype_ = 1;
};

}  // namespace nvdb
