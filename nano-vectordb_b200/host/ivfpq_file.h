// ivfpq_file.h — on-disk form of an IVF-PQ index built by nvdb_b200_ivfpq_build (header-only, host side).
//
// The reference saves / loads its index with faiss::write_index / read_index (apps/nvdb_ivfpq_build.cpp:131,
// apps/nvdb_ivf_eval.cpp:390); that container format belongs to FAISS and is not reproduced. This file is the arrays of
// nvdb_b200_ivfpq_export behind a 64-byte little-endian header, in the style of the reference's own vecbin / gtbin headers
// (include/nvdb/vecbin_format.h:17-29):
//
//   u64 magic "NVDBIVF1" | u32 version=1 | u32 dim | u32 nlist | u32 pq_m | u32 pq_bits=8 | u32 has_opq | u64 n | pad to 64
//   f32 coarse[nlist*dim] | f32 codebooks[pq_m*256*(dim/pq_m)] | u32 list_off[nlist+1] | u32 ids[n] | u8 codes[n*pq_m]
//   | f32 rotation[dim*dim] (only when has_opq = 1: the OPQ rotation R, x' = x R)
#pragma once
#include <cstdint>
#include <cstring>
#include <fstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/nvdb_b200.h"

namespace nvdb {

#pragma pack(push, 1)
struct IvfPqFileHeader {
  uint64_t magic;
  uint32_t version, dim, nlist, pq_m, pq_bits, has_opq;
  uint64_t n;
  uint8_t reserved[64 - 40];
};
#pragma pack(pop)
static_assert(sizeof(IvfPqFileHeader) == 64, "IvfPqFileHeader must be 64 bytes");
constexpr uint64_t kIvfPqMagic = 0x314656494244564EULL;  // "NVDBIVF1"

struct IvfPqArrays {
  uint32_t dim = 0, nlist = 0, pq_m = 0;
  uint64_t n = 0;
  std::vector<float> coarse, codebooks;
  std::vector<uint32_t> list_off, ids;
  std::vector<uint8_t> codes;
  std::vector<float> rotation;  // empty without OPQ
  void resize() {
    coarse.resize((size_t)nlist * dim);
    codebooks.resize((size_t)pq_m * 256 * (dim / pq_m));
    list_off.resize((size_t)nlist + 1);
    ids.resize(n);
    codes.resize((size_t)n * pq_m);
  }
};

inline void ivfpq_save(const nvdb_b200_ivfpq* ivf, const std::string& path) {
  IvfPqArrays a;
  uint64_t bytes = 0;
  if (nvdb_b200_ivfpq_info(ivf, &a.n, &a.dim, &a.nlist, &a.pq_m, &bytes) != NVDB_B200_OK)
    throw std::runtime_error(nvdb_b200_ivfpq_last_error(ivf));
  a.resize();
  if (nvdb_b200_ivfpq_export(ivf, a.coarse.data(), a.codebooks.data(), a.list_off.data(), a.ids.data(), a.codes.data()) !=
      NVDB_B200_OK)
    throw std::runtime_error(nvdb_b200_ivfpq_last_error(ivf));
  int has_rot = 0;
  if (nvdb_b200_ivfpq_rotation(ivf, nullptr, &has_rot) != NVDB_B200_OK) throw std::runtime_error(nvdb_b200_ivfpq_last_error(ivf));
  if (has_rot) {
    a.rotation.resize((size_t)a.dim * a.dim);
    if (nvdb_b200_ivfpq_rotation(ivf, a.rotation.data(), &has_rot) != NVDB_B200_OK)
      throw std::runtime_error(nvdb_b200_ivfpq_last_error(ivf));
  }
  std::ofstream out(path, std::ios::binary);
  if (!out) throw std::runtime_error("Failed to open output: " + path);
  IvfPqFileHeader h{};
  h.magic = kIvfPqMagic;
  h.version = 1;
  h.dim = a.dim;
  h.nlist = a.nlist;
  h.pq_m = a.pq_m;
  h.pq_bits = 8;
  h.has_opq = has_rot ? 1u : 0u;
  h.n = a.n;
  auto put = [&](const void* p, size_t nbytes) { out.write(reinterpret_cast<const char*>(p), (std::streamsize)nbytes); };
  put(&h, sizeof(h));
  put(a.coarse.data(), a.coarse.size() * 4);
  put(a.codebooks.data(), a.codebooks.size() * 4);
  put(a.list_off.data(), a.list_off.size() * 4);
  put(a.ids.data(), a.ids.size() * 4);
  put(a.codes.data(), a.codes.size());
  if (has_rot) put(a.rotation.data(), a.rotation.size() * 4);
  if (!out) throw std::runtime_error("Write failed: " + path);
}

inline nvdb_b200_ivfpq* ivfpq_load(const std::string& path, int device) {
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("Failed to read index: " + path);
  IvfPqFileHeader h{};
  in.read(reinterpret_cast<char*>(&h), sizeof(h));
  if (!in || h.magic != kIvfPqMagic || h.version != 1 || h.pq_bits != 8 || h.pq_m == 0 || h.dim % h.pq_m != 0)
    throw std::runtime_error("Bad IVF-PQ index header: " + path);
  {  // the arrays are sized from the header: refuse a header the file cannot back before allocating anything
    in.seekg(0, std::ios::end);
    const uint64_t file_bytes = (uint64_t)in.tellg();
    in.seekg(sizeof(h), std::ios::beg);
    const uint64_t dsub = h.dim / h.pq_m;
    const uint64_t want = sizeof(h) + (uint64_t)h.nlist * h.dim * 4 + (uint64_t)h.pq_m * 256 * dsub * 4 +
                          ((uint64_t)h.nlist + 1) * 4 + (uint64_t)h.n * 4 + (uint64_t)h.n * h.pq_m +
                          (h.has_opq ? (uint64_t)h.dim * h.dim * 4 : 0);
    if (h.nlist == 0 || h.n == 0 || file_bytes != want)
      throw std::runtime_error("IVF-PQ index file size does not match its header: " + path);
  }
  IvfPqArrays a;
  a.dim = h.dim;
  a.nlist = h.nlist;
  a.pq_m = h.pq_m;
  a.n = h.n;
  a.resize();
  auto get = [&](void* p, size_t nbytes) { in.read(reinterpret_cast<char*>(p), (std::streamsize)nbytes); };
  get(a.coarse.data(), a.coarse.size() * 4);
  get(a.codebooks.data(), a.codebooks.size() * 4);
  get(a.list_off.data(), a.list_off.size() * 4);
  get(a.ids.data(), a.ids.size() * 4);
  get(a.codes.data(), a.codes.size());
  if (h.has_opq) {
    a.rotation.resize((size_t)a.dim * a.dim);
    get(a.rotation.data(), a.rotation.size() * 4);
  }
  if (!in) throw std::runtime_error("IVF-PQ index file truncated: " + path);
  nvdb_b200_ivfpq* ivf = nullptr;
  if (nvdb_b200_ivfpq_import(device, a.dim, a.nlist, a.pq_m, a.n, a.coarse.data(), a.codebooks.data(), a.list_off.data(),
                             a.ids.data(), a.codes.data(), &ivf) != NVDB_B200_OK)
    throw std::runtime_error(nvdb_b200_ivfpq_last_error(nullptr));
  if (h.has_opq && nvdb_b200_ivfpq_set_rotation(ivf, a.rotation.data()) != NVDB_B200_OK) {
    const std::string msg = nvdb_b200_ivfpq_last_error(ivf);
    nvdb_b200_ivfpq_destroy(ivf);
    throw std::runtime_error(msg);
  }
  return ivf;
}

}  // namespace nvdb
