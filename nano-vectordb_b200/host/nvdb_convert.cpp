// nvdb_convert_f16 / nvdb_quantize_i8 (B200) — the reference's fixture converters with the conversion on the GPU.
//
//   nvdb_convert_f16  <input_f32.vecbin> <output_f16.vecbin>      (tools/nvdb_convert_f16.cpp:132-203)
//   nvdb_quantize_i8  <in.vecbin>        <out_i8.vecbin>          (apps/nvdb_quantize_i8.cpp:18-89)
//
// One binary, behaviour chosen by argv[0] (the Makefile installs it under both names) or by a leading "f16" / "i8"
// argument. Output files are byte-identical to the reference tools' (vecbin64 header, payload, then the per-row f32
// scales for int8; tests/test_host_tools.py compares against the unmodified reference binaries). Exit codes follow the
// reference: convert_f16: 1 usage, 2 input not f32, 4 cannot open output; quantize_i8: 1 usage, 2 input not f32/f16,
// 3 cannot open output.
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "../../include/nvdb_b200.h"
#include "dataset.h"

namespace {
#pragma pack(push, 1)
struct VecbinHeader {  // include/nvdb/vecbin_format.h:17-29
  uint64_t magic;
  uint32_t version;
  uint32_t dtype;
  uint32_t dim;
  uint32_t reserved0;
  uint64_t count;
  uint8_t reserved[64 - 32];
};
#pragma pack(pop)
static_assert(sizeof(VecbinHeader) == 64, "VecbinHeader must be 64 bytes");

void write_all(std::ofstream& ofs, const char* p, size_t nbytes) {
  constexpr size_t kChunk = 64ull << 20;
  for (size_t off = 0; off < nbytes; off += kChunk) {
    ofs.write(p + off, (std::streamsize)std::min(kChunk, nbytes - off));
    if (!ofs) throw std::runtime_error("Write failed");
  }
}
}  // namespace

int main(int argc, char** argv) {
  std::string mode;
  const std::string self = argv[0];
  if (self.find("quantize_i8") != std::string::npos) mode = "i8";
  else if (self.find("convert_f16") != std::string::npos) mode = "f16";
  int a0 = 1;
  if (mode.empty() && argc >= 2) {
    mode = argv[1];
    a0 = 2;
  }
  if ((mode != "f16" && mode != "i8") || argc < a0 + 2) {
    std::cerr << (mode == "i8" ? "Usage: nvdb_quantize_i8 <in.vecbin> <out_i8.vecbin>\n"
                               : "Usage: nvdb_convert_f16 <input_f32.vecbin> <output_f16.vecbin> [batch_vecs=4096]\n");
    return 1;
  }
  try {
    const std::string in_path = argv[a0], out_path = argv[a0 + 1];
    nvdb::VectorDataset ds;
    ds.load(in_path);
    const uint64_t n = ds.count();
    const uint32_t dim = ds.dim(), dt = ds.dtype();
    if (mode == "f16" && dt != NVDB_B200_F32) {
      std::cerr << "Input dtype must be Float32. Got dtype=" << dt << "\n";
      return 2;
    }
    if (mode == "i8" && dt != NVDB_B200_F32 && dt != NVDB_B200_F16) {
      std::cerr << "Input must be f32/f16. dtype=" << dt << "\n";
      return 2;
    }
    const uint32_t dst = mode == "f16" ? NVDB_B200_F16 : NVDB_B200_I8;
    std::vector<uint8_t> payload((size_t)n * dim * (dst == NVDB_B200_F16 ? 2 : 1));
    std::vector<float> scales(dst == NVDB_B200_I8 ? n : 0);
    const int rc = nvdb_b200_convert(0, dt, ds.data_ptr(), dst, n, dim, payload.data(), scales.empty() ? nullptr : scales.data());
    if (rc != NVDB_B200_OK) {
      std::cerr << "Error: " << nvdb_b200_last_error(nullptr) << "\n";
      return 10;
    }
    std::ofstream ofs(out_path, std::ios::binary);
    if (!ofs) {
      std::cerr << (mode == "f16" ? "Failed to open output: " + out_path : std::string("Failed to open output")) << "\n";
      return mode == "f16" ? 4 : 3;
    }
    VecbinHeader hdr{};
    hdr.magic = nvdb::kMagic;
    hdr.version = nvdb::kVersion;
    hdr.dtype = dst;
    hdr.dim = dim;
    hdr.count = n;
    ofs.write(reinterpret_cast<const char*>(&hdr), sizeof(hdr));
    write_all(ofs, reinterpret_cast<const char*>(payload.data()), payload.size());
    if (dst == NVDB_B200_I8) write_all(ofs, reinterpret_cast<const char*>(scales.data()), scales.size() * sizeof(float));
    ofs.close();
    if (mode == "f16")
      std::cout << "Wrote FP16 vecbin64: " << out_path << " count=" << n << " dim=" << dim << " (payload bytes=" << payload.size()
                << ")\n";
    else
      std::cerr << "Wrote int8 vecbin64: N=" << n << " D=" << dim << "\n";
    return 0;
  } catch (const std::exception& e) {
    std::cerr << "Error: " << e.what() << "\n";
    return 10;
  }
}
