// exchange.h — host side of the peer-memory exchange handle, shared by exchange.cu and capi.cu (internal).
#pragma once
#include <string>

#include "scan.h"

struct nvdb_b200_exchange {
  int device = 0;
  uint32_t world = 1, rank = 0, epoch = 0;
  size_t slot_bytes = 0, window_bytes = 0;
  unsigned char* window = nullptr;                    // own window (cudaMalloc: CUDA IPC needs a plain allocation)
  unsigned char* peers[nvdb_b200::kXchgMaxWorld] = {};  // windows as mapped here; peers[rank] == window
  bool opened[nvdb_b200::kXchgMaxWorld] = {};         // mapped with cudaIpcOpenMemHandle (to be closed)
  bool connected = false;
  uint32_t* counter = nullptr;                        // [0] CTA arrival counter, [1] sticky status
  unsigned char* staging = nullptr;                   // slot_bytes: packed local result of the two-launch form
  int num_sms = 148;
  unsigned long long timeout_ns = 30ull * 1000000000ull;
  std::string err;
};

namespace nvdb_b200 {
// Start the next epoch and describe it for a kernel tail (scan_gemv.cu) — the caller must launch that kernel.
void exchange_begin_epoch(nvdb_b200_exchange* x, ScanXchg* out);
}  // namespace nvdb_b200
