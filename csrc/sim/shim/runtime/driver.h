// Host stand-in for csrc/runtime/driver.h, used by the protocol simulator: the one driver entry point the host code
// compiled into the simulator needs (csrc/mega/builder.cc encodes a tensor map per linear op).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

namespace tdist {

struct DriverApi {
  bool ok = true;
  std::string err;
  CUresult (*cuTensorMapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill) = sim_tensor_map_encode_tiled;
};
inline const DriverApi& driver() {
  static const DriverApi d;
  return d;
}
void set_error(const std::string& msg);
inline std::string cu_err(CUresult r) { return "simulated driver error " + std::to_string(r); }

}  // namespace tdist

#define TDIST_RT_TRY(expr)                                                           \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      ::tdist::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e)); \
      return -1;                                                                     \
    }                                                                                \
  } while (0)
