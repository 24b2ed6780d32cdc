#include "driver.h"

#include <mutex>

namespace tdist {

static thread_local std::string g_err;
void set_error(const std::string& msg) { g_err = msg; }
const char* last_error() { return g_err.c_str(); }

template <typename F>
static bool resolve(const char* name, F*& out, std::string& err) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
  if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || fn == nullptr) {
    cudaGetLastError();
    err = std::string("driver entry point not found: ") + name;
    return false;
  }
  out = reinterpret_cast<F*>(fn);
  return true;
}

const DriverApi& driver() {
  static DriverApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    DriverApi& a = api;
    bool ok = true;
#define R(field, name) ok = ok && resolve(name, a.field, a.err)
    R(cuGetErrorString, "cuGetErrorString");
    R(cuDeviceGet, "cuDeviceGet");
    R(cuDeviceGetAttribute, "cuDeviceGetAttribute");
    R(cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity");
    R(cuMemCreate, "cuMemCreate");
    R(cuMemRelease, "cuMemRelease");
    R(cuMemExportToShareableHandle, "cuMemExportToShareableHandle");
    R(cuMemImportFromShareableHandle, "cuMemImportFromShareableHandle");
    R(cuMemAddressReserve, "cuMemAddressReserve");
    R(cuMemAddressFree, "cuMemAddressFree");
    R(cuMemMap, "cuMemMap");
    R(cuMemUnmap, "cuMemUnmap");
    R(cuMemSetAccess, "cuMemSetAccess");
    R(cuStreamWriteValue32, "cuStreamWriteValue32");
    R(cuStreamWaitValue32, "cuStreamWaitValue32");
    R(cuStreamWriteValue64, "cuStreamWriteValue64");
    R(cuStreamWaitValue64, "cuStreamWaitValue64");
    R(cuTensorMapEncodeTiled, "cuTensorMapEncodeTiled");
#undef R
    // multicast is optional (absent on old drivers): failure here only disables NVLS
    std::string ignore;
    resolve("cuMulticastCreate", a.cuMulticastCreate, ignore);
    resolve("cuMulticastAddDevice", a.cuMulticastAddDevice, ignore);
    resolve("cuMulticastBindMem", a.cuMulticastBindMem, ignore);
    resolve("cuMulticastGetGranularity", a.cuMulticastGetGranularity, ignore);
    resolve("cuMulticastUnbind", a.cuMulticastUnbind, ignore);
    a.ok = ok;
  });
  return api;
}

std::string cu_err(CUresult r) {
  const DriverApi& d = driver();
  const char* s = nullptr;
  if (d.cuGetErrorString && d.cuGetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "CUresult " + std::to_string(static_cast<int>(r));
}

}  // namespace tdist
