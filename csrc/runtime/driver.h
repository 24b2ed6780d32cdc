// driver.h — lazily resolved CUDA *driver* entry points.
//
// The library never links libcuda.so: the shared object must also load on a
// box without a GPU (CPU-only unit tests, `build()` check). Each symbol is
// looked up with dlsym on first use; a missing driver yields a clean error
// string instead of a loader failure.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include <string>

namespace tdist {

struct DriverApi {
  bool ok = false;
  std::string err;
  CUresult (*cuGetErrorString)(CUresult, const char**) = nullptr;
  CUresult (*cuDeviceGet)(CUdevice*, int) = nullptr;
  CUresult (*cuDeviceGetAttribute)(int*, CUdevice_attribute, CUdevice) = nullptr;
  CUresult (*cuMemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*,
                                            CUmemAllocationGranularity_flags) = nullptr;
  CUresult (*cuMemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*,
                          unsigned long long) = nullptr;
  CUresult (*cuMemRelease)(CUmemGenericAllocationHandle) = nullptr;
  CUresult (*cuMemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType,
                                           unsigned long long) = nullptr;
  CUresult (*cuMemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*,
                                             CUmemAllocationHandleType) = nullptr;
  CUresult (*cuMemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long) = nullptr;
  CUresult (*cuMemAddressFree)(CUdeviceptr, size_t) = nullptr;
  CUresult (*cuMemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long) = nullptr;
  CUresult (*cuMemUnmap)(CUdeviceptr, size_t) = nullptr;
  CUresult (*cuMemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t) = nullptr;
  CUresult (*cuMulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*) = nullptr;
  CUresult (*cuMulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice) = nullptr;
  CUresult (*cuMulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t,
                                 size_t, unsigned long long) = nullptr;
  CUresult (*cuMulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t) = nullptr;
  CUresult (*cuMulticastGetGranularity)(size_t*, const CUmulticastObjectProp*,
                                        CUmulticastGranularity_flags) = nullptr;
  CUresult (*cuStreamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;
  CUresult (*cuStreamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int) = nullptr;
  CUresult (*cuStreamWriteValue64)(CUstream, CUdeviceptr, cuuint64_t, unsigned int) = nullptr;
  CUresult (*cuStreamWaitValue64)(CUstream, CUdeviceptr, cuuint64_t, unsigned int) = nullptr;
  CUresult (*cuTensorMapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                     CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                     CUtensorMapFloatOOBfill) = nullptr;
};

// Returns the process-wide table; `ok` is false (and `err` set) without a driver.
const DriverApi& driver();

// Last error message of this thread (C API: tdist_last_error()).
void set_error(const std::string& msg);
const char* last_error();

std::string cu_err(CUresult r);

}  // namespace tdist

#define TDIST_CU_TRY(expr)                                                                        \
  do {                                                                                            \
    CUresult _r = (expr);                                                                         \
    if (_r != CUDA_SUCCESS) {                                                                     \
      ::tdist::set_error(std::string(#expr) + " failed: " + ::tdist::cu_err(_r));                 \
      return -1;                                                                                  \
    }                                                                                             \
  } while (0)

#define TDIST_RT_TRY(expr)                                                                        \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess) {                                                                      \
      ::tdist::set_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e));              \
      return -1;                                                                                  \
    }                                                                                             \
  } while (0)
