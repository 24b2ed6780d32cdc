// Host stand-in for <cuda.h>: the driver types the sources embed or pass to cuTensorMapEncodeTiled. A simulated tensor map
// keeps the arguments of the encode call in the 128 opaque bytes; csrc/sim/shim/device/ptx.cuh (tma_load_3d) interprets them.
#pragma once
#include <stdint.h>
#include <string.h>
struct alignas(64) CUtensorMap {
  uint64_t opaque[16];
};
typedef int CUresult;
constexpr CUresult CUDA_SUCCESS = 0;
typedef uint64_t cuuint64_t;
typedef uint32_t cuuint32_t;
enum CUtensorMapDataType { CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 = 9 };
enum CUtensorMapInterleave { CU_TENSOR_MAP_INTERLEAVE_NONE = 0 };
enum CUtensorMapSwizzle { CU_TENSOR_MAP_SWIZZLE_NONE = 0, CU_TENSOR_MAP_SWIZZLE_128B = 3 };
enum CUtensorMapL2promotion { CU_TENSOR_MAP_L2_PROMOTION_L2_256B = 3 };
enum CUtensorMapFloatOOBfill { CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE = 0 };

struct SimTmap {  // <= 128 bytes
  char* base;
  uint64_t dims[3];     // elements, dim 0 contiguous
  uint64_t strides[2];  // bytes, of dims 1 and 2
  uint32_t box[3];
  uint32_t rank, elt_bytes, swizzle;
};
static_assert(sizeof(SimTmap) <= sizeof(CUtensorMap), "SimTmap must fit in the opaque tensor map");
// same checks the driver makes that matter to the callers here: 16-byte aligned base and strides, box row of at most 128 B
// for the 128 B swizzle, at most 256 elements per box dimension
inline CUresult sim_tensor_map_encode_tiled(CUtensorMap* out, CUtensorMapDataType, cuuint32_t rank, void* base, const cuuint64_t* dims,
                                            const cuuint64_t* strides, const cuuint32_t* box, const cuuint32_t* elem_strides,
                                            CUtensorMapInterleave, CUtensorMapSwizzle swizzle, CUtensorMapL2promotion,
                                            CUtensorMapFloatOOBfill) {
  if (rank != 3 || (reinterpret_cast<uintptr_t>(base) & 15u) != 0) return 1;
  SimTmap t;
  memset(&t, 0, sizeof(t));
  t.base = static_cast<char*>(base);
  t.rank = rank;
  t.elt_bytes = 2;
  t.swizzle = uint32_t(swizzle);
  for (int i = 0; i < 3; ++i) {
    if (box[i] == 0 || box[i] > 256 || elem_strides[i] != 1 || dims[i] == 0) return 1;
    t.dims[i] = dims[i];
    t.box[i] = box[i];
  }
  for (int i = 0; i < 2; ++i) {
    if (strides[i] % 16 != 0) return 1;
    t.strides[i] = strides[i];
  }
  if (swizzle == CU_TENSOR_MAP_SWIZZLE_128B && box[0] * t.elt_bytes > 128) return 1;
  memset(out, 0, sizeof(*out));
  memcpy(out, &t, sizeof(t));
  return CUDA_SUCCESS;
}
