// Host stand-in for <cuda.h>: the one driver type the GEMM parameter block embeds (the simulator never touches tensor maps).
#pragma once
#include <stdint.h>
struct alignas(64) CUtensorMap {
  uint64_t opaque[16];
};
