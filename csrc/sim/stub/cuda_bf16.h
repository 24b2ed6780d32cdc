// Host stand-in: the bf16 type lives in the simulator's <cuda_runtime.h>.
#pragma once
#include <cuda_runtime.h>
