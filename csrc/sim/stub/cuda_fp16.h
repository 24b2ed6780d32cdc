// Host stand-in: the fp16 type lives in the simulator's <cuda_runtime.h>.
#pragma once
#include <cuda_runtime.h>
