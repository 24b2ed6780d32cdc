// Host stand-in for <math_constants.h> (CUDART_INF_F lives in the cuda_runtime.h stand-in).
#pragma once
#include <cuda_runtime.h>
