// gemm_fp8.cu — host side of the block-scaled fp8 GEMM (gemm_fp8_sm100.cuh): tensor maps over the e4m3 operands, launch.
#include "../runtime/driver.h"
#include "gemm_fp8_sm100.cuh"

namespace tdist {

// 2-D K-major byte tensor [rows, cols] with row stride ld (bytes); box = [128 rows, 128 bytes], 128B swizzle
static int make_tmap_u8(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld) {
  const DriverApi& d = driver();
  if (!d.ok) {
    set_error("CUDA driver not available: " + d.err);
    return -1;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kF8BlockK), 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = d.cuTensorMapEncodeTiled(out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (fp8) failed: " + cu_err(r));
    return -1;
  }
  return 0;
}

}  // namespace tdist

using namespace tdist;

extern "C" int tdist_gemm_fp8(const void* a8, const float* sa, long long sa_ld, const void* b8, const float* sb, void* c, int M, int N,
                              int K, long long lda, long long ldb, long long ldc, int out_dtype, const int* block_expert,
                              int num_experts, int num_sms, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) {
    set_error("gemm_fp8: empty problem");
    return -1;
  }
  if (K % 16 != 0 || lda % 16 != 0 || ldb % 16 != 0 || N % 8 != 0 || ldc % 8 != 0) {
    set_error("gemm_fp8: K, lda, ldb must be multiples of 16 bytes, N and ldc multiples of 8");
    return -1;
  }
  if (out_dtype != TDIST_BF16 && out_dtype != TDIST_F16) {
    set_error("gemm_fp8: output must be bf16 or fp16");
    return -1;
  }
  if (sa_ld < (K + kF8BlockK - 1) / kF8BlockK) {
    set_error("gemm_fp8: sa needs one scale per row and 128-wide K block");
    return -1;
  }
  if (block_expert != nullptr && N % kF8BlockN != 0) {
    set_error("gemm_fp8 (grouped): N must be a multiple of 128 (expert weight slabs are tiled by 128 rows)");
    return -1;
  }
  Fp8Params p;
  memset(&p, 0, sizeof(p));
  p.c = c;
  p.ldc = ldc;
  p.M = M;
  p.N = N;
  p.K = K;
  p.sa = sa;
  p.sa_ld = sa_ld;
  p.sb = sb;
  p.block_expert = block_expert;
  p.is_bf16 = out_dtype == TDIST_BF16;
  const int experts = block_expert != nullptr ? (num_experts > 0 ? num_experts : 1) : 1;
  if (make_tmap_u8(&p.tma_a, a8, M, K, lda) != 0) return -1;
  if (make_tmap_u8(&p.tma_b, b8, uint64_t(N) * experts, K, ldb) != 0) return -1;
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int use_sms = num_sms > 0 && num_sms < sms ? num_sms : sms;
  const int tiles = ((M + kF8BlockM - 1) / kF8BlockM) * ((N + kF8BlockN - 1) / kF8BlockN);
  static bool attr_set = false;
  if (!attr_set) {
    TDIST_RT_TRY(cudaFuncSetAttribute(gemm_fp8_sm100_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, Fp8Smem::kTotal));
    attr_set = true;
  }
  gemm_fp8_sm100_kernel<<<tiles < use_sms ? tiles : use_sms, kF8Threads, Fp8Smem::kTotal, static_cast<cudaStream_t>(stream)>>>(p);
  return check_launch("gemm_fp8");
}
