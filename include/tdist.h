/* tdist.h — C ABI of libtdist (B200 / sm_100a).
 *
 * Every entry point returns 0 on success and -1 on failure; tdist_last_error()
 * then returns a message for the calling thread. Pointers are raw device
 * pointers, `stream` is a cudaStream_t. Nothing here synchronises the device
 * unless stated. Citations name the reference component an entry point stands
 * in for (paths relative to the reference checkout).
 */
#ifndef TDIST_H_
#define TDIST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TDIST_MAX_RANKS 16
#define TDIST_MAX_CTAS 160 /* flag rows reserved per collective context */

enum { TDIST_BF16 = 0, TDIST_F16 = 1, TDIST_F32 = 2, TDIST_FP8_E4M3 = 3 };

/* Symmetric-heap view handed BY VALUE to kernels (mirrors tdist::SymmCtx). */
typedef struct tdist_symm_ctx {
  char* base[TDIST_MAX_RANKS]; /* base[r]: where rank r's heap is mapped in this process */
  char* mc_base;               /* NVLS multicast alias of the heap, NULL if unavailable */
  unsigned long long bytes;    /* bytes per rank */
  int rank;
  int world;
} tdist_symm_ctx;

typedef struct tdist_heap tdist_heap;

const char* tdist_last_error(void);

/* ---- symmetric heap (replaces nvshmem.core.tensor / get_peer_tensor,
 *      python/triton_dist/utils.py:99-142) ------------------ */
/* Collective over the ranks of one node. VMM allocation, fd exchange over
 * abstract UNIX sockets named after job_id, peers mapped at base + p*bytes. */
int tdist_heap_create(int rank, int world, int device, size_t bytes, const char* job_id, tdist_heap** out);
/* cudaMalloc + cudaIpc fallback: step 1 returns this rank's 64-byte handle ... */
int tdist_heap_create_ipc(int rank, int world, int device, size_t bytes, tdist_heap** out, void* handle64);
/* ... step 2 opens all peers (world*64 bytes, rank order). */
int tdist_heap_open_ipc(tdist_heap* h, const void* all_handles);
/* NVLS multicast binding in two collective phases; barrier between them. */
int tdist_heap_mc_phase_a(tdist_heap* h);
int tdist_heap_mc_phase_b(tdist_heap* h);
void* tdist_heap_base(tdist_heap* h, int peer);
void* tdist_heap_mc_base(tdist_heap* h);
size_t tdist_heap_bytes(tdist_heap* h);
int tdist_heap_mode(tdist_heap* h); /* 0 local, 1 VMM, 2 cudaIpc */
int tdist_heap_fill_ctx(tdist_heap* h, tdist_symm_ctx* out);
int tdist_heap_destroy(tdist_heap* h);

/* ---- stream memory ops (kernels/nvidia/common_ops.py:347-389) ------------- */
int tdist_stream_write_value32(void* stream, void* addr, unsigned int value);
int tdist_stream_wait_value32(void* stream, void* addr, unsigned int value, int cmp_geq);
int tdist_stream_write_value64(void* stream, void* addr, unsigned long long value);
int tdist_stream_wait_value64(void* stream, void* addr, unsigned long long value, int cmp_geq);

/* ---- collectives. flags: symmetric uint32[TDIST_MAX_CTAS*TDIST_MAX_RANKS];
 *      phase/done: device uint32 counters (zero initialised) ---------------- */
/* stream-ordered barrier across ranks (utils.py:162-164, common_ops.py:141-209) */
int tdist_barrier_all(const tdist_symm_ctx* ctx, uint32_t* flags, uint32_t* phase, uint32_t* done, void* stream);

enum { TDIST_AR_ONE_SHOT = 1, TDIST_AR_TWO_SHOT = 2, TDIST_AR_ONE_SHOT_MC = 5, TDIST_AR_TWO_SHOT_MC = 6 };
/* kernels/nvidia/allreduce.py:332-682. ws is symmetric; in_place means `in`
 * already lives in ws (multimem methods only). nbytes % 16 == 0. */
int tdist_allreduce(const tdist_symm_ctx* ctx, int method, int dtype, const void* in, void* out, size_t nbytes,
                    void* ws, size_t ws_bytes, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas,
                    int in_place, void* stream);

enum { TDIST_AG_PULL = 0, TDIST_AG_PUSH = 1, TDIST_AG_PUSH_MC = 2, TDIST_AG_LL = 3, TDIST_AG_LL_MC = 4 };
/* kernels/nvidia/low_latency_allgather.py:47-777. symm: [W][shard] (LL modes:
 * [2][W][shard/8] 16-byte packets). out may be NULL (result stays in symm). */
int tdist_allgather(const tdist_symm_ctx* ctx, int mode, const void* in, void* symm, void* out, size_t shard_bytes,
                    size_t symm_bytes /* LL modes: size of symm; its parity halves are symm_bytes/2 apart */, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas, void* stream);

/* kernels/nvidia/reduce_scatter.py:285-866. in: [W][seg]; scatter_ws (symm): two parity halves ws_bytes/2 apart, each
 * [W][seg]; out: [seg] */
int tdist_reduce_scatter(const tdist_symm_ctx* ctx, int dtype, const void* in, void* scatter_ws, size_t ws_bytes, void* out,
                         size_t seg_bytes, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_ctas,
                         void* stream);
/* out = sum of `slots` segments (ring_reduce, reduce_scatter.py:673-810) */
int tdist_slot_reduce(int dtype, const void* in, void* out, size_t seg_bytes, int slots, int begin, int max_ctas,
                      void* stream);
/* kernels/nvidia/p2p.py:30-85 */
int tdist_p2p_copy(void* dst, const void* src, size_t nbytes, int num_ctas, void* stream);
/* one-thread signal kernels (graph-capturable twin of the stream mem-ops) */
int tdist_signal_set(uint32_t* flag, uint32_t value, int add, void* stream);
int tdist_signal_wait(const uint32_t* flag, uint32_t value, int cmp, void* stream);

/* ---- tcgen05 GEMM engine ------------------------------------ */
typedef struct tdist_gemm_args {
  const void* a; /* [M, K] row-major (AG: this rank's shard [M/W, K]) */
  const void* b; /* [N, K] row-major (nn.Linear weight layout) */
  void* c;       /* [M, N] row-major, or [N, M] when transpose_out */
  int M, N, K;
  long long lda, ldb, ldc; /* in elements */
  int dtype;               /* TDIST_BF16 / TDIST_F16 */
  int block_n;             /* 0 = heuristic */
  int group_m;             /* 0 = default */
  int num_sms;             /* 0 = all */
  int transpose_out;       /* store C^T (swap-AB skinny GEMM) */
  /* grouped (MoE) GEMM: b is [num_experts * N, K]; block_expert[i] (device) = expert of 128-row block i of A,
   * negative = padding block (skipped). NULL = ordinary GEMM. (allgather_group_gemm.py:534-608, moe_reduce_rs.py:166-244) */
  const int* block_expert;
  int num_experts;
  int cta_group; /* 0 = auto, 1 = one CTA per 128-row tile, 2 = CTA pairs on 256-row tiles (tcgen05 cta_group::2) */
} tdist_gemm_args;

/* C = A * B^T (kernels/nvidia/allgather_gemm.py:656-720 gemm_persistent) */
int tdist_gemm(const tdist_gemm_args* a, void* stream);

/* Block-scaled fp8 GEMM (csrc/kernels/gemm_fp8_sm100.cuh; the reference has no fp8 GEMM — fp8 appears there only as the
 * all-to-all payload, tutorials/04-deepseek-infer-all2all.py:334-427): C[M,N] (bf16/fp16) = dequant(A8) * dequant(B8)^T with
 * A8 [M,K] / B8 [N,K] e4m3 K-major, sa [M][sa_ld] one fp32 scale per row and 128-wide K block, sb [E][ceil(N/128)][ceil(K/128)]
 * one per 128 x 128 weight block. Grouped (MoE) form: block_expert[i] = expert of 128-row block i of A8 (negative: padding
 * block, skipped), B8 = [num_experts * N, K]. lda / ldb in bytes (= elements), ldc in elements. */
int tdist_gemm_fp8(const void* a8, const float* sa, long long sa_ld, const void* b8, const float* sb, void* c, int M, int N,
                   int K, long long lda, long long ldb, long long ldc, int out_dtype, const int* block_expert, int num_experts,
                   int num_sms, void* stream);

enum { TDIST_GEMM_AG = 1, TDIST_GEMM_RS = 2, TDIST_GEMM_AR = 3 };
typedef struct tdist_fused_args {
  tdist_symm_ctx ctx;
  int mode;
  void* ws;           /* AG: symm gather workspace [2][M][K]; RS: symm scatter [2][W][M/W][N]; AR: symm tiles [2][M][ldc] */
  void* out;          /* RS: [M/W, N]; AR: [M, ldc] (symmetric when !ar_one_shot) */
  uint32_t* flags;    /* symmetric, zero initialised, >= W*ceil(M/W/128) (AG/RS point-to-point) or W * #tiles (AR) entries */
  uint32_t* flags2;   /* AR: symmetric, zero initialised, >= 64 words (closing flags + a local counter) */
  uint32_t* counters; /* RS: local, zero initialised, W*ceil(M/W/128) entries */
  uint32_t* phase;    /* device counter pair, zero initialised */
  uint32_t* done;
  int num_comm_ctas;  /* <0 = default */
  int ar_one_shot;
  int use_mc;
  int variant;        /* protocol: 0 = auto (NVLS when the heap has a multicast mapping), 1 = point-to-point (AG: unicast
                         push ring, RS: scatter + slot reduction), 2 = NVLS (AG: multimem.st broadcast push, RS: partial
                         tiles stay local and the owner sums them with multimem.ld_reduce). NVLS layouts: AG flags
                         >= W*ceil(M/W/128)*128 words; RS ws = [2][max_M][N] partial buffer, flags >= W*ceil(M/W/128)*ceil(N/64) */
  size_t ws_bytes;    /* size of ws; its two parity halves are ws_bytes/2 apart whatever the size of this call */
  size_t flags_words; /* capacity of flags (checked against the layout of the chosen protocol); 0 = unchecked */
} tdist_fused_args;

/* AG+GEMM (allgather_gemm.py:534-653), GEMM+RS (gemm_reduce_scatter.py:508-583),
 * GEMM+AR (gemm_allreduce.py:509-575): one launch each. */
int tdist_gemm_fused(const tdist_gemm_args* a, const tdist_fused_args* f, void* stream);
/* host-side view of the single-CTA kernel's tile walk (the same tile_coord the kernel calls, no CUDA call):
 * out3[3t..] = (shard, 128-row block inside the shard, n block); mode 0 plain, 1 AG, 2 RS; returns the number of tiles */
int tdist_gemm_tile_order(int mode, int M, int N, int world, int rank, int block_n, int group_m, int* out3, int cap_tiles);

/* Self-test of the NVSHMEM-style device primitives (csrc/kernels/primitives_selftest.cu; no counterpart in the reference,
 * which tests them through Triton). `buf`: symmetric, >= (3 * world + 2) * 16 bytes, zeroed once; `flags`: symmetric,
 * >= 2 * world + 1 words, zeroed once; `epoch` must grow with every call. Result: the int at buf + 3 * world * 16 equals
 * `epoch` on success, -1 otherwise. Call on every rank. */
int tdist_primitives_selftest(const tdist_symm_ctx* ctx, void* buf, uint32_t* flags, uint32_t epoch, void* stream);

/* Bandwidth probes of the data paths the fused kernels are built from (csrc/kernels/link_probe.cu; pure data movers, no
 * cross-rank waits). mode: 0 local copy, 1 NVLS broadcast push, 2 unicast push to the peers, 3 pull from the peers, 4 NVLS
 * pull-reduce, 5 bulk-copy ring to the multicast address, 6 bulk-copy ring pulling from a peer. `src` / `out`: local
 * buffers of `bytes` (x world for the pull modes, see the source); `symm`: symmetric [world][bytes]. bytes % 16 == 0. */
int tdist_link_probe(const tdist_symm_ctx* ctx, int mode, const void* src, void* symm, void* out, size_t bytes, int grid,
                     int threads, int unroll, void* stream);

/* ---- MoE helpers (csrc/kernels/moe.cu; reference csrc/lib/moe_utils.cu, kernels/nvidia/moe_utils.py) ---- */
/* sort (token,k) pairs by expert, pad every expert to block_m rows. meta[0] = padded rows, meta[1] = blocks. */
int tdist_moe_align_sort(const int* topk_ids, int num_pairs, int num_experts, int block_m, int* sorted_pair_ids,
                         int* pair_to_row, int* block_expert, int* meta, int max_rows, void* stream);
/* AG+MoE tile order: per 128-row block the range of arrival stages it needs and a stable block order by last stage
 * (stage s = shard of rank (rank + s) % world). stage_count[s] = blocks whose last stage is s. */
int tdist_moe_tile_swizzle(const int* sorted_pair_ids, const int* meta, int num_pairs, int topk, int tokens_per_rank, int rank,
                           int world, int block_m, int* block_order, int* seg_start, int* seg_end, int* stage_count,
                           int max_blocks, void* stream);
int tdist_moe_gather_rows(const void* src, void* dst, const int* sorted_pair_ids, const int* meta, int num_pairs,
                          int topk, int row_bytes, int max_rows, void* stream);
/* Same gather restricted to the row blocks with block_stage[row / block_m] == stage (other rows of dst are not written):
 * the stage-gated AG + grouped GEMM fills one buffer stage by stage as the token shards arrive. */
int tdist_moe_gather_rows_stage(const void* src, void* dst, const int* sorted_pair_ids, const int* meta, const int* block_stage,
                                int stage, int block_m, int num_pairs, int topk, int row_bytes, int max_rows, void* stream);
int tdist_moe_topk_reduce(int dtype, const void* y, void* out, const int* pair_to_row, const float* weights, int tokens,
                          int topk, int row_bytes, void* stream);
/* fused top-k reduce + reduce-scatter over token segments (moe_reduce_rs.py:292-376) */
int tdist_moe_reduce_rs(const tdist_symm_ctx* ctx, int dtype, const void* y, void* scatter_ws, size_t ws_bytes, void* out,
                        const int* pair_to_row, const float* weights, int tokens, int topk, int row_bytes, uint32_t* flags,
                        uint32_t* phase, uint32_t* done, int max_ctas, void* stream);

/* ---- expert-parallel all-to-all (csrc/kernels/ep_a2a.cu) ------------------ */
/* fixed-slot low-latency all-to-all (kernels/nvidia/low_latency_all_to_all.py:35-257). send rows are sorted by
 * expert, splits_cumsum has num_experts+1 entries; recv_data (symm) [2][W][max_m][row_bytes], recv_splits (symm)
 * [2][W*epr], recv_scale (symm, optional) [2][W][max_m]; flags (symm) uint32[2*W] zero initialised. */
int tdist_a2a(const tdist_symm_ctx* ctx, const void* send, const int* splits_cumsum, const float* send_scale, void* recv_data,
              int* recv_splits, float* recv_scale, uint32_t* flags, uint32_t* phase, uint32_t* done, int max_m, int row_bytes,
              int experts_per_rank, int ctas_per_peer, void* stream);
/* token-level dispatch / combine (kernels/nvidia/ep_a2a.py:36-334, layers/nvidia/ep_a2a_layer.py:118-248) */
int tdist_ep_counts(const tdist_symm_ctx* ctx, const int* topk_ids, int num_pairs, int num_experts, int* full_counts,
                    uint32_t* flags, uint32_t* phase, uint32_t* done, void* stream);
int tdist_ep_dispatch(const tdist_symm_ctx* ctx, int dtype, const void* x, const int* topk_ids, const int* full_counts,
                      void* recv_buf, float* recv_scale, int* recv_src_pair, int* scatter_idx, int* cursors,
                      int* recv_expert_counts, int* num_recv, int tokens, int topk, int hidden, int num_experts, int fp8,
                      uint32_t* flags, uint32_t* grid_counter, uint32_t* phase, uint32_t* done, int num_ctas, void* stream);
int tdist_ep_combine(const tdist_symm_ctx* ctx, int dtype, const void* expert_out, const int* topk_ids, const int* scatter_idx,
                     const float* weights, void* out, int tokens, int topk, int hidden, int num_experts, uint32_t* flags,
                     uint32_t* grid_counter, uint32_t* phase, uint32_t* done, int num_ctas, void* stream);

/* ---- decode-path kernels (csrc/kernels/decode.cu) -------------------------
 * Reference: kernels/nvidia/flash_decode.py:129-532 and the mega_triton_kernel task kernels;
 * norm / rope / sampling are flashinfer calls in the reference's layers. */
/* y[M,N] = x[M,K] w[N,K]^T (+bias), 1 <= M <= 16: weight-streaming GEMV */
int tdist_gemv(int dtype, const void* x, const void* w, const void* bias, void* y, int M, int N, int K, long long ldx,
               long long ldw, long long ldy, void* stream);
/* y = rmsnorm(x + residual) * weight; residual (optional, contiguous [rows, hidden]) is updated to x + residual */
int tdist_rmsnorm(int dtype, const void* x, void* residual, const void* weight, void* y, int rows, int hidden, float eps,
                  long long ldx, long long ldy, void* stream);
/* out[t, :] = silu(gate_up[t, :I]) * gate_up[t, I:] */
int tdist_silu_mul(int dtype, const void* gate_up, void* out, long long tokens, int inter, void* stream);
int tdist_add(int dtype, const void* a, const void* b, void* out, long long numel, void* stream);
/* per-head q/k RMSNorm (weights optional) + NeoX RoPE + KV-cache append; qkv [T, (Hq+2Hkv)*128];
 * cos_sin fp32 [max_pos, 128] (cos | sin); cache dense (block_table NULL: strides batch/seq/head) or paged
 * (kv_batch_stride = page stride). */
int tdist_qk_norm_rope_kv(int dtype, const void* qkv, const void* q_norm_w, const void* k_norm_w, const float* cos_sin,
                          const int* positions, const int* batch_idx, void* q_out, void* k_cache, void* v_cache,
                          int tokens, int Hq, int Hkv, int head_dim, float eps, long long kv_batch_stride,
                          long long kv_seq_stride, long long kv_head_stride, const int* block_table, int page_size,
                          int max_pages, void* stream);
/* GQA split-KV flash decode: partial [B,Hq,splits,129] fp32 (out|lse), then (if out16 or out_lse given) the
 * intra-rank combine into out16 [B,Hq,128] and/or out_lse [B,Hq,129]. */
int tdist_flash_decode(int dtype, const void* q, const void* k_cache, const void* v_cache, const int* kv_lens,
                       float* partial, void* out16, float* out_lse, int B, int Hq, int Hkv, int head_dim, int splits,
                       float sm_scale, long long kv_batch_stride, long long kv_seq_stride, long long kv_head_stride,
                       const int* block_table, int page_size, int max_pages, void* stream);
/* log-sum-exp merge of `splits` partials with arbitrary strides (floats): inter-rank combine of distributed decode */
int tdist_flash_decode_combine(int dtype, const float* partial, void* out16, float* out_lse, int B, int H, int splits,
                               long long stride_b, long long stride_h, long long stride_s, void* stream);
/* Decode attention for callers that may not allocate device memory (no counterpart in the reference: it was added
 * for this library's own users). Same computation as tdist_flash_decode(), but the split-KV scratch is a fixed array
 * inside the library (about 2 MiB per GPU, reserved by the driver when the library is loaded and held until the process
 * exits), so the call performs no cudaMalloc/cudaFree and pulls in no allocating library. Meant for code that runs
 * where allocation is forbidden or unavailable: inside a serving loop whose pool owns all device memory, in a plugin
 * that must not use its host's allocator, or from C with no allocator wired up.
 * Limit: B * Hq * splits <= 4096 partial rows, otherwise -1 with a message (use tdist_flash_decode with your own buffer).
 * The first call on each GPU also resolves that GPU's copy of the scratch (cudaGetSymbolAddress on the current device):
 * make that first call outside any stream capture. All calls on a GPU share the one scratch, which still holds the
 * previous call's partials; calls ordered on one stream are safe, calls that may overlap on the GPU race. */
int tdist_flash_decode_noalloc(int dtype, const void* q, const void* k_cache, const void* v_cache, const int* kv_lens,
                               void* out16, int B, int Hq, int Hkv, int head_dim, int splits, float sm_scale,
                               long long kv_batch_stride, long long kv_seq_stride, long long kv_head_stride,
                               const int* block_table, int page_size, int max_pages, void* stream);
int tdist_embedding(int dtype, const long long* ids, const void* table, void* out, int tokens, int hidden, void* stream);
int tdist_argmax(int dtype, const void* logits, long long* out, int rows, int vocab, long long ld, void* stream);

/* ---- sequence-parallel prefill attention (csrc/kernels/sp_attn.cu; reference
 *      kernels/nvidia/sp_ag_attention_intra_node.py:186-425) ------------------------------------------------ */
/* varlen causal GQA flash-attention of this rank's query shard over the all-gathered K/V (global token order).
 * q/out [total_q_local, Hq, 128]; k/v gathered [total_k_global, Hkv, 128]; cu_q local / cu_k global offsets [B+1]. */
/* The same attention on tcgen05 / TMEM / TMA (csrc/kernels/sp_attn_sm100.cu): 128-row query tiles, S = Q K^T and P V on
 * tcgen05.mma with TMEM accumulators, K/V tiles by TMA; V is consumed as an MN-major operand from its [keys][d] layout.
 * total_q / total_k: rows of the q / k / v buffers; q_row_stride / kv_row_stride: elements between consecutive tokens
 * (>= heads * 128, so views into a fused qkv tensor or a KV cache work). Rows of k/v behind a sequence need not be
 * initialised. world = 1, rank = 0 gives plain varlen attention whose queries are the LAST rows of each sequence (prefill,
 * chunked prefill: TP_Attn). */
int tdist_attn_fwd_tc(const void* q, const void* k, const void* v, void* out, const int* cu_q, const int* cu_k, int B,
                      int max_q_local, int total_q, int total_k, int Hq, int Hkv, int head_dim, long long q_row_stride,
                      long long kv_row_stride, int k_batch_stride /* > 0: keys of sequence b start at row b * stride (dense
                      cache), 0: at cu_k[b] */, int rank, int world, int causal, int zigzag, float sm_scale, void* stream);
int tdist_sp_attn_fwd(const void* q, const void* k_gathered, const void* v_gathered, void* out, const int* cu_q, const int* cu_k,
                      int B, int max_q_local, int Hq, int Hkv, int head_dim, int rank, int world, int causal, int zigzag,
                      float sm_scale, void* stream);
/* scatter this rank's K (or V) shard rows into global order in a gathered buffer (local or peer-mapped) */
int tdist_sp_kv_scatter(const void* src, void* dst, const int* cu_k_global, int B, int max_local_rows, int row_bytes, int rank,
                        int world, int zigzag, void* stream);

/* ---- persistent decode megakernel (csrc/mega; reference python/triton_dist/mega_triton_kernel) ---- */
enum {
  TDIST_MEGA_LINEAR = 0, TDIST_MEGA_RMS_NORM = 1, TDIST_MEGA_SILU_MUL = 2, TDIST_MEGA_ADD = 3, TDIST_MEGA_QK_NORM_ROPE_KV = 4,
  TDIST_MEGA_ATTN_SPLIT = 5, TDIST_MEGA_ATTN_COMBINE = 6, TDIST_MEGA_ALLREDUCE = 7, TDIST_MEGA_BARRIER = 8,
  TDIST_MEGA_PREFETCH = 9, TDIST_MEGA_EMBEDDING = 10, TDIST_MEGA_ARGMAX = 11
};
/* Generic op record; the meaning of every slot per op type is documented in csrc/mega/megakernel.cu. */
typedef struct tdist_mega_op {
  const void* in[6];
  void* out[3];
  int M, N, K;
  int i[6];
  long long l[3];
  float f[2];
} tdist_mega_op;
int tdist_mega_create(int rank, int world, int num_ctas, void** out);
/* returns the op index (>= 0) or -1. dep0: -2 previous op, -1 none, >= 0 explicit; dep1: -1 none or explicit. */
int tdist_mega_add_op(void* builder, int type, const tdist_mega_op* desc, int dep0, int dep1, int chain);
/* ar_flags: symmetric uint32[(#allreduce + #barrier ops) * 32 * TDIST_MAX_RANKS], zero initialised (NULL if world == 1) */
int tdist_mega_compile(void* builder, const tdist_symm_ctx* ctx, uint32_t* ar_flags);
int tdist_mega_run(void* builder, void* stream);
int tdist_mega_info(void* builder, int* out6); /* ops, tasks, max queue length, smem bytes, CTAs, #cross-rank ops */
int tdist_mega_op_tiles(void* builder, int op_index);
/* host-only planning query (no CUDA call): out12 = cols_per_tile, tiles, row blocks, rows per block, k-slabs per stage,
 * k per stage, stages along K, stages per tile, scratch bytes, stage buffers that fit, stage buffer bytes, smem limit */
int tdist_mega_linear_plan(int M, int N, int K, int pair, int has_norm, int num_ctas, int* out12);
/* per-task timeline: stamp globaltimer at the start/end of every task of the following launches; get_trace copies
 * [num_ctas][max_tasks][6] uint64 nanosecond stamps of the last launch to the host (begin, end, linear tasks:
 * dependencies satisfied, activations staged, weight stream consumed, spare) */
int tdist_mega_set_trace(void* builder, int enable);
int tdist_mega_get_trace(void* builder, unsigned long long* out, size_t cap_words);
int tdist_mega_get_queue(void* builder, int cta, int* out_8ints_per_task, int cap);
int tdist_mega_destroy(void* builder);

/* ---- tutorial-size kernels (csrc/tutorials; run and explained by tutorials/01, 07, 08 — not used by the library) ----
 * queue: rank r streams n_items items of item_vecs 16-byte vectors into the `out` of rank r+1 through a queue of `slots`
 *   slots in r+1's heap (queue [slots][item_vecs] int4, full / empty [slots] u32: symmetric, zeroed once; base = sum of
 *   ceil(n_items / slots) over earlier calls on these buffers).
 * ag_gemm: C[W * rows, N] = all_gather(a[rows, K]) * b[N, K]^T; ws symmetric, two halves ws_bytes / 2 apart, each >=
 *   W * rows * K elements; flags symmetric [W][flag_blocks], flag_blocks >= ceil(rows / 64).
 * gemm_rs: out[M/W, N] = reduce_scatter(a[M, Kl] * b[N, Kl]^T); part symmetric, two halves part_bytes / 2 apart, each >=
 *   M * N elements; flags symmetric [W][flag_blocks][ceil(N/64)], flag_blocks >= ceil(M / W / 64).
 *   ws_bytes / part_bytes / flag_blocks (and N for gemm_rs) are constants of the buffers: calls of different sizes on the
 *   same buffers must see the same layout (csrc/sim/protocol_sim.cc shows what happens otherwise). */
int tdist_tutorial_queue(const tdist_symm_ctx* ctx, const void* in, void* out, void* queue, uint32_t* full, uint32_t* empty,
                         int n_items, int item_vecs, int slots, uint32_t base, void* stream);
int tdist_tutorial_ag_gemm(const tdist_symm_ctx* ctx, const void* a, const void* b, void* c, void* ws, size_t ws_bytes,
                           uint32_t* flags, int flag_blocks, uint32_t* phase, uint32_t* done, int rows, int N, int K,
                           int push_ctas, int ctas, void* stream);
int tdist_tutorial_gemm_rs(const tdist_symm_ctx* ctx, const void* a, const void* b, void* out, void* part, size_t part_bytes,
                           uint32_t* flags, int flag_blocks, uint32_t* phase, uint32_t* done, int M, int N, int Kl, int ctas,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TDIST_H_ */
