// gemm.cu — host side of the tcgen05 GEMM engine: tensor-map construction,
// tile-shape heuristic, launch. C entry points: tdist_gemm (plain GEMM, also the
// swap-AB skinny path) and tdist_gemm_fused (AG+GEMM / GEMM+RS / GEMM+AR).
#include <mutex>
#include <unordered_map>

#include "../runtime/driver.h"
#include "gemm_sm100_2cta.cuh"

namespace tdist {

// ---- tensor-map cache ---------------------------------------------------------
struct TmapKey {
  const void* ptr;
  uint64_t rows, cols, ld;
  uint32_t box_rows;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows;
  }
};
struct TmapHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h = h * 1000003u ^ k.rows;
    h = h * 1000003u ^ k.cols;
    h = h * 1000003u ^ k.ld;
    h = h * 1000003u ^ k.box_rows;
    return h;
  }
};

// 2-D K-major 16-bit tensor [rows, cols] with row stride ld (elements); box = [box_rows, 64], 128B swizzle
static int make_tmap(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  static std::mutex mu;
  static std::unordered_map<TmapKey, CUtensorMap, TmapHash> cache;
  const TmapKey key{ptr, rows, cols, ld, box_rows};
  {
    std::lock_guard<std::mutex> g(mu);
    auto it = cache.find(key);
    if (it != cache.end()) {
      *out = it->second;
      return 0;
    }
  }
  const DriverApi& d = driver();
  if (!d.ok) {
    set_error("CUDA driver not available: " + d.err);
    return -1;
  }
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(kBlockK), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = d.cuTensorMapEncodeTiled(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides,
                                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: " + cu_err(r));
    return -1;
  }
  std::lock_guard<std::mutex> g(mu);
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return 0;
}

static int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
  }
  return n;
}

template <int BN, int ST>
static int launch_inst(const GemmParams& p, int grid, cudaStream_t s, bool cooperative) {
  using L = GemmSmem<BN, ST>;
  static bool attr_set = false;
  auto kern = gemm_sm100_kernel<BN, ST>;
  if (!attr_set) {
    TDIST_RT_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = cooperative ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TDIST_RT_TRY(cudaLaunchKernelEx(&cfg, kern, p));
  return 0;
}

template <int BN, int ST>
static int launch_inst2(const GemmParams& p, int grid, cudaStream_t s) {
  using L = GemmSmem2<BN, ST>;
  static bool attr_set = false;
  auto kern = gemm_sm100_2cta_kernel<BN, ST>;
  if (!attr_set) {
    TDIST_RT_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kTotal));
    attr_set = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(kGemmThreads);
  cfg.dynamicSmemBytes = L::kTotal;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  TDIST_RT_TRY(cudaLaunchKernelEx(&cfg, kern, p));
  return 0;
}

// CTA-pair kernel: tile widths and stage counts (per-CTA stage = 16 KB of A + BLOCK_N * 64 B of B)
static int dispatch2(const GemmParams& p, int bn, int grid, cudaStream_t s) {
  switch (bn) {
    case 256: return launch_inst2<256, 6>(p, grid, s);
    case 224: return launch_inst2<224, 6>(p, grid, s);
    case 192: return launch_inst2<192, 7>(p, grid, s);
    case 160: return launch_inst2<160, 7>(p, grid, s);
    case 128: return launch_inst2<128, 8>(p, grid, s);
    case 112: return launch_inst2<112, 8>(p, grid, s);
    case 96: return launch_inst2<96, 8>(p, grid, s);
    case 64: return launch_inst2<64, 8>(p, grid, s);
  }
  set_error("gemm (cta pair): unsupported BLOCK_N");
  return -1;
}

// tile-width choice for the CTA-pair kernel: 256-row tiles on `pairs` CTA pairs. Narrow tiles stay efficient here
// because each CTA only stages half of B (see gemm_sm100_2cta.cuh), so the model is dominated by wave quantisation.
static int pick_block_n2(int ptiles_m, int N, int pairs, int forced) {
  static const int cands[] = {256, 224, 192, 160, 128, 112, 96, 64};
  // relative mainloop efficiency measured on B200 (M=8192 N=14336 K=4096, profiles/README.md): like the single-CTA
  // kernel, narrow tiles re-read the A operand from shared memory more often per FLOP
  static const double eff[] = {1.0, 0.78, 0.68, 0.56, 0.44, 0.42, 0.36, 0.25};
  if (forced > 0)
    for (int c : cands)
      if (c == forced) return c;
  double best_cost = -1;
  int best = 256;
  for (size_t i = 0; i < sizeof(cands) / sizeof(cands[0]); ++i) {
    const int bn = cands[i];
    const long tiles = long(ptiles_m) * ((N + bn - 1) / bn);
    const long waves = (tiles + pairs - 1) / pairs;
    const double cost = double(waves) * (bn / eff[i] + 6.0);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  return best;
}

static int want_cta_pair(const tdist_gemm_args* a) {
  if (a->cta_group == 1) return 0;
  if (a->cta_group == 2) return 1;
  static int env = -1;
  if (env < 0) {
    const char* e = getenv("TDIST_GEMM_CTA_GROUP");
    env = e ? atoi(e) : 0;
  }
  if (env == 1) return 0;
  if (env == 2) return 1;
  // default: 256-row pair tiles (each CTA stages half of B, 6 pipeline stages instead of 4: cuBLAS-level throughput)
  // whenever a full pair tile of rows exists; callers still fall back to single-CTA tiles for grouped / transposed /
  // all-reduce GEMMs
  return a->M >= 2 * kBlockM;
}

// protocol variant of a fused op: explicit argument > environment (TDIST_AG_VARIANT / TDIST_RS_VARIANT: 1 = point to
// point, 2 = NVLS) > NVLS whenever the heap has a multicast mapping
static int pick_variant(const tdist_fused_args* f, const SymmCtx& c, const char* env_name, int min_world_for_nvls) {
  int v = f->variant;
  if (v == 0) {
    const char* e = getenv(env_name);
    if (e) v = atoi(e);
  }
  if (v == 0) v = (c.mc_base != nullptr && c.world >= min_world_for_nvls) ? 2 : 1;
  if (v == 2 && c.mc_base == nullptr) v = -1;
  return v;
}
static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

static double g_last_cost = 0;
static int pick_block_n(int M_tiles, int N, int ctas, int forced) {
  static const int cands[] = {256, 224, 192, 160, 128, 112, 96, 64, 32};
  if (forced > 0) {
    for (int c : cands)
      if (c == forced) return c;
  }
  // relative mainloop efficiency of each tile width, measured on B200 (profiles/gemm_tile_sweep.md): narrow
  // tiles re-read the 128-row A operand from shared memory more often per FLOP and become smem-bound.
  static const double eff[] = {1.0, 0.943, 0.826, 0.70, 0.585, 0.52, 0.45, 0.32, 0.17};
  double best_cost = -1;
  int best = 128;
  for (size_t i = 0; i < sizeof(cands) / sizeof(cands[0]); ++i) {
    const int bn = cands[i];
    const long tiles = long(M_tiles) * ((N + bn - 1) / bn);
    const long waves = (tiles + ctas - 1) / ctas;
    const double cost = double(waves) * (bn / eff[i] + 6.0);  // +6: per-tile hand-off
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = bn;
    }
  }
  g_last_cost = best_cost;
  return best;
}

static int dispatch(const GemmParams& p, int bn, int grid, cudaStream_t s, bool coop) {
  switch (bn) {
    case 256: return launch_inst<256, 4>(p, grid, s, coop);
    case 224: return launch_inst<224, 4>(p, grid, s, coop);
    case 192: return launch_inst<192, 5>(p, grid, s, coop);
    case 160: return launch_inst<160, 5>(p, grid, s, coop);
    case 128: return launch_inst<128, 6>(p, grid, s, coop);
    case 112: return launch_inst<112, 6>(p, grid, s, coop);
    case 96: return launch_inst<96, 7>(p, grid, s, coop);
    case 64: return launch_inst<64, 8>(p, grid, s, coop);
    case 32: return launch_inst<32, 8>(p, grid, s, coop);
  }
  set_error("gemm: unsupported BLOCK_N");
  return -1;
}

static int check_common(const tdist_gemm_args* a) {
  if (a->dtype != TDIST_BF16 && a->dtype != TDIST_F16) {
    set_error("gemm: dtype must be bf16 or fp16");
    return -1;
  }
  if (a->K % 8 != 0 || a->N % 8 != 0 || a->lda % 8 != 0 || a->ldb % 8 != 0 || a->ldc % 8 != 0) {
    set_error("gemm: K, N and leading dimensions must be multiples of 8 (16-byte rows)");
    return -1;
  }
  if (a->M <= 0 || a->N <= 0 || a->K <= 0) {
    set_error("gemm: empty problem");
    return -1;
  }
  return 0;
}

}  // namespace tdist

using namespace tdist;

extern "C" {

int tdist_gemm(const tdist_gemm_args* a, void* stream) {
  if (check_common(a) != 0) return -1;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.mode = GEMM_PLAIN;
  p.M = a->M;
  p.N = a->N;
  p.K = a->K;
  p.c = a->c;
  p.ldc = a->ldc;
  p.is_bf16 = a->dtype == TDIST_BF16;
  p.rows_per_rank = a->M;
  p.transpose_out = a->transpose_out;
  p.block_expert = a->block_expert;
  p.ctx.world = 1;
  const int sms = a->num_sms > 0 ? a->num_sms : sm_count();
  p.group_m = a->group_m > 0 ? a->group_m : 8;
  p.num_comm_ctas = 0;
  if (a->block_expert == nullptr && !a->transpose_out && want_cta_pair(a)) {
    const int pairs = sms / 2;
    const int ptiles_m = (a->M + 2 * kBlockM - 1) / (2 * kBlockM);
    const int bn2 = pick_block_n2(ptiles_m, a->N, pairs, a->block_n);
    const int ptiles = ptiles_m * ((a->N + bn2 - 1) / bn2);
    if (make_tmap(&p.tma_a[0], a->a, a->M, a->K, a->lda, kBlockM) != 0) return -1;
    if (make_tmap(&p.tma_b, a->b, a->N, a->K, a->ldb, bn2 / 2) != 0) return -1;
    const int use_pairs = ptiles < pairs ? ptiles : pairs;
    return dispatch2(p, bn2, 2 * use_pairs, static_cast<cudaStream_t>(stream));
  }
  const int tiles_m = (a->M + kBlockM - 1) / kBlockM;
  const int bn = pick_block_n(tiles_m, a->N, sms, a->block_n);
  const int tiles = tiles_m * ((a->N + bn - 1) / bn);
  if (make_tmap(&p.tma_a[0], a->a, a->M, a->K, a->lda, kBlockM) != 0) return -1;
  const int b_rows = a->block_expert != nullptr ? a->N * (a->num_experts > 0 ? a->num_experts : 1) : a->N;
  if (make_tmap(&p.tma_b, a->b, b_rows, a->K, a->ldb, bn) != 0) return -1;
  const int grid = tiles < sms ? tiles : sms;
  return dispatch(p, bn, grid, static_cast<cudaStream_t>(stream), false);
}

// Host-side view of the single-CTA kernel's tile walk (the SAME tile_coord the kernel calls): for tile index t writes
// (shard, 128-row block inside the shard, n block). mode: 0 plain, 1 AG, 2 RS. Returns the number of tiles.
int tdist_gemm_tile_order(int mode, int M, int N, int world, int rank, int block_n, int group_m, int* out3, int cap_tiles) {
  if (M < 1 || N < 1 || world < 1 || rank < 0 || rank >= world || block_n < 8 || (mode != 0 && M % world != 0)) {
    set_error("gemm_tile_order: bad arguments");
    return -1;
  }
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.mode = mode;
  p.M = M;
  p.N = N;
  p.ctx.world = world;
  p.ctx.rank = rank;
  p.group_m = group_m > 0 ? group_m : 8;
  p.rows_per_rank = mode == 0 ? M : M / world;
  const int tiles_n = (N + block_n - 1) / block_n;
  const int tiles_m = mode == 0 ? (M + kBlockM - 1) / kBlockM : ((p.rows_per_rank + kBlockM - 1) / kBlockM) * world;
  const int tiles = tiles_m * tiles_n;
  for (int t = 0; t < tiles && t < cap_tiles; ++t) {
    const TileCoord tc = tile_coord(p, t, tiles_m, tiles_n);
    out3[3 * t] = tc.shard;
    out3[3 * t + 1] = tc.blk;
    out3[3 * t + 2] = tc.n_blk;
  }
  return tiles;
}

int tdist_gemm_fused(const tdist_gemm_args* a, const tdist_fused_args* f, void* stream) {
  if (check_common(a) != 0) return -1;
  const SymmCtx c = to_ctx(&f->ctx);
  const int W = c.world;
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.mode = f->mode;
  p.ctx = c;
  p.M = a->M;
  p.N = a->N;
  p.K = a->K;
  p.is_bf16 = a->dtype == TDIST_BF16;
  p.flags = f->flags;
  p.flags2 = f->flags2;
  p.counters = f->counters;
  p.ps = PhaseState{f->phase, f->done};
  p.group_m = a->group_m > 0 ? a->group_m : 8;
  const int sms = a->num_sms > 0 ? a->num_sms : sm_count();
  int comm = f->num_comm_ctas;
  bool comm_given = comm >= 0;
  int tiles_m;
  if (W == 1 && (f->mode == TDIST_GEMM_AG || f->mode == TDIST_GEMM_RS)) {
    // a single rank has nothing to gather or scatter: run the plain GEMM straight from / into the caller's tensors
    tdist_gemm_args g = *a;
    if (f->mode == TDIST_GEMM_RS) {
      g.c = f->out;
      g.ldc = a->N;
    }
    return tdist_gemm(&g, stream);
  }
  if (W == 1 && f->mode == TDIST_GEMM_AR) {  // nothing to reduce: the plain GEMM, written to the caller's output
    tdist_gemm_args g = *a;
    g.c = f->out;
    return tdist_gemm(&g, stream);
  }
  if (f->mode == TDIST_GEMM_AG) {
    // a->a = own shard [M/W, K]; f->ws = symmetric workspace [2][M][K]; a->M = full M
    if (a->M % W != 0) {
      set_error("ag_gemm: M must be divisible by world size");
      return -1;
    }
    p.rows_per_rank = a->M / W;
    p.a_local = a->a;
    p.a_lda = a->lda;
    p.a_ws = f->ws;
    p.c = a->c;
    p.ldc = a->ldc;
    if (W == 1) comm = 0;
    if (comm < 0) comm = env_int("TDIST_AG_PUSH_CTAS", -1);
    comm_given = comm >= 0;
    // measured (profiles/README.md, Llama-3-8B MLP shape): with one peer the broadcast has nothing to save and the
    // unicast ring wins (0.317 vs 0.349 ms); at 4 ranks they tie, at 8 the broadcast wins (0.149 vs 0.219 ms)
    const int v = pick_variant(f, c, "TDIST_AG_VARIANT", 3);
    if (v < 0) {
      set_error("ag_gemm: the NVLS variant needs a heap with a multicast mapping");
      return -1;
    }
    p.variant = v == 2 ? VARIANT_NVLS : VARIANT_LEGACY;
    const size_t par_need = size_t(a->M) * a->K * 2;
    p.ws_parity_bytes = f->ws_bytes > 0 ? f->ws_bytes / 2 : par_need;
    if (p.ws_parity_bytes < par_need) {
      set_error("ag_gemm: workspace too small (need 2 * M * K elements)");
      return -1;
    }
    const int bps_ag = (p.rows_per_rank + kBlockM - 1) / kBlockM;
    if (p.variant == VARIANT_NVLS) {
      // `slices` push CTAs split the rows of a 128-row block, which is pushed (and flagged, and consumed) in `slabs`
      // K slabs. Measured at 8 ranks (profiles/README.md): ONE group of push CTAs, so that blocks complete strictly in
      // order; 4 CTAs already saturate the NVLS ingress, 8 leave headroom.
      int slices = env_int("TDIST_AG_SLICES", 8);
      if (slices < 1) slices = 1;
      if (slices > 32) slices = 32;  // one polling lane per slice
      while (slices & (slices - 1)) slices &= slices - 1;  // power of two (divides 128)
      int slabs = env_int("TDIST_AG_KSLABS", 1);
      const int num_kb = (a->K + kBlockK - 1) / kBlockK;
      if (slabs > kAgSlots / slices) slabs = kAgSlots / slices;
      if (slabs > num_kb) slabs = num_kb;
      if (slabs < 1) slabs = 1;
      p.ag_slices = slices;
      p.ag_slab_kblocks = (num_kb + slabs - 1) / slabs;
      p.ag_kslabs = (num_kb + p.ag_slab_kblocks - 1) / p.ag_slab_kblocks;
      if (f->flags_words > 0 && f->flags_words < size_t(W) * bps_ag * kAgSlots) {
        set_error("ag_gemm: flag area too small for the NVLS protocol (need W * ceil(M/W/128) * 128 words)");
        return -1;
      }
      if (comm < 0) {
        // one group of `slices` CTAs per ~8 MiB of shard: a group moves ~100 GB/s, the push has to finish well inside
        // the GEMM (measured: 8 CTAs at 8 ranks / 8 MiB, 16 at 4 ranks / 16 MiB; more groups than needed delay the
        // first blocks, because groups work on different blocks at the same time)
        const size_t shard_bytes = size_t(p.rows_per_rank) * a->K * 2;
        int groups = int((shard_bytes + (6u << 20)) / (8u << 20));
        groups = groups < 1 ? 1 : groups > 4 ? 4 : groups;
        comm = slices * groups;
        comm_given = true;
      }
    } else if (a->lda != a->K) {
      set_error("ag_gemm: the point-to-point push needs a contiguous shard (lda == K)");
      return -1;
    } else if (f->flags_words > 0 && f->flags_words < size_t(W) * bps_ag * kAgParts) {
      set_error("ag_gemm: flag area too small (need W * ceil(M/W/128) * 4 words)");
      return -1;
    }
    if (comm < 0) {
      // as many push CTAs as the wave count of the GEMM part tolerates (they are free until they add a wave)
      const int tm = ((a->M / W + kBlockM - 1) / kBlockM) * W;
      double best = -1;
      comm = 12;
      for (int c2 = 12; c2 <= 32; c2 += 4) {
        pick_block_n(tm, a->N, sms - c2, a->block_n);
        if (best < 0 || g_last_cost <= best * 1.001) {
          best = best < 0 ? g_last_cost : (g_last_cost < best ? g_last_cost : best);
          comm = c2;
        }
      }
    }
    // opt-in, as for RS / AR below: one flag row of constant length per source rank (the size checks above still apply: the
    // call's own rows fit in such a row whenever they fit in the area)
    if (env_int("TDIST_FIXED_FLAG_LAYOUT", 0) && f->flags_words >= size_t(W)) p.flag_row = int(f->flags_words / W);
    if (make_tmap(&p.tma_a_local, a->a, p.rows_per_rank, a->K, a->lda, kBlockM) != 0) return -1;
    if (make_tmap(&p.tma_a[0], f->ws, a->M, a->K, a->K, kBlockM) != 0) return -1;
    if (make_tmap(&p.tma_a[1], static_cast<char*>(f->ws) + p.ws_parity_bytes, a->M, a->K, a->K, kBlockM) != 0) return -1;
    tiles_m = ((p.rows_per_rank + kBlockM - 1) / kBlockM) * W;
  } else if (f->mode == TDIST_GEMM_RS) {
    // a->a = [M, K/W] local; f->ws = symmetric scatter buffer [2][W][M/W][N]; f->out = [M/W, N]
    if (a->M % W != 0) {
      set_error("gemm_rs: M must be divisible by world size");
      return -1;
    }
    p.rows_per_rank = a->M / W;
    p.c = f->ws;
    p.out = f->out;
    p.ldc = a->N;
    comm = 0;
    // measured: the NVLS pull wins at 2, 4 and 8 ranks (0.341 / 0.176 / 0.130 ms against 0.367 / 0.254 / 0.230 ms)
    const int v = pick_variant(f, c, "TDIST_RS_VARIANT", 2);
    p.variant = v == 2 ? VARIANT_NVLS : VARIANT_LEGACY;  // without multicast the pull variant sums plain peer loads
    if (v < 0) p.variant = VARIANT_LEGACY;
    p.use_mc = c.mc_base != nullptr;
    p.debug = env_int("TDIST_FUSED_DEBUG", 0);
    p.rs_unit_rows = env_int("TDIST_RS_UNIT_ROWS", 16);
    if (p.rs_unit_rows != 8 && p.rs_unit_rows != 16 && p.rs_unit_rows != 32) p.rs_unit_rows = 16;
    // both layouts ([W][M/W][N] scatter slots, [M][N] partial buffer) need M * N elements per parity
    const size_t par_need = size_t(a->M) * a->N * 2;
    p.ws_parity_bytes = f->ws_bytes > 0 ? f->ws_bytes / 2 : par_need;
    if (p.ws_parity_bytes < par_need) {
      set_error("gemm_rs: workspace too small (need 2 * M * N elements)");
      return -1;
    }
    if (make_tmap(&p.tma_a[0], a->a, a->M, a->K, a->lda, kBlockM) != 0) return -1;
    tiles_m = ((p.rows_per_rank + kBlockM - 1) / kBlockM) * W;
    // opt-in: one flag row per source rank whose length is a constant of the context (see GemmParams::flag_row); the size
    // checks below still apply to the call's own layout, which fits in such a row whenever it fits in the area
    if (env_int("TDIST_FIXED_FLAG_LAYOUT", 0) && f->flags_words >= size_t(W)) p.flag_row = int(f->flags_words / W);
  } else if (f->mode == TDIST_GEMM_AR) {
    // a->a = [M, K/W]; f->ws = symmetric tile buffer [2][M][ldc]; f->out = [M, ldc]
    p.rows_per_rank = a->M;
    p.c = f->ws;
    p.out = f->out;
    p.ldc = a->ldc;
    p.ar_one_shot = f->ar_one_shot;
    p.use_mc = c.mc_base != nullptr && f->use_mc;
    p.ws_parity_bytes = f->ws_bytes > 0 ? f->ws_bytes / 2 : size_t(a->M) * a->ldc * 2;
    if (p.ws_parity_bytes < size_t(a->M) * a->ldc * 2) {
      set_error("gemm_ar: workspace too small (need 2 * M * ldc elements)");
      return -1;
    }
    // no SM is reserved: warps 2 / 3 of every GEMM CTA reduce / publish (f->num_comm_ctas is accepted and ignored)
    comm = 0;
    p.rs_unit_rows = env_int("TDIST_RS_UNIT_ROWS", 16);
    if (p.rs_unit_rows != 8 && p.rs_unit_rows != 16 && p.rs_unit_rows != 32) p.rs_unit_rows = 16;
    if (f->flags2 == nullptr) {
      set_error("gemm_ar: flags2 (>= 64 symmetric words) is required");
      return -1;
    }
    p.counters = f->flags2 + 32;  // a local word behind the W closing flags
    if (env_int("TDIST_FIXED_FLAG_LAYOUT", 0) && f->flags_words >= size_t(W)) p.flag_row = int(f->flags_words / W);
    if (make_tmap(&p.tma_a[0], a->a, a->M, a->K, a->lda, kBlockM) != 0) return -1;
    tiles_m = (a->M + kBlockM - 1) / kBlockM;
  } else {
    set_error("gemm_fused: unknown mode");
    return -1;
  }
  // Tile kind. Point-to-point AG at 8 ranks measured 0.159 ms on single-CTA tiles against 0.213 ms on pair tiles (a
  // 256-row tile waits for twice as many arriving row blocks), so that variant takes pair tiles only from 2048 rows per
  // rank; with the NVLS broadcast all shards arrive block by block and pair tiles are the default.
  const bool pair_ok = f->mode == TDIST_GEMM_RS || a->cta_group == 2 || getenv("TDIST_GEMM_CTA_GROUP") != nullptr ||
                       p.variant == VARIANT_NVLS || p.rows_per_rank >= 16 * kBlockM;
  const bool nvls_ag = f->mode == TDIST_GEMM_AG && p.variant == VARIANT_NVLS;
  auto round_push = [&](int cm) {
    // NVLS push: whole groups of `ag_slices` CTAs, so that each block is pushed by one group
    if (!nvls_ag) return cm;
    cm = cm / p.ag_slices * p.ag_slices;
    return cm < p.ag_slices ? p.ag_slices : cm;
  };
  auto check_rs_flags = [&](int bn) {
    if (f->mode == TDIST_GEMM_AR && f->flags_words > 0 &&
        f->flags_words < size_t(W) * ((a->M + kBlockM - 1) / kBlockM) * ((a->N + bn - 1) / bn)) {
      set_error("gemm_ar: flag area too small (need W * ceil(M/128) * ceil(N/BLOCK_N) words)");
      return -1;
    }
    if (f->mode != TDIST_GEMM_RS || p.variant != VARIANT_NVLS || f->flags_words == 0) return 0;
    const size_t need = size_t(W) * ((p.rows_per_rank + kBlockM - 1) / kBlockM) * ((a->N + bn - 1) / bn);
    if (f->flags_words >= need) return 0;
    set_error("gemm_rs: flag area too small for the NVLS protocol (need W * ceil(M/W/128) * ceil(N/BLOCK_N) words)");
    return -1;
  };
  if ((f->mode == TDIST_GEMM_AG || f->mode == TDIST_GEMM_RS) && want_cta_pair(a) && pair_ok && p.rows_per_rank >= 2 * kBlockM) {
    // CTA-pair kernel: 256-row tiles on cluster pairs. Push CTAs (AG) are launched in pairs as well.
    const int ptiles_m = ((p.rows_per_rank + 2 * kBlockM - 1) / (2 * kBlockM)) * W;
    if (f->mode == TDIST_GEMM_AG && W > 1) {
      if (!comm_given) {
        // largest even push-CTA count that does not add a GEMM wave
        long best_waves = -1;
        comm = 12;
        for (int c2 = 8; c2 <= 32; c2 += 4) {
          const int prs = (sms - c2) / 2;
          const int bnc = pick_block_n2(ptiles_m, a->N, prs, a->block_n);
          const long tl = long(ptiles_m) * ((a->N + bnc - 1) / bnc);
          const long cost = ((tl + prs - 1) / prs) * bnc;
          if (best_waves < 0 || cost <= best_waves) {
            best_waves = cost;
            comm = c2;
          }
        }
      }
      comm = round_push((comm + 1) / 2 * 2);
    }
    p.num_comm_ctas = comm;
    const int pairs = (sms - comm) / 2;
    if (pairs < 1) {
      set_error("gemm_fused: no SMs left for the GEMM");
      return -1;
    }
    const int bn2 = pick_block_n2(ptiles_m, a->N, pairs, a->block_n);
    if (check_rs_flags(bn2) != 0) return -1;
    if (make_tmap(&p.tma_b, a->b, a->N, a->K, a->ldb, bn2 / 2) != 0) return -1;
    const long ptiles = long(ptiles_m) * ((a->N + bn2 - 1) / bn2);
    const int use_pairs = (f->mode == TDIST_GEMM_RS) ? pairs : int(ptiles < pairs ? ptiles : pairs);
    return dispatch2(p, bn2, comm + 2 * use_pairs, static_cast<cudaStream_t>(stream));
  }
  if (f->mode == TDIST_GEMM_AG) comm = round_push(comm);
  p.num_comm_ctas = comm;
  const int gemm_ctas_max = sms - comm;
  if (gemm_ctas_max < 1) {
    set_error("gemm_fused: no SMs left for the GEMM");
    return -1;
  }
  const int bn = pick_block_n(tiles_m, a->N, gemm_ctas_max, a->block_n);
  if (check_rs_flags(bn) != 0) return -1;
  if (make_tmap(&p.tma_b, a->b, a->N, a->K, a->ldb, bn) != 0) return -1;
  const int tiles = tiles_m * ((a->N + bn - 1) / bn);
  // RS: the whole grid takes part in the reduction (tail or NVLS reducer warps), so always launch every SM
  const int gemm_ctas = (f->mode == TDIST_GEMM_RS) ? gemm_ctas_max : (tiles < gemm_ctas_max ? tiles : gemm_ctas_max);
  return dispatch(p, bn, comm + gemm_ctas, static_cast<cudaStream_t>(stream), true);
}

}  // extern "C"
