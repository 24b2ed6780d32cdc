"""Reference primitive name -> native device function (header, signature, semantics)."""
from __future__ import annotations

PRIMITIVES = {
    # ---- the six core primitives (docs/primitives.md:5-17)
    "rank": ("csrc/device/symm.cuh", "int rank(const SymmCtx&)", "this process' rank (kernel parameter, no call into a runtime)"),
    "num_ranks": ("csrc/device/symm.cuh", "int num_ranks(const SymmCtx&)", "world size"),
    "symm_at": ("csrc/device/symm.cuh", "T* symm_at(const SymmCtx&, T* local, int peer)",
                "address of the same symmetric object on `peer`: base[peer] + (local - base[rank])"),
    "wait": ("csrc/device/symm.cuh", "wait_eq_sys / wait_ge_sys / wait_warp(flags, n, value)",
             "spin with ld.acquire.sys until the flag matches; wait_warp = first n lanes poll n flags then __syncwarp "
             "(the reference's lowering, DistributedOpToLLVM.cpp:179-206)"),
    "consume_token": ("-", "-", "not needed: the acquire load in wait() already orders the data loads that follow it"),
    "notify": ("csrc/device/symm.cuh", "notify(const SymmCtx&, uint32_t* flag, int peer, uint32_t value, SignalOp)",
               "st.release.sys (SET) or red.release.sys.add (ADD) on the peer's copy of the flag; call from one thread "
               "after a CTA barrier so that the CTA's payload stores are ordered before it"),
    # ---- SIMT dialect: CUDA is already SIMT
    "simt_exec_region": ("-", "-", "not needed (plain CUDA threads)"),
    "extract": ("-", "-", "not needed"),
    "insert": ("-", "-", "not needed"),
    # ---- PTX helpers (backends/nvidia/language/cuda/language_extra.py:108-978)
    "ld_acquire": ("csrc/device/ptx.cuh", "ld_acquire_sys / ld_acquire_gpu", ""),
    "st": ("csrc/device/ptx.cuh", "st_release_sys / st_release_gpu / st_relaxed_sys", ""),
    "atomic_add": ("csrc/device/ptx.cuh", "atom_add_release_sys / atom_add_acq_rel_gpu / red_add_release_sys", ""),
    "atomic_cas": ("csrc/device/ptx.cuh", "atom_cas_acq_rel_sys", ""),
    "__fence": ("csrc/device/ptx.cuh", "fence_acq_rel_gpu / fence_acq_rel_sys / fence_sc_sys / fence_proxy_async", ""),
    "multimem_ld_reduce_v4": ("csrc/device/ptx.cuh", "multimem_ld_reduce_{bf16x8,f16x8,f32x4}", "NVLS in-switch reduction, fp32 accumulate"),
    "multimem_st_v4": ("csrc/device/ptx.cuh", "multimem_st_v4 / multimem_st_b32 / multimem_red_add_u32", "NVLS broadcast store / reduction"),
    "load_v4_u32": ("csrc/device/ptx.cuh", "ld_v4 / ld_nc_v4 / ld_volatile_v4 / ld_relaxed_sys_v4", "16-byte loads"),
    "st_v4_u32": ("csrc/device/ptx.cuh", "st_v4 / st_na_v4 / st_relaxed_sys_v4", "16-byte stores"),
    "tid": ("CUDA", "threadIdx.x", ""), "ntid": ("CUDA", "blockDim.x", ""), "laneid": ("csrc/device/ptx.cuh", "laneid()", ""),
    "__syncthreads": ("CUDA", "__syncthreads()", ""),
    "barrier_on_this_grid": ("csrc/device/symm.cuh", "grid_barrier(uint32_t* counter)", "sign-bit grid barrier, all CTAs co-resident"),
    "barrier_all_intra_node": ("csrc/device/symm.cuh", "barrier_all_block(const SymmCtx&, uint32_t* flags, uint32_t phase)",
                               "monotone-phase cross-GPU barrier executed by one CTA"),
}


def describe(name: str) -> str:
    header, sig, note = PRIMITIVES[name]
    return f"{name}: {sig}  [{header}]  {note}".rstrip()
