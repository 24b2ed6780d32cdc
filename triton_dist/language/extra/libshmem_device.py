"""NVSHMEM device API names used by the reference's kernels (``language/extra/libshmem_device.py:28-342``) and their
native equivalents over the VMM symmetric heap (no proxy thread, no NIC: a put is a store to a mapped peer address)."""

NVSHMEM_CMP_EQ, NVSHMEM_CMP_NE, NVSHMEM_CMP_GT, NVSHMEM_CMP_LE, NVSHMEM_CMP_LT, NVSHMEM_CMP_GE = range(6)
NVSHMEM_SIGNAL_SET, NVSHMEM_SIGNAL_ADD = 9, 10

EQUIVALENTS = {
    "my_pe": "rank(ctx)", "n_pes": "num_ranks(ctx)", "team_my_pe": "rank(ctx)", "team_n_pes": "num_ranks(ctx)",
    "remote_ptr": "symm_at(ctx, ptr, pe)", "remote_mc_ptr": "mc_ptr(ctx, ptr)",
    "int_p": "ptx::st_relaxed_sys(symm_at(ctx, ptr, pe), value)",
    "putmem_block": "putmem_block(ctx, dst, src, bytes, pe)  (16-byte vector stores by the whole CTA)",
    "putmem_nbi_block": "putmem_block (stores are posted; completion = fence() / quiet())",
    "putmem_warp": "putmem_warp(ctx, dst, src, bytes, pe)", "putmem_nbi_warp": "putmem_warp",
    "getmem_block": "getmem_block(ctx, dst, src, bytes, pe)", "getmem_nbi_block": "getmem_block",
    "putmem_signal_block": "putmem_signal_block(ctx, dst, src, bytes, flag, value, op, pe)",
    "putmem_signal_nbi_block": "putmem_signal_block", "putmem_signal_nbi_warp": "putmem_warp + notify",
    "signal_op": "notify(ctx, flag, pe, value, op)", "signal_wait_until": "signal_wait_until(flag, cmp, value)",
    "fence": "fence()  (fence.acq_rel.sys)", "quiet": "quiet()  (fence.sc.sys)",
    "barrier_all_block": "barrier_all_block(ctx, flags, phase)", "barrier_all": "barrier_all_block (one CTA)",
    "sync_all_block": "barrier_all_block", "broadcast_block": "multimem_st_v4 loop over mc_ptr(ctx, buf), or putmem_block to every pe",
    "fcollect_block": "allgather_kernel (csrc/kernels/collectives.cu)",
    "putmem_rma_block": "n/a on NVSwitch (exists in the reference to force the NIC path on PCIe boxes)",
}
