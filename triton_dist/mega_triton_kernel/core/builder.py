"""ModelBuilder: Python face of the native task-graph builder (csrc/mega/builder.cc).

Same ``make_*`` surface as the reference's ``mega_triton_kernel/models/model_builder.py:83-404``; each call appends
one op to the native graph (tiles, dependencies and the per-SM static schedule are computed in C++), ``compile()``
uploads the schedule and ``run()`` is a single cooperative kernel launch. Dependencies are linear as in the reference
(an op waits for all tiles of the previous op); ``make_prefetch`` ops carry no dependency.

B200 additions (optional keyword arguments, all default to the reference behaviour): ``rms_weight=`` folds an RMSNorm
into a linear's prologue, ``fuse_silu=True`` makes fc1 emit silu(gate)*up directly, ``residual=`` folds the residual add
into the all-reduce. ``Qwen3Model`` uses them to cut a layer from 13 dependent ops to 7.

Without a GPU the builder records the same graph and ``run()`` executes it op by op with torch (used by CPU tests)."""
from __future__ import annotations

import ctypes
from typing import Callable, List, Optional

import torch

from ... import _lib
from ...kernels.nvidia import elementwise as ew
from ...kernels.nvidia.flash_decode import torch_paged_decode
from ...runtime import symm as _symm

(OP_LINEAR, OP_RMS_NORM, OP_SILU_MUL, OP_ADD, OP_QK_NORM_ROPE_KV, OP_ATTN_SPLIT, OP_ATTN_COMBINE, OP_ALLREDUCE, OP_BARRIER,
 OP_PREFETCH, OP_EMBEDDING, OP_ARGMAX, OP_ATTN_FUSED) = range(13)
MAX_AR_TILES = 32
NUM_KV_SPLITS_DEFAULT = 16


def _p(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


class ModelBuilder:

    def __init__(self, rank: int = 0, world_size: int = 1, local_world_size: int = 1, num_sms: int = 0):
        assert world_size % local_world_size == 0 and world_size == local_world_size, "single NVSwitch node only"
        self.rank, self.world_size, self.local_world_size = rank, world_size, local_world_size
        self.native = _lib.use_native()
        self._keep: List[torch.Tensor] = []  # tensors referenced by raw pointer must outlive the graph
        self._torch_ops: List[Callable[[], None]] = []
        self._metrics = {"memory": 0}
        self.all_symm_tensors: List[torch.Tensor] = []
        self._num_cross_rank_ops = 0
        self._handle = ctypes.c_void_p()
        self._ar_flags = None
        self._compiled = False
        self._step = 0
        if self.native:
            _lib.check(_lib.lib().tdist_mega_create(rank, world_size, num_sms, ctypes.byref(self._handle)), "mega_create")

    # ------------------------------------------------------------------ helpers
    def create_symm_tensor(self, shape, dtype) -> torch.Tensor:
        t = _symm.get_heap().alloc(tuple(shape), dtype)
        self.all_symm_tensors.append(t)
        return t

    def get_memory_size(self) -> int:
        return self._metrics["memory"]

    def _count(self, *tensors):
        self._metrics["memory"] += sum(t.numel() * t.element_size() for t in tensors if t is not None)

    def _add(self, op_type: int, ins, outs, M=0, N=0, K=0, ints=(), longs=(), floats=(), dep0=-2, dep1=-1, chain=True) -> int:
        self._keep.extend(t for t in list(ins) + list(outs) if t is not None)
        if not self.native:
            return len(self._torch_ops)
        d = _lib.MegaOp()
        for k, t in enumerate(ins):
            d.inp[k] = _p(t)
        for k, t in enumerate(outs):
            d.out[k] = _p(t)
        d.M, d.N, d.K = M, N, K
        for k, v in enumerate(ints):
            d.i[k] = int(v)
        for k, v in enumerate(longs):
            d.l[k] = int(v)
        for k, v in enumerate(floats):
            d.f[k] = float(v)
        idx = _lib.lib().tdist_mega_add_op(self._handle, op_type, ctypes.byref(d), dep0, dep1, int(chain))
        if idx < 0:
            _lib.check(-1, "mega_add_op")
        return idx

    # ------------------------------------------------------------------ linears
    def _make_fc(self, input, weight, output, layer_id, rms_weight=None, rms_eps=1e-6, fuse_silu=False, bias=None,
                 parity_stride_bytes: int = 0):
        x = input.reshape(-1, input.shape[-1])
        M, K = x.shape
        N = weight.shape[0] // 2 if fuse_silu else weight.shape[0]
        y = output.reshape(-1, output.shape[-1])
        assert weight.shape[1] == K and y.shape[-1] >= N and x.is_contiguous() and weight.is_contiguous()
        self._count(x, weight, y)

        def ref():
            xin = x if rms_weight is None else ew.rmsnorm(x, rms_weight, rms_eps)
            h = torch.nn.functional.linear(xin, weight, None)
            if fuse_silu:
                h = ew.silu_mul(h)
            if bias is not None:
                h = h + bias
            y[:, :N].copy_(h)

        self._torch_ops.append(ref)
        return self._add(OP_LINEAR, [x, weight, rms_weight, bias], [y], M=M, N=N, K=K,
                         ints=[int(fuse_silu), y.stride(0), parity_stride_bytes], floats=[rms_eps])

    def make_fc1(self, input, weight, output, layer_id: int = 0, **kw):
        return self._make_fc(input, weight, output, layer_id, **kw)

    def make_fc2(self, input, weight, output, layer_id: int = 0, **kw):
        return self._make_fc(input, weight, output, layer_id, **kw)

    def make_qkv_proj(self, input, weight, output, layer_id: int = 0, **kw):
        return self._make_fc(input, weight, output, layer_id, **kw)

    def make_o_proj(self, input, weight, output, layer_id: int = 0, **kw):
        return self._make_fc(input, weight, output, layer_id, **kw)

    def make_linear(self, input, weight, output, layer_id: int = 0, **kw):
        return self._make_fc(input, weight, output, layer_id, **kw)

    # ------------------------------------------------------------------ attention
    def make_attn(self, query, key_cache, value_cache, block_tables, kv_lens, output, sm_scale=None, soft_cap=0.0,
                  layer_id: int = 0, num_kv_splits: int = NUM_KV_SPLITS_DEFAULT):
        """query/output: [B, 1, Hq, D]; caches [pages, page_size, Hkv, D]; block_tables [B, max_pages] int32;
        kv_lens [B] int32 holds the LAST position (kv length - 1), i.e. the position written by this step."""
        B, q_len, Hq, D = query.shape
        assert q_len == 1 and D == 128 and soft_cap == 0.0
        Hkv, page = key_cache.shape[2], key_cache.shape[1]
        scale = D ** -0.5 if sm_scale is None else sm_scale
        partial = torch.empty((B, Hq, num_kv_splits, D + 1), dtype=torch.float32, device=query.device)
        self._count(query, output)

        def ref():
            o, _ = torch_paged_decode(query.view(B, Hq, D), key_cache, value_cache, kv_lens + 1, block_tables, scale)
            output.view(B, Hq, D).copy_(o.to(output.dtype))

        self._torch_ops.append(ref)
        self._torch_ops.append(lambda: None)
        ints = [Hq, Hkv, page, block_tables.shape[1], num_kv_splits]
        strides = [key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)]
        self._add(OP_ATTN_SPLIT, [query, key_cache, value_cache, block_tables, kv_lens], [partial], M=B, ints=ints, longs=strides,
                  floats=[scale])
        return self._add(OP_ATTN_COMBINE, [partial], [output], M=B, ints=ints)

    @staticmethod
    def supports_fused_attn(num_q_heads: int, num_kv_heads: int) -> bool:
        return num_kv_heads > 0 and num_q_heads % num_kv_heads == 0 and 1 <= num_q_heads // num_kv_heads <= 8

    def make_attn_fused(self, qkv, key_cache, value_cache, block_tables, kv_lens, q_rms_weight, k_rms_weight, cos_sin_cache,
                        output, q_rms_eps: float = 1e-6, sm_scale=None, layer_id: int = 0,
                        num_kv_splits: int = NUM_KV_SPLITS_DEFAULT):
        """B200 addition: ``make_qk_norm_rope_update_kvcache`` + ``make_attn`` as ONE task type. Every split task
        normalises / rotates its own query heads from ``qkv`` [B, 1, Hq + 2*Hkv, D], the split that owns the newest
        position writes K/V into the paged cache, and the last split to finish does the LSE combine into ``output``
        [B, 1, Hq, D]: two scoreboard hops and two small ops fewer per layer."""
        B = qkv.shape[0]
        Hq, D = output.shape[2], output.shape[3]
        Hkv, page = key_cache.shape[2], key_cache.shape[1]
        assert D == 128 and self.supports_fused_attn(Hq, Hkv)
        scale = D ** -0.5 if sm_scale is None else sm_scale
        # partial results [B, Hq, S, D + 1] fp32 followed by one arrival counter per (batch, kv head)
        blob = torch.zeros(B * Hq * num_kv_splits * (D + 1) + B * Hkv, dtype=torch.float32, device=qkv.device)
        norm_w = None
        if q_rms_weight is not None:
            norm_w = torch.stack([q_rms_weight, k_rms_weight]).to(qkv.dtype).contiguous()
        self._count(qkv, output)
        q_tmp = torch.empty((B, Hq, D), dtype=qkv.dtype, device=qkv.device)

        def ref():
            ew.qk_norm_rope_update_kvcache(qkv.reshape(B, -1), q_rms_weight, k_rms_weight, cos_sin_cache, kv_lens, key_cache,
                                           value_cache, Hq, Hkv, D, q_rms_eps,
                                           batch_idx=torch.arange(B, dtype=torch.int32, device=qkv.device),
                                           block_table=block_tables, q_out=q_tmp)
            o, _ = torch_paged_decode(q_tmp, key_cache, value_cache, kv_lens + 1, block_tables, scale)
            output.view(B, Hq, D).copy_(o.to(output.dtype))

        self._torch_ops.append(ref)
        ints = [Hq, Hkv, page, block_tables.shape[1], num_kv_splits]
        strides = [key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)]
        return self._add(OP_ATTN_FUSED, [qkv, key_cache, value_cache, block_tables, kv_lens, cos_sin_cache], [blob, output, norm_w],
                         M=B, ints=ints, longs=strides, floats=[scale, q_rms_eps])

    def make_qk_norm_rope_update_kvcache(self, qkv, key_cache, value_cache, block_tables, kv_lens, q_rms_weight, k_rms_weight,
                                         cos_sin_cache, q_norm_rope, q_rms_eps: float = 1e-6, k_rms_eps: float = 1e-6,
                                         rope_theta: int = 1000000, layer_id: int = 0):
        """qkv [B, 1, Hq + 2*Hkv, D] -> q_norm_rope [B, 1, Hq, D]; K/V of the new token are written into the paged cache at
        position kv_lens[b]. ``cos_sin_cache``: fp32 [max_pos, D] (cos | sin)."""
        B = qkv.shape[0]
        Hq = q_norm_rope.shape[2]
        Hkv = key_cache.shape[2]
        D = qkv.shape[-1]
        self._count(qkv, q_norm_rope)

        def ref():
            ew.qk_norm_rope_update_kvcache(qkv.reshape(B, -1), q_rms_weight, k_rms_weight, cos_sin_cache, kv_lens, key_cache,
                                           value_cache, Hq, Hkv, D, q_rms_eps,
                                           batch_idx=torch.arange(B, dtype=torch.int32, device=qkv.device),
                                           block_table=block_tables, q_out=q_norm_rope.view(B, Hq, D))

        self._torch_ops.append(ref)
        return self._add(OP_QK_NORM_ROPE_KV, [qkv, key_cache, value_cache, q_rms_weight, k_rms_weight, cos_sin_cache],
                         [q_norm_rope, kv_lens, block_tables], M=B, ints=[Hq, Hkv, key_cache.shape[1], block_tables.shape[1]],
                         longs=[key_cache.stride(0), key_cache.stride(1), key_cache.stride(2)], floats=[q_rms_eps])

    # ------------------------------------------------------------------ elementwise
    def make_silu_mul_up(self, fc1_out, act_out, layer_id: int = 0):
        x = fc1_out.reshape(-1, fc1_out.shape[-1])
        y = act_out.reshape(-1, act_out.shape[-1])
        self._count(x, y)
        self._torch_ops.append(lambda: y.copy_(ew.silu_mul(x)))
        return self._add(OP_SILU_MUL, [x], [y], M=x.shape[0], N=y.shape[1])

    def make_rms_norm(self, input, rms_weight, output, rms_eps: float = 1e-6, layer_id: int = 0, residual=None):
        x = input.reshape(-1, input.shape[-1])
        y = output.reshape(-1, input.shape[-1])
        self._count(x, y)
        self._torch_ops.append(lambda: y.copy_(ew.rmsnorm(x, rms_weight, rms_eps, residual=residual)))
        return self._add(OP_RMS_NORM, [x, None, rms_weight, residual], [y], M=x.shape[0], K=x.shape[1], floats=[rms_eps])

    def make_add(self, lhs, rhs, output, layer_id: int = 0):
        self._count(lhs, rhs, output)
        self._torch_ops.append(lambda: torch.add(lhs, rhs, out=output))
        return self._add(OP_ADD, [lhs, rhs], [output], N=lhs.numel())

    def make_embedding(self, ids, table, output, layer_id: int = 0):
        self._torch_ops.append(lambda: output.view(-1, table.shape[1]).copy_(table[ids.view(-1)]))
        return self._add(OP_EMBEDDING, [ids, table], [output], M=ids.numel(), K=table.shape[1])

    def make_argmax(self, logits, output, layer_id: int = 0):
        l2 = logits.reshape(-1, logits.shape[-1])
        self._torch_ops.append(lambda: output.view(-1).copy_(l2.float().argmax(-1)))
        return self._add(OP_ARGMAX, [l2], [output], M=l2.shape[0], N=l2.shape[1], ints=[0, l2.stride(0)])

    # ------------------------------------------------------------------ cross-rank
    def make_barrier_all_intra_node(self, layer_id: int = 0):
        self._num_cross_rank_ops += 1
        self._torch_ops.append(lambda: None)
        return self._add(OP_BARRIER, [], [])

    def make_allreduce(self, input, output, double_input_buffer: bool = False, layer_id: int = 0, residual=None,
                       parity_stride_bytes: int = 0):
        """Sum the symmetric ``input`` over ranks into ``output`` (and/or add it into ``residual`` in place).
        ``parity_stride_bytes`` > 0: ``input`` is the first of two buffers alternated per step by the producer
        (the reference's double_input_buffer contract)."""
        assert (input.dtype == torch.bfloat16 or not self.native) and input.numel() % 8 == 0
        if self.native and self.world_size > 1:
            assert _symm.get_heap().is_symmetric(input), "all-reduce input must be a symmetric tensor"
        self._num_cross_rank_ops += 1
        self._count(input, output)

        def ref():
            import torch.distributed as dist

            s = input.clone()
            if self.world_size > 1:
                dist.all_reduce(s)
            if residual is not None:
                residual.add_(s.view(residual.shape))
                if output is not None:
                    output.copy_(residual.view(output.shape))
            elif output is not None:
                output.copy_(s.view(output.shape))

        self._torch_ops.append(ref)
        return self._add(OP_ALLREDUCE, [input, None, None, residual], [output], N=input.numel(), ints=[0, 0, parity_stride_bytes])

    def make_prefetch(self, weight, layer_id: int = 0):
        self._torch_ops.append(lambda: None)
        return self._add(OP_PREFETCH, [None, weight], [], N=weight.shape[0], K=weight.shape[1], dep0=-1, chain=False)

    # ------------------------------------------------------------------ compile / run
    def compile(self):
        if self.native:
            heap = _symm.get_heap() if _symm.heap_initialized() else None
            ctx_ptr = None
            flags_ptr = None
            if heap is not None and isinstance(heap, _symm.SymmetricHeap):
                n = max(self._num_cross_rank_ops, 1) * MAX_AR_TILES * _lib.MAX_RANKS
                self._ar_flags = heap.alloc((n,), torch.int32)
                heap.barrier_all()
                torch.cuda.synchronize()
                ctx_ptr = ctypes.byref(heap.ctx)
                flags_ptr = self._ar_flags.data_ptr()
            else:
                assert self.world_size == 1, "multi-rank megakernel needs initialize_distributed()"
            _lib.check(_lib.lib().tdist_mega_compile(self._handle, ctx_ptr, flags_ptr), "mega_compile")
        self._compiled = True

    def info(self) -> dict:
        if not self.native:
            return {"ops": len(self._torch_ops), "tasks": len(self._torch_ops), "max_queue": len(self._torch_ops), "smem": 0,
                    "ctas": 1, "cross_rank_ops": self._num_cross_rank_ops}
        arr = (ctypes.c_int * 6)()
        _lib.lib().tdist_mega_info(self._handle, ctypes.byref(arr))
        return dict(zip(["ops", "tasks", "max_queue", "smem", "ctas", "cross_rank_ops"], list(arr)))

    def run(self, stream=None):
        assert self._compiled, "compile() first"
        self._step += 1
        if not self.native:
            for f in self._torch_ops:
                f()
            return
        _lib.check(_lib.lib().tdist_mega_run(self._handle, _lib.stream_ptr(stream)), "mega_run")

    # ------------------------------------------------------------------ per-task timeline
    OP_NAMES = ["linear", "rms_norm", "silu_mul", "add", "qk_norm_rope_kv", "attn_split", "attn_combine", "allreduce", "barrier",
                "prefetch", "embedding", "argmax", "attn_fused"]

    def enable_profiling(self, on: bool = True):
        """Stamp %globaltimer at the start and end of every task of the following launches (thread 0 of each CTA).

        The reference profiles its megakernel from outside with torch.profiler (test/models/bench_qwen3.py:34); a
        persistent kernel is one opaque launch there, so this build records the task timeline inside the kernel."""
        assert self._compiled, "compile() first"
        if self.native:
            _lib.check(_lib.lib().tdist_mega_set_trace(self._handle, 1 if on else 0), "mega_set_trace")

    def get_trace(self):
        """Returns a list of (cta, index, op, type_name, tile, begin_ns, end_ns, phases) of the last profiled launch;
        ``phases`` = (deps_satisfied_ns, staged_ns, consumed_ns) for linear tasks (zeros otherwise)."""
        import numpy as np
        info = self.info()
        ctas, mq = info["ctas"], info["max_queue"]
        buf = np.zeros(ctas * mq * 6 + ctas * 16 * 4, dtype=np.uint64)
        torch.cuda.synchronize()
        _lib.check(_lib.lib().tdist_mega_get_trace(self._handle, buf.ctypes.data, buf.size), "mega_get_trace")
        # per-warp cycle totals of the linear main loop: [cta][warp][wait stage, ldmatrix+mma, issue next, block barrier]
        self.last_warp_cycles = buf[ctas * mq * 6:].reshape(ctas, 16, 4).copy()
        buf = buf[:ctas * mq * 6].reshape(ctas, mq, 6)
        out = []
        q = (ctypes.c_int * (8 * mq))()
        for c in range(ctas):
            n = _lib.lib().tdist_mega_get_queue(self._handle, c, q, mq)
            for i in range(max(n, 0)):
                op, typ, tile = q[8 * i], q[8 * i + 1], q[8 * i + 2]
                out.append((c, i, op, self.OP_NAMES[typ], tile, int(buf[c, i, 0]), int(buf[c, i, 1]),
                            (int(buf[c, i, 2]), int(buf[c, i, 3]), int(buf[c, i, 4]))))
        return out

    def dump_trace(self, path: str):
        """Write the last profiled launch as a chrome://tracing JSON (one row per CTA)."""
        import json
        ev = self.get_trace()
        t0 = min((e[5] for e in ev if e[5] > 0), default=0)
        js = [{"name": f"{e[3]}#{e[2]}", "ph": "X", "pid": self.rank, "tid": e[0], "ts": (e[5] - t0) / 1e3,
               "dur": max(e[6] - e[5], 0) / 1e3, "args": {"tile": e[4], "queue_index": e[1]}} for e in ev if e[5] > 0]
        with open(path, "w") as f:
            json.dump({"traceEvents": js, "displayTimeUnit": "ns"}, f)
        return path

    def profile_summary(self) -> dict:
        """Per op type: summed busy time over CTAs (us) and count, plus the span of the whole launch."""
        ev = [e for e in self.get_trace() if e[5] > 0]
        if not ev:
            return {}
        span = (max(e[6] for e in ev) - min(e[5] for e in ev)) / 1e3
        agg: dict = {}
        for e in ev:
            d = agg.setdefault(e[3], [0.0, 0])
            d[0] += (e[6] - e[5]) / 1e3
            d[1] += 1
        ctas = self.info()["ctas"]
        return {"span_us": span, "per_type": {k: {"busy_us_per_cta": v[0] / ctas, "tasks": v[1]} for k, v in agg.items()}}

    def finalize(self):
        if self.native and self._handle:
            _lib.lib().tdist_mega_destroy(self._handle)
            self._handle = ctypes.c_void_p()
        if _symm.heap_initialized():
            heap = _symm.get_heap()
            if self._ar_flags is not None:
                heap.free(self._ar_flags)
            for t in self.all_symm_tensors:
                heap.free(t)
        self._keep.clear()
