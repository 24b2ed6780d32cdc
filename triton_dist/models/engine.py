"""Engine: prefill + CUDA-graph decode loop (reference ``models/engine.py:37-189``).

``serve`` prefillls with the torch path, switches the model to the selected backend, captures ONE decode step in a
CUDA graph (all native kernels and their cross-GPU flag protocols are graph-capturable: no host barriers, no
stream forks) and replays it per generated token."""
from __future__ import annotations

import time
from typing import Optional

import torch
import torch.distributed as dist

from ..kernels.allreduce import AllReduceMethod
from .config import ModelConfig
from .kv_cache import KV_Cache
from .utils import logger, sample_token


class Engine:

    def __init__(self, model_config: ModelConfig, temperature: float = 0.0, top_p: float = 0.95, verbose: bool = False, group=None,
                 model=None, tokenizer=None):
        self.logger = logger
        self.model_config, self.group = model_config, group
        self.temperature, self.top_p, self.verbose = temperature, top_p, verbose
        self.kv_cache: Optional[KV_Cache] = None
        self.no_graph = False
        self.backend = "torch"
        self.enable_profile = False
        self.tokenizer = tokenizer
        if model is not None:
            self.model = model
        else:
            self._init_model()
        self.last_decode_ms_per_token = None

    def _init_model(self):
        from . import AutoLLM, AutoTokenizer

        self.logger.log(f"Initializing model {self.model_config}...", "info")
        self.model = AutoLLM.from_pretrained(self.model_config, self.group)
        if self.tokenizer is None:
            try:
                self.tokenizer = AutoTokenizer.from_pretrained(self.model_config)
            except Exception as e:  # offline without a cached tokenizer: serving token ids still works
                self.logger.log(f"tokenizer unavailable ({type(e).__name__}); serving token ids only", "warning")

    def _init_kv_cache(self, bsz: int):
        self.kv_cache = KV_Cache(num_layers=self.model.num_layers, kv_heads=self.model.num_key_value_heads,
                                 head_dim=self.model.head_dim, batch_size=bsz, dtype=self.model.dtype,
                                 max_length=self.model.max_length, world_size=self.model.world_size)

    def _device(self):
        return "cuda" if torch.cuda.is_available() else "cpu"

    def _init_cuda_graph(self, bsz: int = 1):
        dev = self._device()
        in_bsz = bsz // self.model.world_size if self.backend == "triton_dist" else bsz
        static_input_ids = torch.full((in_bsz, 1), 1225, dtype=torch.long, device=dev)
        static_position_ids = torch.zeros((bsz, 1), dtype=torch.long, device=dev)
        static_position_ids.copy_(self.get_ctx(static_position_ids))
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                logits = self.model.inference(input_ids=static_input_ids, position_ids=static_position_ids, kv_cache=self.kv_cache)
            s.synchronize()
        torch.cuda.current_stream().wait_stream(s)
        if dist.is_initialized():
            dist.barrier()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            logits = self.model.inference(input_ids=static_input_ids, position_ids=static_position_ids, kv_cache=self.kv_cache)

        def run(input_ids, position_ids):
            static_input_ids.copy_(input_ids)
            static_position_ids.copy_(position_ids)
            graph.replay()
            return logits

        self._graph = graph
        return run

    def get_ctx(self, input_ids: torch.Tensor) -> torch.Tensor:
        input_len = input_ids.size(1)
        past_len = self.kv_cache.get_kv_len()
        return past_len[:, None].long() + torch.arange(input_len, device=past_len.device).long()[None, :]

    def set_backend(self, backend: str, bsz: int):
        self.backend = backend
        if backend == "triton_dist":
            self.model.set_fwd(mode="triton_dist")
            self.model.init_triton_dist_ctx(max_M=max(bsz, self.model.world_size))
        elif backend == "triton_dist_AR":
            self.model.set_fwd(mode="triton_dist_AR")
            self.model.init_triton_dist_AR_ctx(max_M=bsz, ar_method=AllReduceMethod.Unknown)
        elif backend == "triton_dist_gemm_ar":
            self.model.set_fwd(mode="triton_dist_gemm_ar")
            self.model.init_triton_dist_gemm_ar_ctx(max_M=bsz)
        elif backend == "torch":
            self.model.set_fwd(mode="torch")
        else:
            raise ValueError(f"unknown backend {backend}")

    @torch.inference_mode()
    def serve(self, input_ids: torch.Tensor, gen_len: int) -> torch.Tensor:
        dev = self._device()
        bsz = input_ids.shape[0]
        self.kv_cache = None
        self._init_kv_cache(bsz=bsz)
        self.kv_cache.clear()
        # prefill always runs the torch path (as in the reference, engine.py:119-124)
        backend = self.backend
        self.model.set_fwd(mode="torch")
        input_ids = input_ids.to(dev)
        logits = self.model.inference(input_ids=input_ids, position_ids=self.get_ctx(input_ids), kv_cache=self.kv_cache)
        next_token = sample_token(logits[:, -1, :], temperature=self.temperature, top_p=self.top_p)
        self.kv_cache.kv_offset.fill_(input_ids.shape[-1])
        self.set_backend(backend, bsz)
        if self.backend == "triton_dist":
            next_token = next_token.split(bsz // self.model.world_size, dim=0)[self.model.rank]
        if self.no_graph or not torch.cuda.is_available():
            def run(ids, pos):
                return self.model.inference(input_ids=ids, position_ids=pos, kv_cache=self.kv_cache)
            self.model_launch = run
        else:
            self.model_launch = self._init_cuda_graph(bsz)
        output_ids = []
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier()
        prof = None
        if self.enable_profile:  # profile (at most) the first 64 decode steps, merged over ranks (reference engine.py:153-179)
            from ..utils import group_profile

            prof = group_profile(f"engine_{self.backend}", do_prof=True, group=self.group)
            prof.__enter__()
        t0 = time.perf_counter()
        for step_i in range(gen_len):
            if prof is not None and step_i == 64:
                prof.__exit__(None, None, None)
                prof = None
            position_ids = self.get_ctx(torch.zeros((bsz, 1), dtype=torch.long, device=dev))
            logits = self.model_launch(next_token, position_ids)
            full_logits = logits[:, -1, :]
            next_token = sample_token(full_logits, temperature=self.temperature, top_p=self.top_p)
            self.kv_cache.inc_offset()
            output_ids.append(next_token)
        if prof is not None:
            prof.__exit__(None, None, None)
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        if dist.is_initialized():
            dist.barrier()
        total = time.perf_counter() - t0
        self.last_decode_ms_per_token = total * 1e3 / max(gen_len, 1)
        self.logger.log(f"Decoding finished: {gen_len} tokens, {self.last_decode_ms_per_token:.3f} ms/token (host clock)")
        out = torch.cat(output_ids, dim=1).cpu()
        if self.verbose and self.tokenizer is not None:
            print(self.tokenizer.batch_decode(out, skip_special_tokens=True))
        self.model_launch = None
        return out
