"""Minimal generation server around the megakernel model (mirrors reference
mega_triton_kernel/test/models/model_server.py + chat.py): rank 0 listens on a TCP port for newline-delimited JSON
requests {"prompt_ids": [...], "max_new_tokens": n}, broadcasts them to all ranks, runs greedy decode with one
megakernel launch per token and streams back {"token_ids": [...]}.
    torchrun --nproc-per-node 8 -m triton_dist.mega_triton_kernel.test.models.model_server --model Qwen/Qwen3-32B --port 9178"""
import argparse
import json
import socket

import torch
import torch.distributed as dist

from triton_dist.mega_triton_kernel import ModelBuilder, Qwen3Model
from triton_dist.models import ModelConfig
from triton_dist.test.common import setup
from triton_dist.utils import finalize_distributed


def generate(mega: Qwen3Model, prompt_ids, max_new_tokens: int):
    dev = mega.input_ids.device
    mega.set_kv_len(0)
    out = []
    tok = None
    for t in prompt_ids:  # the megakernel is decode-only: the prompt is consumed token by token
        mega.mega_forwrad(torch.tensor([t], device=dev))
        tok = int(mega.next_ids[0])
    for _ in range(max_new_tokens):
        out.append(tok)
        mega.mega_forwrad(torch.tensor([tok], device=dev))
        tok = int(mega.next_ids[0])
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="Qwen/Qwen3-32B")
    ap.add_argument("--port", type=int, default=9178)
    ap.add_argument("--max-length", type=int, default=4096)
    ap.add_argument("--max-requests", type=int, default=0, help="exit after this many requests (0 = serve forever)")
    args = ap.parse_args()
    group, rank, world = setup()
    cfg = ModelConfig(model_name=args.model, max_length=args.max_length, dtype=torch.bfloat16, rank=rank, world_size=world)
    mega = Qwen3Model(1, cfg, ModelBuilder(rank, world, world), group=group)
    srv = None
    if rank == 0:
        srv = socket.socket()
        srv.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        srv.bind(("127.0.0.1", args.port))
        srv.listen(1)
        print(f"model_server listening on 127.0.0.1:{args.port}", flush=True)
    served = 0
    while True:
        req = [None]
        conn = None
        if rank == 0:
            conn, _ = srv.accept()
            line = conn.makefile().readline()
            req = [json.loads(line) if line.strip() else {"quit": True}]
        dist.broadcast_object_list(req, src=0)  # prompt broadcast, as the reference does (model_server.py:153)
        r = req[0]
        if r.get("quit"):
            if conn:
                conn.close()
            break
        toks = generate(mega, r["prompt_ids"], int(r.get("max_new_tokens", 16)))
        if rank == 0:
            conn.sendall((json.dumps({"token_ids": toks}) + "\n").encode())
            conn.close()
        served += 1
        if args.max_requests and served >= args.max_requests:
            break
    mega.finalize()
    finalize_distributed()
