"""Command-line client of model_server.py (mirrors reference mega_triton_kernel/test/models/chat.py).
    python -m triton_dist.mega_triton_kernel.test.models.chat --port 9178 --prompt-ids 1,2,3 --max-new-tokens 8
With --tokenizer <local HF dir> plain text prompts are tokenised / detokenised."""
import argparse
import json
import socket

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--port", type=int, default=9178)
    ap.add_argument("--prompt-ids", default="")
    ap.add_argument("--prompt", default="")
    ap.add_argument("--tokenizer", default="")
    ap.add_argument("--max-new-tokens", type=int, default=16)
    ap.add_argument("--quit", action="store_true")
    args = ap.parse_args()
    tok = None
    if args.tokenizer:
        from transformers import AutoTokenizer

        tok = AutoTokenizer.from_pretrained(args.tokenizer, local_files_only=True)
    if args.quit:
        req = {"quit": True}
    else:
        ids = tok(args.prompt)["input_ids"] if tok and args.prompt else [int(x) for x in args.prompt_ids.split(",") if x]
        req = {"prompt_ids": ids, "max_new_tokens": args.max_new_tokens}
    s = socket.create_connection(("127.0.0.1", args.port))
    s.sendall((json.dumps(req) + "\n").encode())
    if not args.quit:
        resp = json.loads(s.makefile().readline())
        print(tok.decode(resp["token_ids"]) if tok else resp["token_ids"])
    s.close()
