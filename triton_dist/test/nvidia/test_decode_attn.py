"""Single-GPU GQA decode attention (paged and dense cache, split-KV + combine) against torch (mirrors reference
test/nvidia/test_decode_attn.py).
    python -m triton_dist.test.nvidia.test_decode_attn --kv-len 8192 --bsz 4"""
import torch

from triton_dist.kernels.nvidia.flash_decode import (dense_decode, gqa_fwd_batch_decode, gqa_fwd_batch_decode_intra_rank,
                                                     torch_paged_decode)
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, setup
from triton_dist.utils import dist_print

if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("--bsz", type=int, default=3)
    p.add_argument("--heads", type=int, default=32)
    p.add_argument("--kv-heads", type=int, default=8)
    p.add_argument("--kv-len", type=int, default=1500)
    p.add_argument("--page", type=int, default=16)
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    B, Hq, Hkv, D, page = args.bsz, args.heads, args.kv_heads, 128, args.page
    scale = D ** -0.5
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(77 + it + rank)
        kv_lens = torch.randint(1, args.kv_len + 1, (B,), device=dev, dtype=torch.int32)
        kv_lens[0] = args.kv_len
        pages_per_seq = (args.kv_len + page - 1) // page
        n_pages = B * pages_per_seq
        perm = torch.randperm(n_pages, device=dev).int().view(B, pages_per_seq)  # scattered pages
        kc = torch.randn(n_pages, page, Hkv, D, device=dev).to(dtype)
        vc = torch.randn(n_pages, page, Hkv, D, device=dev).to(dtype)
        q = torch.randn(B, Hq, D, device=dev).to(dtype)
        ref, ref_lse = torch_paged_decode(q, kc, vc, kv_lens, perm, scale)
        out = gqa_fwd_batch_decode(q, kc, vc, None, None, kv_lens, perm, scale)
        good = check_close(out, ref.float(), dtype if dev == "cuda" else torch.float16, f"paged decode iter {it}")
        part = gqa_fwd_batch_decode_intra_rank(q, kc, vc, None, None, kv_lens, perm, scale)
        good = check_close(part[..., :D], ref.float(), dtype if dev == "cuda" else torch.float16, f"paged decode (out|lse) iter {it}") and good
        good = bool(((part[..., D] - ref_lse.float()).abs() < 2e-2).all()) and good
        # dense cache [B, S, Hkv, D]
        kd = torch.randn(B, args.kv_len, Hkv, D, device=dev).to(dtype)
        vd = torch.randn(B, args.kv_len, Hkv, D, device=dev).to(dtype)
        bt = torch.arange(B, dtype=torch.int32, device=dev)[:, None]
        refd, _ = torch_paged_decode(q, kd, vd, kv_lens, bt, scale)
        good = check_close(dense_decode(q, kd, vd, kv_lens, scale), refd.float(), dtype if dev == "cuda" else torch.float16,
                           f"dense decode iter {it}") and good
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        kv_lens = torch.full((B,), args.kv_len, device=dev, dtype=torch.int32)
        t = bench(lambda: gqa_fwd_batch_decode(q, kc, vc, None, None, kv_lens, perm, scale), args)
        nbytes = 2.0 * B * args.kv_len * Hkv * D * 2
        dist_print(f"decode attn B={B} Hq={Hq} Hkv={Hkv} kv_len={args.kv_len}: {t * 1e3:.1f} us, KV read {nbytes / t / 1e9:.2f} TB/s")
    finish(ok)
