"""Expert-parallel MoE layer end to end: router -> EP dispatch (optionally fp8) -> local experts (SwiGLU MLP per expert)
-> EP combine with the routing weights, against a dense per-expert torch loop over the replicated experts (mirrors the
DistributedMoELayer demo of reference test/nvidia/test_ep_moe_inference.py:317).
    torchrun --nproc-per-node 8 -m triton_dist.test.nvidia.test_ep_moe_inference -M 128 --hidden 7168 --inter 2048 -G 256 --topk 8"""
import torch
import torch.nn.functional as F

from triton_dist.layers.nvidia import EPAll2AllLayer
from triton_dist.test.common import DTYPES, all_ok, base_parser, bench, check_close, finish, rand, setup, skew
from triton_dist.utils import dist_print


class DistributedMoELayer:
    """Each rank owns G / W experts. forward(x [T, H]) -> [T, H]."""

    def __init__(self, group, rank, world, max_tokens, hidden, inter, num_experts, topk, dtype, fp8=False):
        self.rank, self.world, self.topk, self.E, self.epr = rank, world, topk, num_experts, num_experts // world
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        g = torch.Generator(device=dev).manual_seed(4242)  # all experts generated everywhere, each rank keeps its slice
        self.router = (torch.randn(num_experts, hidden, device=dev, generator=g) * 0.05).to(dtype)
        self.w1_all = (torch.randn(num_experts, 2 * inter, hidden, device=dev, generator=g) * 0.03).to(dtype)
        self.w2_all = (torch.randn(num_experts, hidden, inter, device=dev, generator=g) * 0.03).to(dtype)
        self.w1 = self.w1_all[rank * self.epr:(rank + 1) * self.epr]
        self.w2 = self.w2_all[rank * self.epr:(rank + 1) * self.epr]
        self.a2a = EPAll2AllLayer(group, max_tokens, hidden, topk, rank, num_experts, world, world, dtype=dtype, fp8_dispatch=fp8)
        self.fp8, self.dtype, self.hidden = fp8, dtype, hidden
        # fp8 dispatch on a GPU: the experts run on the block-scaled fp8 GEMM and consume the dispatched e4m3 rows + 1x128
        # scales as they arrive (weights quantised once, one scale per 128 x 128 block)
        self.fp8_gemm = fp8 and dev == "cuda" and hidden % 128 == 0 and inter % 128 == 0
        if self.fp8_gemm:
            from triton_dist.kernels.nvidia.gemm_fp8 import quant_fp8_128x128

            self.w1_8, self.s1 = quant_fp8_128x128(self.w1)
            self.w2_8, self.s2 = quant_fp8_128x128(self.w2)

    def route(self, x):
        logits = F.linear(x.float(), self.router.float())
        w, ids = torch.softmax(logits, -1).topk(self.topk, dim=-1)
        return (w / w.sum(-1, keepdim=True)), ids.int()

    def forward(self, x):
        w, ids = self.route(x)
        buf, counts, n = self.a2a.dispatch(x, ids, exact=False)
        n_recv = int(n.item())  # (the reference reads the count back on the host as well, ep_a2a_layer.py:173-193)
        if self.fp8_gemm:
            from triton_dist.kernels.nvidia.gemm_fp8 import FP8, gemm_fp8, quant_fp8_1x128

            a8, sa = buf[:n_recv].view(FP8), self.a2a.dispatch_scales()[:n_recv]
            out = torch.empty((n_recv, self.hidden), dtype=self.dtype, device=x.device)
            offs = torch.cumsum(counts, 0) - counts
            for e in range(self.epr):
                c, o = int(counts[e]), int(offs[e])
                if c:
                    gu = gemm_fp8(a8[o:o + c], sa[o:o + c], self.w1_8[e], self.s1[e], out_dtype=self.dtype)
                    act = F.silu(gu[:, :gu.shape[1] // 2]) * gu[:, gu.shape[1] // 2:]
                    act8, sact = quant_fp8_1x128(act)
                    gemm_fp8(act8, sact, self.w2_8[e], self.s2[e], out=out[o:o + c])
            return self.a2a.combine(out, w)
        if self.fp8 and x.is_cuda:
            rows = buf[:n_recv].view(torch.float8_e4m3fn).float().view(n_recv, self.hidden // 128, 128)
            h = (rows * self.a2a.dispatch_scales()[:n_recv, :, None]).view(n_recv, self.hidden).to(self.dtype)
        else:
            h = buf[:n_recv].to(self.dtype)
        out = torch.empty_like(h)
        offs = torch.cumsum(counts, 0) - counts
        for e in range(self.epr):  # received rows are grouped by local expert
            c, o = int(counts[e]), int(offs[e])
            if c:
                gu = F.linear(h[o:o + c], self.w1[e])
                act = F.silu(gu[:, :gu.shape[1] // 2]) * gu[:, gu.shape[1] // 2:]
                out[o:o + c] = F.linear(act, self.w2[e])
        return self.a2a.combine(out, w)

    def torch_forward(self, x):
        w, ids = self.route(x)
        out = torch.zeros(x.shape, device=x.device)
        for k in range(self.topk):
            for e in ids[:, k].unique().tolist():
                m = ids[:, k] == e
                gu = F.linear(x[m], self.w1_all[e])
                act = F.silu(gu[:, :gu.shape[1] // 2]) * gu[:, gu.shape[1] // 2:]
                out[m] += w[m, k:k + 1] * F.linear(act, self.w2_all[e]).float()
        return out


if __name__ == "__main__":
    p = base_parser(__doc__)
    p.add_argument("-M", type=int, default=64, help="tokens per rank")
    p.add_argument("--hidden", type=int, default=1024)
    p.add_argument("--inter", type=int, default=512)
    p.add_argument("-G", type=int, default=16, help="total experts")
    p.add_argument("--topk", type=int, default=4)
    p.add_argument("--fp8", action="store_true")
    args = p.parse_args()
    group, rank, world = setup()
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    dtype = DTYPES[args.dtype] if dev == "cuda" else torch.float32
    layer = DistributedMoELayer(group, rank, world, args.M, args.hidden, args.inter, args.G, args.topk, dtype, fp8=args.fp8)
    ok = True
    for it in range(args.verify_iters):
        torch.manual_seed(600 + it * 11 + rank)
        x = rand((args.M, args.hidden), dtype)
        skew(args.stress)
        out = layer.forward(x)
        ref = layer.torch_forward(x)
        tol = (0.12 if layer.fp8_gemm else 0.08) if args.fp8 else None
        if tol is None:
            good = check_close(out, ref, dtype if dev == "cuda" else torch.float16, f"ep moe iter {it}")
        else:
            err, scale = float((out.float() - ref).abs().max()), float(ref.abs().max())
            good = err <= tol * scale
            if not good:
                dist_print(f"ep moe (fp8 dispatch) iter {it}: max err {err:.4g} vs scale {scale:.4g}", allowed_ranks="all", prefix=True)
        ok = all_ok(good)
        if not ok:
            break
    if ok and not args.check and dev == "cuda":
        x = rand((args.M, args.hidden), dtype)
        t = bench(lambda: layer.forward(x), args, flush_l2=False)
        dist_print(f"ep moe layer M={args.M}/rank H={args.hidden} I={args.inter} G={args.G} topk={args.topk} W={world} fp8={args.fp8}: {t:.3f} ms")
    layer.a2a.finalize()
    finish(ok)
