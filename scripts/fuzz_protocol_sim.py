#!/usr/bin/env python
"""Random configurations of the CPU protocol simulator (csrc/sim) for a given time: ops x ranks x sizes x CTA counts x calls x
alternating sizes x stragglers, each run under ThreadSanitizer. Any data race, wait bound, mismatch or crash is printed with
its command line.    python scripts/fuzz_protocol_sim.py [--seconds 300] [--seed 1]
(tests/test_cpu_protocol_sim.py holds the fixed regression cases; this script is the search.)"""
import argparse
import os
import random
import subprocess
import sys
import tempfile
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tests"))
import test_cpu_protocol_sim as T  # noqa: E402  (staging + build helpers)

OPS = ["ar1", "ar2", "ar1mc", "ar2mc", "ar1mc_inplace", "ar2mc_inplace", "ag_pull", "ag_push", "ag_mc", "ag_ll", "ag_llmc", "rs", "rsmc",
       "barrier", "selftest"]


def random_case(rng):
    W = rng.choice([1, 2, 2, 3, 4, 5, 8])
    kind = rng.choice(["coll"] * 6 + ["moe_rs", "a2a", "ep", "queue", "ag_gemm", "gemm_rs", "fused", "decode", "mega", "mega"])
    c = rng.choice
    if kind == "decode":  # one simulated GPU
        hkv = c([1, 2, 3])
        page = c([0, 0, 4, 16])
        return ["decode", c([1, 2, 5, 16]), hkv * c([1, 2, 3, 4, 7, 8, 12]), hkv, c([1, 9, 40, 90]), c([1, 2, 5, 11]), page,
                c([64, 128, 200]), c([8, 40, 64]), c([3, 50, 333]), c([0, 1])], 1
    if kind == "mega":
        W = c([1, 2, 2, 3, 4])
        graph = c([0, 0, 1, 2])
        return ["mega", W, c([1, 2, 3, 5, 9]), c([1, 1, 2]), c([1, 2, 3, 8]), c([1, 2, 3]), graph, c([0, 1]), c([1, 2]),
                c([1, 2, 4, 8] if graph == 2 else [1, 2, 4, 8, 9]), c([64, 128, 192]), c([1, 2, 3, 7, 16])], W
    if kind == "fused" and W > 1 and rng.random() < 0.3:
        pair = rng.random() < 0.5  # the CTA-pair kernel: even CTA counts
        push = c([2, 4, 8]) if pair else c([1, 2, 3, 8])
        return ["fused", "ag2" if pair else "ag", W, c([64, 128, 200, 256, 384]), c([64, 128, 192]), c([64, 128, 320, 512]), push,
                push + (c([2, 4]) if pair else c([1, 2, 4])),
                c([1, 2, 4]), c([0, 0, 1, 2]), c([1, 2, 4, 8]), c([1, 2, 4, 8])], W
    if kind == "fused":
        mode = c(["rs", "rsl", "rs2", "ar1", "ar2"])
        M = W * c([128, 256, 384]) if mode != "rs2" else W * c([256, 384, 512])
        ctas = c([2, 4, 6]) if mode == "rs2" else c([1, 2, 3, 5])  # the CTA-pair form needs whole pairs
        return ["fused", mode, W, M, c([64, 128, 192]), ctas, c([1, 2, 4]), c([0, 1]), c([0, 0, 1, 2]), 1], W
    if kind == "coll":
        return ["coll", c(OPS), W, c([1, 2, 3, 7, 16, 33, 96, 200]), c([1, 2, 3, 5]), c([1, 2, 3, 5]), c([0, 0, 1, 2])], W
    if kind == "moe_rs":
        return ["moe_rs", W, W * c([1, 2, 4, 6]) * 2, c([1, 2, 3]), c([1, 2, 5]), c([1, 2, 4]), c([2, 3, 5]), c([0, 1, 2])], W
    if kind == "a2a":
        return ["a2a", W, c([4, 9, 24]), c([1, 3]), c([1, 2, 3]), c([1, 2]), c([2, 4])], W
    if kind == "ep":
        return ["ep", W, c([2, 5, 8]), c([256, 512]), W * c([1, 2, 3]), c([1, 2, 3]), c([1, 2, 3]), c([0, 0, 1])], W
    if kind == "queue":
        return ["queue", W, c([1, 3, 10, 25]), c([1, 16, 100]), c([1, 2, 5]), c([1, 3, 5])], W
    if kind == "ag_gemm":
        ctas = c([2, 3, 5])
        return ["ag_gemm", W, c([8, 64, 100, 192]), c([8, 40, 72]), c([8, 16, 40]), rng.randrange(1, ctas), ctas, c([1, 3, 5]), c([0, 1, 2]), c([2, 3])], W
    return ["gemm_rs", W, W * c([8, 64, 100, 192]), c([8, 40, 72]), c([8, 16, 40]), c([1, 2, 4]), c([1, 3, 5]), c([0, 1, 2]), 2], W


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=300)
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()
    rng = random.Random(args.seed)
    with tempfile.TemporaryDirectory() as d:
        d = Path(d)
        T._stage(d / "stage")
        T._build(d / "stage", d / "sim")
        t_end, n, bad = time.time() + args.seconds, 0, 0
        while time.time() < t_end:
            case, W = random_case(rng)
            env = {}
            if W > 1 and rng.random() < 0.3:
                env = {"TDIST_SIM_SLOW_RANK": str(rng.randrange(W)), "TDIST_SIM_SLOW_MS": str(rng.choice([1, 3, 8]))}
            rc, out = T._run(d / "sim", case, env)
            n += 1
            ok = rc == 0 and "ThreadSanitizer" not in out and "timed out" not in out and (
                "data ok" in out or ": ok" in out or "needs more than one rank" in out)
            if not ok:
                bad += 1
                print("FAIL", " ".join(map(str, case)), env, "\n", out[-1500:], flush=True)
        print(f"fuzz_protocol_sim: {n} runs, {bad} failures (seed {args.seed})")
        sys.exit(1 if bad else 0)
