"""Protocol simulator: device code of this library is compiled for the CPU and run with 1-8 simulated ranks in one process
under ThreadSanitizer (csrc/sim/, docs/simulator.md), on top of the REAL symm.cuh and common.cuh:
  * the PRODUCTION collectives of csrc/kernels/collectives.cu (barrier, 4 all-reduces, 5 all-gathers incl. the LL packet
    protocol, 2 reduce-scatters; NVLS emulated by reading / writing every rank's copy), primitives_selftest.cu (every
    NVSHMEM-style primitive of csrc/device/symm.cuh), the MoE kernels of moe.cu and the expert-parallel kernels of ep_a2a.cu;
  * the communication side of the fused tcgen05 GEMMs, both kernels and every form: push bodies and the producer warp's
    arrival wait of AG + GEMM, reducers, publishers and tile orders of GEMM + RS / GEMM + AR, around stand-ins for the TMA
    loads and the tensor-core epilogue;
  * the decode megakernel (csrc/mega/megakernel.cu) with its native builder (builder.cc) as they are — TMA tiled loads,
    mbarrier phases, ldmatrix and mma.sync emulated — on a small Qwen3-style decoder, against a double-precision model;
  * the decode-path kernels of decode.cu (functional, against host oracles) and the tutorial-size kernels (csrc/tutorials).
Only csrc/device/ptx.cuh, csrc/runtime/driver.h and the CUDA headers are replaced by host stand-ins.

What a pass means: for every payload access of every rank there is a release / acquire chain through the flags (no data
race), nobody waits for a flag that never comes (no wait bound), and the results equal a host computation — over several
back-to-back calls per rank without any barrier in between, with calls of different sizes on the same buffers and with one
rank straggling per flag write. The negative controls (flags written without release ordering) must be reported as races
and the mutation controls (one expression of a kernel changed) as wrong results, otherwise a pass would prove nothing.
The cases run through a small process pool (see _POOL). No GPU and no CUDA toolkit needed: g++ with -fsanitize=thread."""
import os
import re
import shutil
import subprocess
from pathlib import Path

from concurrent.futures import ThreadPoolExecutor

import pytest

ROOT = Path(__file__).resolve().parent.parent
SIM = ROOT / "csrc" / "sim"


def _stage(dst: Path) -> None:
    """Lay the sources out as in the repository, with the host ptx.cuh in place of the device one. The kernel files are
    copied with two mechanical edits: cut before their launch wrappers (they use <<< >>>), and `__shared__ T a[..], b[..];`
    becomes references to per-CTA buffers. csrc/kernels/common.cuh and csrc/device/symm.cuh are the REAL ones."""
    for d in ("include", "csrc/device", "csrc/kernels", "csrc/tutorials", "csrc/mega", "csrc/runtime"):
        (dst / d).mkdir(parents=True)
    shutil.copy(ROOT / "include" / "tdist.h", dst / "include" / "tdist.h")
    shutil.copy(SIM / "shim" / "device" / "ptx.cuh", dst / "csrc" / "device" / "ptx.cuh")

    def shared(m):
        out = []
        for decl in m.group(2).split(","):
            decl = decl.strip()
            name = re.match(r"\w+", decl).group(0)
            out.append(f"typedef {m.group(1)} {name}_sim_t{decl[len(name):]}; SIM_SHARED({name}_sim_t, {name});")
        return " ".join(out)

    def copy(rel, cut, asm=()):
        src = (ROOT / "csrc" / rel).read_text()
        if cut:
            src = src[:src.index(cut if isinstance(cut, str) else '\nextern "C"')] + (cut if isinstance(cut, str) else "")
        src = re.sub(r"extern\s+__shared__\s+(?:__align__\(\d+\)\s+)?(\w+)\s+(\w+)\[\];", r"SIM_DYN_SHARED(\1, \2);", src)  # dynamic shared memory
        for name in asm:  # a function whose body is one inline-asm statement: call the host emulation of that instruction instead
            a = src.index("asm volatile(", src.index(f" void {name}("))
            args = re.findall(r"[&\s(]\(?&?(\w+)\)?(?:\[\d+\])?\s*[,)]", src[src.index(f" void {name}("):a])
            src = src[:a] + f"ptx::sim_{name}({', '.join(args)});" + src[src.index("\n}", a):]
        (dst / "csrc" / rel).write_text(re.sub(r"__shared__\s+(\w+)\s+([^;]+);", shared, src))

    copy("device/symm.cuh", False)
    copy("kernels/common.cuh", False)
    for rel in ("tutorials/notify_wait_tutorial.cu", "tutorials/overlap_tutorial.cu", "kernels/collectives.cu",
                "kernels/primitives_selftest.cu", "kernels/moe.cu", "kernels/ep_a2a.cu"):
        copy(rel, True)
    copy("kernels/decode.cu", "\n}  // namespace tdist\n")  # the kernels only: the launch helpers that follow use the runtime API
    # the decode megakernel with its native builder: the interpreter kernel up to its launch wrapper, ldmatrix / mma.sync
    # through the host emulations of csrc/sim/shim/device/ptx.cuh; builder.cc and mega.h as they are, on a stand-in driver.h
    copy("mega/megakernel.cu", True, asm=("ldmatrix_x4", "mma_bf16_16816"))
    for rel in ("mega/mega.h", "mega/builder.cc"):
        shutil.copy(ROOT / "csrc" / rel, dst / "csrc" / rel)
    shutil.copy(SIM / "shim" / "runtime" / "driver.h", dst / "csrc" / "runtime" / "driver.h")
    # the fused tcgen05 GEMM: parameter block, tile order and communication bodies (everything before the kernel template),
    # plus the publisher loops of the kernel and the arrival wait of its TMA producer warp in AG mode, i.e. the text between
    # its [sim:...] markers, with a yield in the publishers' spin on the shared tile counter (csrc/sim/fused_gemm_scenarios.h
    # wraps them)
    gemm = (ROOT / "csrc" / "kernels" / "gemm_sm100.cuh").read_text()
    cut = gemm.index("\ntemplate <int BLOCK_N, int STAGES>\n__global__")
    (dst / "csrc" / "kernels" / "gemm_sm100.cuh").write_text(gemm[:cut] + "\n}  // namespace tdist\n")
    spin = re.compile(r"(while \(\*reinterpret_cast<volatile uint32_t\*>\(stored_count\) <= k\) \{)\s*\}")
    pair = (ROOT / "csrc" / "kernels" / "gemm_sm100_2cta.cuh").read_text()  # the CTA-pair kernel: tile order + publisher
    cut = pair.index("\ntemplate <int BLOCK_N, int STAGES>\n__global__")
    (dst / "csrc" / "kernels" / "gemm_sm100_2cta.cuh").write_text(pair[:cut] + "\n}  // namespace tdist\n")
    for name, text in (("rs_publish", gemm), ("ar_publish", gemm), ("rs_publish_pair", pair), ("ag_wait", gemm), ("ag_wait_pair", pair),
                       ("epi_publish", gemm)):
        a = text.index("\n", text.index(f"[sim:{name} begin]")) + 1
        snippet = text[a:text.index(f"// [sim:{name} end]")]
        assert spin.search(snippet) or name.startswith("ag_wait") or name == "epi_publish", name  # (the AG wait polls flags, which yield by themselves)
        (dst / "csrc" / "kernels" / f"gemm_{name}.inc").write_text(spin.sub(r"\1 sim_yield(); }", snippet))


def _build(stage_dir: Path, out: Path, *defines: str) -> None:
    # (the device wait bound of symm.cuh is compiled in: 45 s here instead of 60; several cases share the cores)
    cmd = ["g++", "-std=c++20", "-O1", "-g", "-fsanitize=thread", "--param=tsan-instrument-func-entry-exit=0", "-DTDIST_SPIN_TIMEOUT_S=45", *defines,
           f"-I{stage_dir / 'csrc'}", f"-I{SIM / 'stub'}", f"-I{SIM}", "-x", "c++", str(SIM / "protocol_sim.cc"), "-o", str(out),
           "-pthread"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr


@pytest.fixture(scope="module")
def sims(tmp_path_factory):
    if shutil.which("g++") is None:
        pytest.skip("g++ not available")
    d = tmp_path_factory.mktemp("protocol_sim")
    _stage(d / "stage")
    probe = subprocess.run(["g++", "-fsanitize=thread", "-x", "c++", "-", "-o", str(d / "probe")], input="int main(){}",
                           capture_output=True, text=True)
    if probe.returncode != 0 or subprocess.run([str(d / "probe")], capture_output=True).returncode != 0:
        pytest.skip("ThreadSanitizer runtime not available (or it cannot run on this kernel's address-space layout)")
    builds = [_POOL.submit(_build, d / "stage", d / "sim"), _POOL.submit(_build, d / "stage", d / "sim_broken", "-DTDIST_SIM_BREAK_RELEASE")]
    for b in builds:
        b.result()
    good, broken = d / "sim", d / "sim_broken"
    for case in CLEAN:  # run ahead, in the order the tests will ask for them
        _submit(good, case)
    for case, env in AHEAD:
        _submit(good, case, env)
    for case in NEGATIVE:
        _submit(broken, case)
    for name, (rel, old, new, expect, case, env) in MUTANTS.items():  # mutants: stage, build and run in the pool as well
        _jobs[("mutant", name)] = _POOL.submit(_mutant, d / f"mut_{name}", rel, old, new, case, env)
    yield good, broken
    for f in _jobs.values():  # a run with -k leaves most of the queue unasked for: drop it instead of finishing it at exit
        f.cancel()


def _mutant(dst, rel, old, new, case, env):
    _stage(dst / "stage")
    f = dst / "stage" / "csrc" / rel
    src = f.read_text()
    assert old in src, f"the mutated expression is no longer in {rel}: update MUTANTS"
    f.write_text(src.replace(old, new))
    _build(dst / "stage", dst / "sim_mut")
    return _run_now(dst / "sim_mut", case, env)


def _run_now(exe, args, env_extra=None):
    env = dict(os.environ)
    env.update(env_extra or {})
    r = subprocess.run([str(exe)] + [str(a) for a in args], capture_output=True, text=True, timeout=900, env=env)
    return r.returncode, r.stdout + r.stderr


# The cases are independent processes: a small pool runs them ahead of the tests that look at them (the `sims` fixture submits
# everything once both binaries exist), which cuts the wall time of this file to a third. Not more workers than half the
# cores: a case is itself ranks x CTAs spinning threads, and its waits are bounded in wall-clock time.
_CORES = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 2)
_POOL = ThreadPoolExecutor(max_workers=max(1, min(4, _CORES // 2)))
_jobs = {}


def _key(exe, args, env_extra):
    return str(exe), tuple(str(a) for a in args), tuple(sorted((env_extra or {}).items()))


def _submit(exe, args, env_extra=None):
    k = _key(exe, args, env_extra)
    if k not in _jobs:
        _jobs[k] = _POOL.submit(_run_now, exe, args, env_extra)
    return _jobs[k]


def _run(exe, args, env_extra=None, fresh=False):
    """(rc, output) of one simulator run; the result of a run that was submitted ahead is reused unless `fresh`."""
    if fresh:
        return _run_now(exe, args, env_extra)
    fut = _submit(exe, args, env_extra)
    if fut.cancel():  # still queued (e.g. a run with -k asks for a late case first): run it here instead of waiting for its turn
        return _run_now(exe, args, env_extra)
    return fut.result()


CLEAN = [
    # queue: ranks, items, vecs per item, slots, calls
    ("queue", 3, 11, 300, 3, 3), ("queue", 5, 23, 17, 2, 6), ("queue", 1, 9, 33, 2, 3), ("queue", 4, 20, 40, 1, 4),
    # ag_gemm: ranks, rows per rank, N, K, push CTAs, CTAs, calls, alternate (1 odd / 2 even calls smaller), divisor
    ("ag_gemm", 3, 96, 72, 32, 2, 5, 4), ("ag_gemm", 8, 64, 24, 16, 2, 4, 4), ("ag_gemm", 1, 70, 40, 16, 1, 3, 2),
    ("ag_gemm", 4, 130, 40, 24, 1, 3, 5), ("ag_gemm", 3, 128, 40, 16, 1, 3, 6, 2), ("ag_gemm", 4, 128, 40, 16, 1, 3, 7, 1),
    # gemm_rs: ranks, M, N, K per rank, CTAs, calls, alternate, divisor
    ("gemm_rs", 3, 288, 72, 32, 4, 4), ("gemm_rs", 8, 512, 24, 16, 5, 4), ("gemm_rs", 1, 70, 40, 16, 2, 2),
    ("gemm_rs", 4, 520, 40, 24, 3, 5), ("gemm_rs", 3, 384, 40, 16, 3, 6, 2), ("gemm_rs", 2, 256, 72, 16, 2, 8, 2),
]


# production collectives: op, ranks, 16-byte vectors, CTAs, calls, alternate (1 / 2: odd / even calls use a third of the size)
COLLECTIVES = [("coll", op, 3, 96, 3, 4) for op in ("barrier", "ar1", "ar2", "ar1mc", "ar2mc", "ar1mc_inplace", "ar2mc_inplace",
                                                     "ag_pull", "ag_push", "ag_mc", "ag_ll", "ag_llmc", "rs", "rsmc", "selftest")]
COLLECTIVES += [("coll", op, 4, 120, 3, 6, alt) for op in ("ar1", "ar2", "ar1mc", "ar2mc", "ag_ll", "ag_llmc", "rs", "rsmc")
                for alt in (1, 2)]
COLLECTIVES += [("coll", "ar2mc", 8, 64, 2, 3, 1), ("coll", "selftest", 8, 64, 1, 3), ("coll", "ar1", 1, 40, 2, 2),
                ("coll", "rs", 2, 8, 4, 3)]
CLEAN += COLLECTIVES
# production MoE / EP kernels (argument lists: csrc/sim/protocol_sim.cc)
CLEAN += [("moe_utils", 96, 2, 5, 3, 2), ("moe_utils", 256, 4, 16, 4, 1), ("moe_rs", 3, 48, 2, 3, 3, 4), ("moe_rs", 4, 64, 2, 2, 3, 6, 2),
          ("moe_rs", 8, 64, 2, 2, 2, 4, 1), ("a2a", 3, 24, 3, 2, 2, 4), ("a2a", 8, 8, 2, 2, 1, 3), ("ep", 3, 12, 256, 6, 2, 3),
          ("ep", 4, 8, 256, 8, 3, 3), ("ep", 8, 8, 256, 16, 2, 2), ("ep", 2, 8, 256, 4, 2, 2, 1), ("ep", 4, 8, 512, 8, 2, 2, 1)]
# communication side of the production fused GEMM: mode, ranks, M, N, GEMM CTAs, calls, in-switch reduction, alternate,
# fixed flag rows (GemmParams::flag_row, TDIST_FIXED_FLAG_LAYOUT=1)
CLEAN += [("fused", "rs", 3, 768, 128, 3, 3, 1), ("fused", "rs", 3, 768, 128, 3, 3, 0), ("fused", "rs", 4, 1024, 64, 2, 4, 1, 2),
          ("fused", "rs", 8, 2048, 64, 2, 3, 1), ("fused", "ar1", 3, 256, 128, 2, 3, 1), ("fused", "ar1", 4, 256, 128, 2, 4, 1, 1),
          ("fused", "ar2", 3, 256, 128, 2, 3, 1), ("fused", "ar2", 2, 256, 128, 2, 2, 0), ("fused", "ar2", 8, 256, 64, 2, 2, 1),
          ("fused", "rs", 3, 768, 128, 3, 4, 1, 2, 1), ("fused", "rs", 8, 2048, 64, 2, 3, 1, 1, 1), ("fused", "rs", 3, 768, 128, 3, 3, 0, 0, 1),
          ("fused", "ar1", 4, 256, 128, 2, 4, 1, 1, 1), ("fused", "ar2", 4, 256, 128, 2, 4, 1, 2, 1),
          # "rsl": the scatter form of GEMM + RS (no multicast mapping): stores into the owner's slots, the epilogue's publish
          # lambda (tile counter per (owner, block), then the flag), rs_reduce_body; last argument: fixed flag rows
          ("fused", "rsl", 3, 768, 128, 3, 3, 0), ("fused", "rsl", 2, 512, 64, 2, 4, 0, 1), ("fused", "rsl", 8, 2048, 64, 4, 2, 0),
          ("fused", "rsl", 4, 1024, 192, 5, 4, 0, 2), ("fused", "rsl", 4, 1024, 64, 2, 4, 0, 1, 1),
          # "rs2": the CTA-pair kernel's publisher and tile order (the flagship's default), incl. shards of 1.5 pair tiles
          ("fused", "rs2", 2, 1024, 128, 2, 2, 1), ("fused", "rs2", 3, 1536, 128, 4, 3, 1), ("fused", "rs2", 4, 1536, 64, 2, 4, 1, 2),
          ("fused", "rs2", 4, 1536, 64, 4, 4, 1, 1, 1), ("fused", "rs2", 8, 2048, 64, 2, 2, 1), ("fused", "rs2", 3, 1152, 128, 2, 3, 0, 0, 1)]

# the production AG + GEMM in its in-switch broadcast form (real ag_push_mc_body, real arrival wait of the TMA producer warp, real
# tile order): "ag", ranks, rows per rank, N, K, push CTAs, CTAs, calls, alternate, row slices, K slabs
CLEAN += [("fused", "ag", 3, 192, 128, 192, 2, 5, 3, 0, 2, 2), ("fused", "ag", 8, 128, 64, 128, 8, 11, 2, 0, 8, 1),
          ("fused", "ag", 2, 300, 128, 64, 1, 2, 4, 1, 4, 3), ("fused", "ag", 4, 256, 192, 256, 3, 6, 4, 2, 2, 4),
          ("fused", "ag", 4, 200, 64, 512, 5, 9, 3, 1, 1, 8),
          # slices = 0: the point-to-point form (bulk-copy ring to every peer, 4 flagged parts per block); last: fixed flag rows
          ("fused", "ag", 2, 300, 128, 192, 2, 4, 3, 0, 0, 1), ("fused", "ag", 3, 200, 64, 128, 3, 5, 4, 1, 0, 1), ("fused", "ag2", 2, 512, 128, 64, 2, 4, 4, 2, 0, 1),
          ("fused", "ag2", 4, 300, 64, 320, 2, 6, 3, 1, 0, 1, 1), ("fused", "ag", 4, 256, 192, 256, 3, 6, 4, 2, 2, 4, 1), ("fused", "ag2", 8, 256, 64, 128, 8, 12, 2, 1, 8, 1, 1),
          # "ag2": the CTA-pair kernel's producer wait and tile order (each CTA of a pair waits for its 128-row half)
          ("fused", "ag2", 3, 384, 128, 192, 2, 6, 3, 0, 2, 2), ("fused", "ag2", 8, 256, 64, 128, 8, 12, 2, 0, 8, 1),
          ("fused", "ag2", 2, 300, 128, 64, 2, 4, 4, 1, 4, 3), ("fused", "ag2", 4, 512, 192, 256, 4, 8, 4, 2, 2, 4),
          ("fused", "ag2", 3, 200, 64, 128, 2, 4, 3, 1, 1, 1)]
# one decode step through the production decode-path kernels on one simulated GPU: batch, query heads, KV heads, cached tokens,
# KV splits, page size (0 = dense cache), hidden, intermediate, vocabulary, fp16. The head groups 14 / 2 = 7 and 16 / 2 = 8
# and 1 cover the 4 + 2 + 1, 8 and 1-head passes; sequence 1 is always shorter than the number of splits (empty splits)
CLEAN += [("decode", 3, 14, 2, 70, 5, 0, 128, 64, 300, 0), ("decode", 3, 14, 2, 70, 5, 16, 128, 64, 300, 1),
          ("decode", 2, 16, 2, 45, 3, 8, 64, 40, 130, 0), ("decode", 5, 3, 3, 33, 9, 0, 64, 24, 77, 1),
          ("decode", 16, 4, 1, 20, 2, 4, 64, 24, 50, 0), ("decode", 1, 6, 2, 200, 4, 32, 128, 64, 1100, 0)]

# the decode megakernel + its native builder: ranks, CTAs, layers, batch, steps, graph (0 fused, 1 op by op + barrier op,
# 2 fused + one-task attention), in-switch all-reduce, KV heads per rank, query heads per KV head, MLP columns per rank, KV splits
CLEAN += [("mega", 2, 4, 2, 2, 3, 0, 1, 1, 2, 128, 3), ("mega", 3, 5, 2, 3, 3, 1, 0, 2, 3, 128, 4), ("mega", 4, 3, 1, 8, 2, 2, 1, 1, 8, 64, 5),
          ("mega", 8, 2, 1, 1, 2, 0, 1, 1, 1, 64, 2), ("mega", 1, 1, 2, 3, 2, 1, 0, 2, 4, 128, 3), ("mega", 2, 7, 1, 4, 4, 2, 0, 2, 4, 64, 4),
          ("mega", 3, 4, 2, 2, 3, 0, 1, 1, 2, 128, 16), ("mega", 1, 2, 1, 2, 1, 0, 0, 1, 12, 64, 3)]


# straggler runs of the tests below (same arguments and environment), submitted ahead with the rest
_S = lambda rank, ms: {"TDIST_SIM_SLOW_RANK": str(rank), "TDIST_SIM_SLOW_MS": str(ms)}  # noqa: E731
AHEAD = [(("gemm_rs", 4, 768, 64, 16, 1, 2, 1, 3), _S(0, 30)), (("ag_gemm", 4, 192, 40, 16, 1, 3, 3, 1, 3), _S(1, 10)),
         (("mega", 3, 3, 1, 2, 3, 0, 1, 1, 2, 64, 2), _S(1, 5)), (("fused", "ag", 4, 256, 64, 128, 2, 4, 4, 1, 2, 2), _S(0, 50)),
         (("fused", "ag2", 4, 512, 64, 128, 2, 6, 4, 1, 2, 2), _S(0, 50))]
AHEAD += [(("coll", op, 8, 64, 2, 3, 1), _S(1, 5)) for op in ("ar2", "rsmc", "ag_llmc")]
AHEAD += [(("fused", "rs", 4, 1024, 64, 1, 2, 1, 1), _S(0, 100)), (("fused", "rs", 4, 1024, 64, 1, 2, 1, 1, 1), _S(0, 100)),
          (("fused", "ag", 8, 256, 64, 128, 2, 4, 4, 1, 0, 1, 0), _S(3, 60))]
AHEAD += [(("fused", "ag", 8, 256, 64, 128, 2, 4, 4, alt, 0, 1, 1), _S(3, 60)) for alt in (1, 2)]
AHEAD += [(("fused", "rsl", 8, 2048, 64, 1, 4, 0, 1, 1), _S(3, 60))]


@pytest.mark.parametrize("case", CLEAN, ids=lambda c: "-".join(map(str, c)))
def test_protocols_are_race_free(sims, case):
    rc, out = _run(sims[0], case)
    assert rc == 0 and ("data ok" in out or ": ok" in out) and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]


def test_straggler_with_changing_sizes(sims):
    """One rank whose every flag write is 30 ms late while the calls alternate between M and M / 3 on the same buffers: with a
    flag layout (or a workspace stride) that follows the size of the CALL, a rank that is one call ahead writes the word a
    slower rank reads under another meaning — the simulator showed the resulting data race and, when the straggler's older
    epoch overwrote a newer one, a waiter that never woke up. Layouts that are constants of the BUFFERS are clean."""
    rc, out = _run(sims[0], ("gemm_rs", 4, 768, 64, 16, 1, 2, 1, 3), {"TDIST_SIM_SLOW_RANK": "0", "TDIST_SIM_SLOW_MS": "30"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]
    rc, out = _run(sims[0], ("ag_gemm", 4, 192, 40, 16, 1, 3, 3, 1, 3), {"TDIST_SIM_SLOW_RANK": "1", "TDIST_SIM_SLOW_MS": "10"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]
    # the megakernel with one rank 5 ms late at every cross-rank flag: 3 steps back to back on double-buffered partials
    rc, out = _run(sims[0], ("mega", 3, 3, 1, 2, 3, 0, 1, 1, 2, 64, 2), {"TDIST_SIM_SLOW_RANK": "1", "TDIST_SIM_SLOW_MS": "5"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]
    # production AG + GEMM, calls alternating between 256 and 128 rows per rank, one rank 50 ms late at every flag
    for case in (("fused", "ag", 4, 256, 64, 128, 2, 4, 4, 1, 2, 2), ("fused", "ag2", 4, 512, 64, 128, 2, 6, 4, 1, 2, 2)):
        rc, out = _run(sims[0], case, {"TDIST_SIM_SLOW_RANK": "0", "TDIST_SIM_SLOW_MS": "50"})
        assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]
    for op in ("ar2", "rsmc", "ag_llmc"):  # production collectives, 8 ranks, alternating sizes, one rank 5 ms late per flag
        rc, out = _run(sims[0], ("coll", op, 8, 64, 2, 3, 1), {"TDIST_SIM_SLOW_RANK": "1", "TDIST_SIM_SLOW_MS": "5"})
        assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-3000:]


@pytest.mark.xfail(strict=False, reason="known formal gap of the default flag layout of gemm_sm100.cuh (profiles/protocol_sim.md); TDIST_FIXED_FLAG_LAYOUT=1 closes it")
def test_fused_gemm_rs_changing_sizes_with_a_straggler(sims):
    """The production GEMM + RS communication code with calls alternating between M and M / 2 on one context while one rank's
    flag writes are 100 ms late: a rank that is a whole call ahead writes the flag word under which a slower owner waits for
    another rank's tile (the DEFAULT flag layout is computed from the call's M) — the owner reduces a tile that has not been
    written. Needs a skew of a whole kernel inside one wave, which the hardware does not produce; the fix (next test) is
    opt-in until it has run on GPUs. Passes once it is the default."""
    rc, out = _run(sims[0], ("fused", "rs", 4, 1024, 64, 1, 2, 1, 1), {"TDIST_SIM_SLOW_RANK": "0", "TDIST_SIM_SLOW_MS": "100"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out, out[-1500:]


def test_fused_gemm_rs_fixed_flag_rows_survive_the_straggler(sims):
    """The same scenario with one flag row per source rank whose length is a constant of the context (GemmParams::flag_row,
    enabled by TDIST_FIXED_FLAG_LAYOUT=1): every flag word has one writer over all calls — clean."""
    rc, out = _run(sims[0], ("fused", "rs", 4, 1024, 64, 1, 2, 1, 1, 1), {"TDIST_SIM_SLOW_RANK": "0", "TDIST_SIM_SLOW_MS": "100"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-1500:]


@pytest.mark.xfail(strict=False, reason="the same formal gap in the point-to-point form of AG + GEMM (unicast flags placed by the call's block count); TDIST_FIXED_FLAG_LAYOUT=1 closes it")
def test_fused_ag_ring_changing_sizes_with_a_straggler(sims):
    """AG + GEMM in its point-to-point form (the default with one peer and on heaps without a multicast mapping) at 8 ranks, calls
    alternating between 256 and 128 rows per rank, rank 3's flag writes 60 ms late: ranks that receive rank 3's parts first finish
    the call and, one call ahead, write their flags where a slower rank still waits for rank 3's blocks. (The in-switch form is
    immune: a block's flag reaches every rank with one multicast store.)"""
    rc, out = _run(sims[0], ("fused", "ag", 8, 256, 64, 128, 2, 4, 4, 1, 0, 1, 0), {"TDIST_SIM_SLOW_RANK": "3", "TDIST_SIM_SLOW_MS": "60"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out, out[-1500:]


def test_fused_ag_ring_fixed_flag_rows_survive_the_straggler(sims):
    for alt in (1, 2):
        rc, out = _run(sims[0], ("fused", "ag", 8, 256, 64, 128, 2, 4, 4, alt, 0, 1, 1), {"TDIST_SIM_SLOW_RANK": "3", "TDIST_SIM_SLOW_MS": "60"})
        assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-1500:]
    # the scatter form of GEMM + RS (unicast flags [source][block]) under the same conditions
    rc, out = _run(sims[0], ("fused", "rsl", 8, 2048, 64, 1, 4, 0, 1, 1), {"TDIST_SIM_SLOW_RANK": "3", "TDIST_SIM_SLOW_MS": "60"})
    assert rc == 0 and "data ok" in out and "ThreadSanitizer" not in out and "timed out" not in out, out[-1500:]


NEGATIVE = [("queue", 2, 6, 64, 2, 2), ("ag_gemm", 2, 128, 64, 16, 1, 3, 2), ("gemm_rs", 2, 128, 64, 16, 2, 2),
                                  ("coll", "ar2", 2, 96, 2, 2), ("coll", "rsmc", 2, 96, 2, 2), ("coll", "selftest", 2, 96, 2, 2),
                                  ("moe_rs", 2, 32, 2, 3, 2, 2), ("a2a", 2, 16, 3, 2, 1, 2), ("ep", 2, 8, 256, 4, 2, 2),
                                  ("fused", "rs", 2, 256, 64, 1, 1, 1), ("fused", "ar2", 2, 128, 64, 1, 1, 1),
                                  ("mega", 2, 3, 1, 2, 2, 0, 1, 1, 2, 64, 2), ("fused", "ag", 2, 256, 64, 128, 2, 4, 2, 0, 2, 2),
            ("fused", "rsl", 2, 256, 64, 1, 1, 0)]


@pytest.mark.parametrize("case", NEGATIVE, ids=lambda c: "-".join(map(str, c[:3])))
def test_negative_control_is_reported(sims, case):
    """Same kernels, flags written and polled WITHOUT release / acquire ordering: the checker must see the payload race."""
    for attempt in range(4):  # whether the unordered accesses actually meet depends on the schedule of that run
        rc, out = _run(sims[1], case, fresh=attempt > 0)
        if "ThreadSanitizer: data race" in out:
            break
    assert "ThreadSanitizer: data race" in out, out[-2000:]
    assert re.search(r"(notify_wait_tutorial|overlap_tutorial|collectives|primitives_selftest|moe|ep_a2a|megakernel)\.cu:\d+|(symm|ptx|gemm_sm100)\.cuh:\d+|fused_gemm_scenarios\.h:\d+", out), \
        "the report should point into the device sources"


_DECODE_CASE = ("decode", 3, 14, 2, 70, 5, 16, 128, 64, 300, 0)
_MEGA_CASE = ("mega", 2, 3, 1, 2, 2, 0, 0, 1, 2, 64, 2)
# name: (file under csrc/, expression, replacement, what the run must print, case, environment)
MUTANTS = {
    # decode path: rotation sign, head offset of the second GQA pass, slot inside a KV page, tie rule of the sampler
    "rope_sign": ("kernels/decode.cu", "lane < 16 ? f[i] * cc[i] - other * sn[i]", "lane < 16 ? f[i] * cc[i] + other * sn[i]", "rope q", _DECODE_CASE, None),
    "gqa_pass_offset": ("kernels/decode.cu", "const int h0 = kvh * g_total + g_offset;", "const int h0 = kvh * g_total;", "flash decode", _DECODE_CASE, None),
    "page_slot": ("kernels/decode.cu", "size_t(pos % page_size) * kv_seq_stride + size_t(kvh) * kv_head_stride;\n    }\n    return",
                  "size_t(pos / page_size) * kv_seq_stride + size_t(kvh) * kv_head_stride;\n    }\n    return", "flash decode", _DECODE_CASE, None),
    "argmax_tie": ("kernels/decode.cu", "(ov == best && oi < bi)", "(ov == best && oi > bi)", "argmax", _DECODE_CASE, None),
    # megakernel: the ldmatrix address ignores the 128 B swizzle the TMA applied to the stage buffer (wrong weights)
    "swizzle": ("mega/megakernel.cu", "const uint8_t* ap = lp + ((uint32_t(s4 * 2) + a_hi) ^ ph) * 16u;",
                "const uint8_t* ap = lp + (uint32_t(s4 * 2) + a_hi) * 16u;", "DATA MISMATCH", _MEGA_CASE, {"TDIST_SIM_WATCHDOG_S": "240"}),
    # tasks wait for the scoreboard value of the PREVIOUS step: consumers run before their producers
    "scoreboard_target": ("mega/megakernel.cu", "const uint32_t tgt0 = step * uint32_t(t.dep0_cnt), tgt1",
                          "const uint32_t tgt0 = (step - 1) * uint32_t(t.dep0_cnt), tgt1", "ThreadSanitizer: data race", _MEGA_CASE,
                          {"TDIST_SIM_WATCHDOG_S": "240"}),
    # the "up" half of a gate | up weight tile is fetched from the gate rows
    "pair_rows": ("mega/megakernel.cu", "c1 = int(e.row0 + (second ? e.pair_rows : 0u) + blk * e.R);", "c1 = int(e.row0 + blk * e.R);",
                  "DATA MISMATCH", _MEGA_CASE, {"TDIST_SIM_WATCHDOG_S": "240"}),
}


@pytest.mark.parametrize("name", list(MUTANTS))
def test_mutated_kernels_are_caught(sims, name):
    """Mutation controls: ONE expression of csrc/kernels/decode.cu or csrc/mega/megakernel.cu changed in a staged copy — the
    host oracle of the kernel concerned (or the sanitizer, for the scoreboard) must object."""
    rc, out = _jobs[("mutant", name)].result()
    assert rc != 0 and MUTANTS[name][3] in out, out[-2000:]
