"""Counterpart of the reference's test_language_extra.py / test_simt.py: those exercise the Triton-side PTX helpers
(`tid`, `ld_acquire`, `atomic_add`, `multimem_*`, SIMT regions). Here the helpers are CUDA device functions; this test checks
that every primitive name of the reference maps to a native function that exists in the headers (no GPU needed)."""
import re
import sys
from pathlib import Path

from triton_dist.language.core import PRIMITIVES, describe

if __name__ == "__main__":
    root = Path(__file__).resolve().parents[3]
    text = {h: (root / h).read_text() for h in ("csrc/device/ptx.cuh", "csrc/device/symm.cuh")}
    ok = True
    for name, (header, sig, _) in PRIMITIVES.items():
        if header not in text:
            continue  # plain CUDA or "not needed"
        for fn in re.findall(r"[A-Za-z_][A-Za-z0-9_]*(?=\s*\(|\s*/|\s*$)", sig.replace("{", " ").replace("}", " ")):
            if fn in ("int", "const", "uint32_t", "T", "SymmCtx", "SignalOp"):
                continue
            if re.search(r"\b" + re.escape(fn), text[header]) is None and "{" not in sig:
                print(f"{name}: {fn} not found in {header}")
                ok = False
        assert describe(name).startswith(name)
    print(f"{len(PRIMITIVES)} primitive names mapped")
    print("PASS" if ok else "FAIL")
    sys.exit(0 if ok else 1)
