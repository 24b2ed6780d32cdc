"""The public API checklist of SURVEY.md section 8: every name a user of the reference imports must exist here."""


def test_public_api_checklist():
    import importlib
    spec = {
     "triton_dist.utils": "initialize_distributed finalize_distributed init_seed init_nvshmem_by_torch_process_group nvshmem_create_tensor nvshmem_create_tensors nvshmem_free_tensor_sync nvshmem_barrier_all_on_stream nvshmem_signal_wait TorchStreamWrapper NVSHMEM_SIGNAL_DTYPE perf_func dist_print group_profile get_torch_prof_ctx assert_allclose generate_data sleep_async CUDA_CHECK is_cuda is_hip has_tma has_fullmesh_nvlink is_nvshmem_multimem_supported supports_p2p_native_atomic requires launch_cooperative_grid_options get_intranode_max_speed get_nvlink_max_speed get_numa_world_size get_device_property get_device_max_shared_memory_size",
     "triton_dist.kernels.allreduce": "AllReduceMethod to_allreduce_method get_allreduce_methods",
     "triton_dist.kernels.common_ops": "wait_eq set_signal BarrierAllContext barrier_all_on_stream",
     "triton_dist.kernels.nvidia": "ag_gemm create_ag_gemm_context gemm_persistent gemm_non_persistent get_auto_all_gather_method AllGatherMethod cp_engine_producer_all_gather_intra_node cp_engine_producer_all_gather_inter_node gemm_rs create_gemm_rs_context create_gemm_ar_context gemm_allreduce_op create_ll_gemm_ar_context low_latency_gemm_allreduce_op fast_allgather create_fast_allgather_context _forward_pull_kernel _forward_push_2d_kernel _forward_push_3d_kernel _forward_push_2d_ll_kernel _forward_push_2d_ll_multimem_kernel _forward_push_numa_2d_kernel _forward_push_numa_2d_ll_kernel _forward_push_numa_2d_ll_multinode_kernel create_all_to_all_context fast_all_to_all all_to_all_post_process create_ag_group_gemm_context ag_group_gemm create_moe_rs_context gqa_fwd_batch_decode gqa_fwd_batch_decode_aot gqa_fwd_batch_decode_persistent gqa_fwd_batch_decode_persistent_aot gqa_fwd_batch_decode_intra_rank gqa_fwd_batch_decode_intra_rank_aot kernel_gqa_fwd_batch_decode_split_kv_persistent kernel_inter_rank_gqa_fwd_batch_decode_combine_kv get_triton_combine_kv_algo_info fused_sp_ag_attn_intra_node fused_sp_ag_attn_inter_node create_sp_ag_attention_context_intra_node create_sp_ag_attention_context_inter_node",
     "triton_dist.kernels.nvidia.allreduce": "create_allreduce_ctx all_reduce get_auto_allreduce_method",
     "triton_dist.kernels.nvidia.reduce_scatter": "create_reduce_scater_2d_ctx reduce_scatter_2d_op ring_reduce",
     "triton_dist.kernels.nvidia.moe_reduce_rs": "run_moe_reduce_rs",
     "triton_dist.kernels.nvidia.common_ops": "BarrierAllContext barrier_all_on_stream",
     "triton_dist.kernels.nvidia.gemm_perf_model": "",
     "triton_dist.kernels.nvidia.comm_perf_model": "",
     "triton_dist.kernels.nvidia.ep_a2a": "",
     "triton_dist.kernels.nvidia.moe_utils": "",
     "triton_dist.layers.nvidia": "TP_MLP TP_Attn TP_MoE GemmARLayer EPAll2AllLayer AllGatherLayer SpGQAFlashDecodeAttention CommOp",
     "triton_dist.models": "ModelConfig AutoLLM AutoTokenizer DenseLLM Qwen3MoE KV_Cache Engine",
     "triton_dist.mega_triton_kernel": "ModelBuilder Qwen3Model PagedKVCache",
     "triton_dist.autotuner": "contextual_autotune",
     "triton_dist.language": "",
    }
    missing = []
    for mod, names in spec.items():
        m = importlib.import_module(mod)
        for n in names.split():
            if not hasattr(m, n):
                missing.append(f"{mod}.{n}")
    from triton_dist.layers.nvidia import TP_MLP, TP_Attn
    for cls in (TP_MLP, TP_Attn):
        for meth in "_init_parameters _init_ctx _init_AR_ctx _init_gemm_ar_ctx torch_fwd dist_triton_fwd dist_triton_AR_fwd dist_triton_gemm_ar_fwd finalize".split():
            if not hasattr(cls, meth): missing.append(f"{cls.__name__}.{meth}")
    for meth in "torch_ag_gemm dist_triton_ag_gemm torch_gemm_rs dist_triton_gemm_rs".split():
        if not hasattr(TP_MLP, meth): missing.append(f"TP_MLP.{meth}")
    from triton_dist.mega_triton_kernel import ModelBuilder
    for meth in "make_fc1 make_fc2 make_qkv_proj make_o_proj make_linear make_attn make_silu_mul_up make_qk_norm_rope_update_kvcache make_rms_norm make_add make_barrier_all_intra_node make_allreduce make_prefetch create_symm_tensor compile run finalize".split():
        if not hasattr(ModelBuilder, meth): missing.append(f"ModelBuilder.{meth}")
    from triton_dist.models import DenseLLM
    for meth in "set_fwd init_triton_dist_ctx init_triton_dist_AR_ctx init_triton_dist_gemm_ar_ctx inference".split():
        if not hasattr(DenseLLM, meth): missing.append(f"DenseLLM.{meth}")
    assert not missing, missing



def test_ctypes_prototypes_match_the_c_header():
    """Every prototype declared in triton_dist/_lib.py has the argument count and the argument KINDS (pointer, 32-bit integer,
    size_t, 64-bit integer, float) of its declaration in include/tdist.h, and every function of the header that Python calls
    through the library has a prototype (an undeclared function would get 32-bit integer arguments for its pointers)."""
    import ctypes
    import re
    from pathlib import Path

    from triton_dist import _lib

    root = Path(__file__).resolve().parent.parent
    header = re.sub(r"/\*.*?\*/", " ", (root / "include" / "tdist.h").read_text(), flags=re.S)
    header = re.sub(r"//[^\n]*", " ", header)
    protos = {}
    for m in re.finditer(r"\b(?:int|void\s*\*|size_t|const\s+char\s*\*)\s+(tdist_\w+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        params = [q.strip() for q in m.group(2).replace("\n", " ").split(",") if q.strip() and q.strip() != "void"]
        protos[m.group(1)] = params
    assert len(protos) > 60, len(protos)

    def kind_c(param):
        if "*" in param or "[" in param:
            return "ptr"
        t = param.rsplit(" ", 1)[0].strip() if " " in param else param
        t = t.replace("const ", "").strip()
        return {"int": "i32", "uint32_t": "i32", "unsigned": "i32", "unsigned int": "i32", "size_t": "i64", "long long": "i64",
                "unsigned long long": "i64", "uint64_t": "i64", "float": "f32"}[t]

    def kind_py(t):
        if t in (ctypes.c_void_p, ctypes.c_char_p) or hasattr(t, "_type_") and not isinstance(t._type_, str):
            return "ptr"  # c_void_p or POINTER(struct)
        return {ctypes.c_int: "i32", ctypes.c_uint32: "i32", ctypes.c_size_t: "i64", ctypes.c_longlong: "i64",
                ctypes.c_ulonglong: "i64", ctypes.c_float: "f32"}[t]

    lib = _lib.load(required=True)
    checked = 0
    for name, params in protos.items():
        fn = getattr(lib, name)
        if fn.argtypes is None:
            continue  # declared where it is used (tutorial and probe entry points); see below
        got = [kind_py(t) for t in fn.argtypes]
        want = [kind_c(q) for q in params]
        assert got == want, f"{name}: ctypes {got} vs header {want}"
        checked += 1
    assert checked > 55, checked
    # functions the package calls: every `lib().tdist_xxx(` / `lib.tdist_xxx(` in the Python sources is declared
    used = set()
    for py in (root / "triton_dist").rglob("*.py"):
        used |= set(re.findall(r"\.(tdist_\w+)\(", py.read_text()))
    undeclared = sorted(n for n in used if n in protos and getattr(lib, n).argtypes is None)
    local_decl = {"tdist_link_probe"}  # declared next to its only caller
    assert set(undeclared) <= local_decl, undeclared


def test_locally_declared_ctypes_prototypes_match_the_c_header():
    """Entry points whose ctypes prototype is written next to the only caller (tutorials 01 / 07 / 08, the link probe, the
    no-alloc test) instead of triton_dist/_lib.py: the `argtypes` expression there, evaluated, has the argument count and
    kinds of the declaration in include/tdist.h."""
    import ctypes
    import re
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    header = re.sub(r"/\*.*?\*/", " ", (root / "include" / "tdist.h").read_text(), flags=re.S)
    protos = {}
    for m in re.finditer(r"\b(?:int|void\s*\*|size_t|const\s+char\s*\*)\s+(tdist_\w+)\s*\(([^;{]*?)\)\s*;", header, flags=re.S):
        protos[m.group(1)] = [q.strip() for q in m.group(2).replace("\n", " ").split(",") if q.strip() and q.strip() != "void"]

    def kind_c(param):
        if "*" in param or "[" in param:
            return "ptr"
        t = param.rsplit(" ", 1)[0].replace("const ", "").strip()
        return {"int": "i32", "uint32_t": "i32", "unsigned": "i32", "size_t": "i64", "long long": "i64", "float": "f32"}[t]

    kind_py = {ctypes.c_void_p: "ptr", ctypes.c_int: "i32", ctypes.c_uint32: "i32", ctypes.c_size_t: "i64", ctypes.c_longlong: "i64",
               ctypes.c_float: "f32"}
    env = {"ctypes": ctypes, "vp": ctypes.c_void_p, "i32": ctypes.c_int, "f32": ctypes.c_float, "ll": ctypes.c_longlong}
    files = list((root / "tutorials").glob("*.py")) + [root / "tests" / "test_zz_no_alloc.py",
                                                       root / "triton_dist" / "benchmark" / "bench_link_probe.py"]
    checked = set()
    for py in files:
        text = py.read_text()
        for m in re.finditer(r"\.(tdist_\w+)\.argtypes\s*=\s*", text):
            # the expression runs to the end of the line on which its brackets are balanced
            end, depth = m.end(), 0
            while True:
                ch = text[end]
                depth += ch in "([" 
                depth -= ch in ")]"
                if ch == "\n" and depth == 0:
                    break
                end += 1
            got = [kind_py[t] for t in eval(text[m.end():end], dict(env))]
            want = [kind_c(q) for q in protos[m.group(1)]]
            assert got == want, f"{py.name}: {m.group(1)}: ctypes {got} vs header {want}"
            checked.add(m.group(1))
    assert {"tdist_tutorial_queue", "tdist_tutorial_ag_gemm", "tdist_tutorial_gemm_rs", "tdist_link_probe", "tdist_flash_decode_noalloc",
            "tdist_flash_decode"} <= checked, checked


def test_ctypes_structs_match_the_c_header(tmp_path):
    """The ctypes mirrors of the C ABI structs (context, GEMM / fused arguments, megakernel op) have the size and the field
    offsets a C compiler gives include/tdist.h: same field order, types and padding."""
    import ctypes
    import re
    import shutil
    import subprocess
    from pathlib import Path

    import pytest

    from triton_dist import _lib

    if shutil.which("gcc") is None:
        pytest.skip("gcc not available")
    root = Path(__file__).resolve().parent.parent
    header = (root / "include" / "tdist.h").read_text()
    pairs = {"tdist_symm_ctx": _lib.SymmCtx, "tdist_gemm_args": _lib.GemmArgs, "tdist_fused_args": _lib.FusedArgs,
             "tdist_mega_op": _lib.MegaOp}
    rename = {("tdist_mega_op", "inp"): "in"}  # `in` is a Python keyword
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "tdist.h"', 'int main(void) {']
    for cname, cls in pairs.items():
        body = re.search(r"typedef struct " + cname + r"\s*\{(.*?)\}\s*" + cname + r"\s*;", header, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", " ", body, flags=re.S)
        c_fields = []
        for decl in (d for d in body.split(";") if d.strip()):  # "int M, N, K" declares three fields
            for part in decl.split(","):
                c_fields.append(re.sub(r"\[.*", "", part.split()[-1].lstrip("*")))
        py_fields = [rename.get((cname, n), n) for n, _ in cls._fields_]
        assert c_fields == py_fields, f"{cname}: header fields {c_fields} vs ctypes {py_fields}"
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for f in c_fields:
            lines.append(f'  printf("{cname} {f} %zu\\n", offsetof({cname}, {f}));')
    lines += ["  return 0;", "}"]
    (tmp_path / "abi.c").write_text("\n".join(lines))
    subprocess.run(["gcc", f"-I{root / 'include'}", str(tmp_path / "abi.c"), "-o", str(tmp_path / "abi")], check=True)
    out = subprocess.run([str(tmp_path / "abi")], capture_output=True, text=True, check=True).stdout
    for line in out.splitlines():
        cname, field, value = line.split()
        cls = pairs[cname]
        if field == "size":
            assert ctypes.sizeof(cls) == int(value), f"{cname}: sizeof {ctypes.sizeof(cls)} vs {value}"
        else:
            py_name = {v: k[1] for k, v in rename.items() if k[0] == cname}.get(field, field)
            assert getattr(cls, py_name).offset == int(value), f"{cname}.{field}: offset {getattr(cls, py_name).offset} vs {value}"


def test_native_call_sites_pass_the_declared_number_of_arguments():
    """Every call of a native entry point in the Python sources (``lib.tdist_xxx(...)``) passes as many positional arguments as
    the ctypes prototype declares: most of these calls sit on GPU-only paths, where a miscount would only show on a GPU."""
    import ast
    from pathlib import Path

    from triton_dist import _lib

    root = Path(__file__).resolve().parent.parent
    lib = _lib.load(required=True)
    checked, problems = 0, []
    files = list((root / "triton_dist").rglob("*.py")) + [root / "bench.py", root / "__graft_entry__.py"]
    for py in files:
        tree = ast.parse(py.read_text())
        for node in ast.walk(tree):
            if not (isinstance(node, ast.Call) and isinstance(node.func, ast.Attribute) and node.func.attr.startswith("tdist_")):
                continue
            name = node.func.attr
            fn = getattr(lib, name, None)
            if fn is None:
                problems.append(f"{py.name}:{node.lineno}: {name} is not exported by the library")
                continue
            if fn.argtypes is None or any(isinstance(a, ast.Starred) for a in node.args) or node.keywords:
                continue
            checked += 1
            if len(node.args) != len(fn.argtypes):
                problems.append(f"{py.relative_to(root)}:{node.lineno}: {name} called with {len(node.args)} arguments, prototype has "
                                f"{len(fn.argtypes)}")
    assert not problems, "\n".join(problems)
    assert checked > 60, checked


def test_no_undefined_names_in_the_python_sources():
    """A name that is read but bound nowhere (module, enclosing function, builtins) in code that only runs on a GPU would first
    fail on a GPU: scan every source file with the compiler's own symbol tables."""
    import builtins
    import symtable
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    files = (list((root / "triton_dist").rglob("*.py")) + [root / "bench.py", root / "__graft_entry__.py", root / "build.py"] +
             list((root / "tutorials").glob("*.py")) + list((root / "scripts").glob("*.py")) + list((root / "tests").glob("*.py")))
    known = set(dir(builtins)) | {"__file__", "__name__", "__doc__", "__builtins__", "__spec__", "__package__", "__path__"}
    problems = []
    for f in files:
        src = f.read_text()
        if "import *" in src:
            continue
        top = symtable.symtable(src, str(f), "exec")
        module_names = {s.get_name() for s in top.get_symbols() if s.is_assigned() or s.is_imported() or s.is_namespace()}

        def walk(table):
            for s in table.get_symbols():
                name = s.get_name()
                if not s.is_referenced() or name in known:
                    continue
                if table.get_type() == "module":
                    if not (s.is_assigned() or s.is_imported() or s.is_namespace()):
                        problems.append(f"{f.relative_to(root)}: {name}")
                elif s.is_global() and not s.is_declared_global() and name not in module_names:
                    problems.append(f"{f.relative_to(root)}: {table.get_name()}(): {name}")
            for child in table.get_children():
                walk(child)

        walk(top)
    assert len(files) > 150 and not problems, problems


def test_module_attribute_references_resolve():
    """`from x import name` and `alias.name` on the package's own modules, anywhere in the sources (also in branches that only
    run on a GPU): the module really has that attribute. Catches a function renamed or removed under a caller that no CPU
    test reaches."""
    import ast
    import importlib
    import types
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    files = (list((root / "triton_dist").rglob("*.py")) + [root / "bench.py", root / "__graft_entry__.py"] +
             list((root / "tutorials").glob("*.py")) + list((root / "tests").glob("*.py")))
    problems, resolved = [], 0

    def has(mod, name):
        if hasattr(mod, name):
            return True
        try:
            importlib.import_module(mod.__name__ + "." + name)
            return True
        except Exception:
            return False

    for f in files:
        tree = ast.parse(f.read_text())
        rel = f.relative_to(root)
        package = ".".join(rel.parts[:-1]) if rel.parts[0] == "triton_dist" else ""
        alias = {}
        for node in ast.walk(tree):
            if isinstance(node, ast.ImportFrom):
                if node.level:
                    if not package:
                        continue
                    base = package.split(".")[:len(package.split(".")) - node.level + 1]
                    modname = ".".join(base + ([node.module] if node.module else []))
                elif node.module and node.module.startswith("triton_dist"):
                    modname = node.module
                else:
                    continue
                try:
                    mod = importlib.import_module(modname)
                except Exception as e:  # noqa: BLE001
                    problems.append(f"{rel}:{node.lineno}: import {modname}: {e}")
                    continue
                for a in node.names:
                    if a.name == "*":
                        continue
                    if not has(mod, a.name):
                        problems.append(f"{rel}:{node.lineno}: {modname} has no attribute {a.name}")
                        continue
                    resolved += 1
                    obj = getattr(mod, a.name, None)
                    if obj is None:
                        obj = importlib.import_module(modname + "." + a.name)
                    if isinstance(obj, types.ModuleType):
                        alias[a.asname or a.name] = obj
            elif isinstance(node, ast.Import):
                for a in node.names:
                    if a.name.startswith("triton_dist") and a.asname:
                        alias[a.asname] = importlib.import_module(a.name)
        for node in ast.walk(tree):
            if isinstance(node, ast.Attribute) and isinstance(node.value, ast.Name) and node.value.id in alias:
                resolved += 1
                if not has(alias[node.value.id], node.attr):
                    problems.append(f"{rel}:{node.lineno}: {node.value.id}.{node.attr}: {alias[node.value.id].__name__} has no such attribute")
    assert resolved > 1000 and not problems, (resolved, problems)


def test_calls_of_package_functions_bind_to_their_signatures():
    """Every call of a function or class of this package through an imported name (`f(...)` after `from x import f`,
    `mod.f(...)` after `from x import mod`), anywhere in the sources, binds to the callee's signature: number of positional
    arguments and keyword names. Calls with * / ** arguments are skipped."""
    import ast
    import importlib
    import inspect
    import types
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    files = (list((root / "triton_dist").rglob("*.py")) + [root / "bench.py", root / "__graft_entry__.py"] +
             list((root / "tutorials").glob("*.py")) + list((root / "tests").glob("*.py")))
    problems, checked = [], 0
    for f in files:
        tree = ast.parse(f.read_text())
        rel = f.relative_to(root)
        package = ".".join(rel.parts[:-1]) if rel.parts[0] == "triton_dist" else ""
        names, rebound = {}, set()
        for node in ast.walk(tree):
            if isinstance(node, ast.ImportFrom):
                if node.level:
                    if not package:
                        continue
                    parts = package.split(".")
                    modname = ".".join(parts[:len(parts) - node.level + 1] + ([node.module] if node.module else []))
                elif node.module and node.module.startswith("triton_dist"):
                    modname = node.module
                else:
                    continue
                try:
                    mod = importlib.import_module(modname)
                except Exception:  # noqa: BLE001  (reported by test_module_attribute_references_resolve)
                    continue
                for a in node.names:
                    obj = getattr(mod, a.name, None)
                    if obj is None and a.name != "*":
                        try:
                            obj = importlib.import_module(modname + "." + a.name)
                        except Exception:  # noqa: BLE001
                            continue
                    names[a.asname or a.name] = obj
            elif isinstance(node, (ast.FunctionDef, ast.ClassDef)):
                rebound.add(node.name)  # a local definition of the same name shadows the import
            elif isinstance(node, ast.Name) and isinstance(node.ctx, ast.Store):
                rebound.add(node.id)
            elif isinstance(node, ast.arg):
                rebound.add(node.arg)
        for node in ast.walk(tree):
            if not isinstance(node, ast.Call):
                continue
            fn = node.func
            target = None
            if isinstance(fn, ast.Name) and fn.id in names and fn.id not in rebound:
                target = names[fn.id]
            elif (isinstance(fn, ast.Attribute) and isinstance(fn.value, ast.Name) and fn.value.id in names and fn.value.id not in rebound
                  and isinstance(names[fn.value.id], types.ModuleType)):
                target = getattr(names[fn.value.id], fn.attr, None)
            if target is None or not (inspect.isfunction(target) or inspect.isclass(target)):
                continue
            if not getattr(target, "__module__", "").startswith("triton_dist"):
                continue
            if any(isinstance(a, ast.Starred) for a in node.args) or any(k.arg is None for k in node.keywords):
                continue
            try:
                sig = inspect.signature(target)
            except (TypeError, ValueError):
                continue
            try:
                sig.bind(*[None] * len(node.args), **{k.arg: None for k in node.keywords})
                checked += 1
            except TypeError as e:
                problems.append(f"{rel}:{node.lineno}: {getattr(target, '__qualname__', target)}{sig}: {e}")
    assert checked > 600 and not problems, (checked, problems)
