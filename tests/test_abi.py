"""CPU: the C-ABI library builds, loads and exports every symbol include/ctr_b200.h declares, with
prototypes that match the ctypes table; the product path refuses to run without a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_protos():
    src = open(os.path.join(ROOT, "include", "ctr_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return re.findall(r"\b(int64_t|int|const char\*)\s+(ctr_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S)


def test_library_exports_every_declared_symbol(lib_built):
    from deep_learning_b200 import _lib
    lib = _lib.load()
    names = [n for _, n, _ in _header_protos()]
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.EXPORTS)
    assert lib.ctr_abi_version() == 4
    assert lib.ctr_sizeof_embed_cfg() == C.sizeof(_lib.EmbedCfg) == 104


def test_ctypes_prototypes_match_header(lib_built):
    from deep_learning_b200 import _lib

    def ctype(a):
        a = a.strip()
        if a in ("void", ""):
            return None
        if "*" in a:
            if "ctr_embed_cfg" in a:
                return _lib._P_CFG
            if "ctr_cross_cfg" in a:
                return _lib._P_CROSS
            if "ctr_adam_hp" in a:
                return _lib._P_HP
            if "ctr_peer_route" in a:
                return _lib._P_RT
            if "ctr_dense_layout" in a:
                return _lib._P_LAY
            if "ctr_handle**" in a.replace(" ", "") or "void**" in a.replace(" ", ""):
                return C.POINTER(C.c_void_p)
            return C.c_void_p
        ty = " ".join(a.split()[:-1])
        return {"int": C.c_int32, "int64_t": C.c_int64, "uint64_t": C.c_uint64, "float": C.c_float,
                "ctr_stream": C.c_void_p, "int32_t": C.c_int32, "uint32_t": C.c_uint32}[ty]
    for ret, name, args in _header_protos():
        want = [t for t in (ctype(x) for x in args.split(",")) if t is not None]
        res, have = _lib._PROTOS[name]
        assert want == (have or []), name
        assert res == {"int": C.c_int32, "int64_t": C.c_int64, "const char*": C.c_char_p}[ret], name


def test_no_cpu_fallback(lib_built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from deep_learning_b200 import _lib
    from deep_learning_b200.engine import CtrEngine
    from deep_learning_b200.spec import spec_from_hparams
    import ctr_helpers as H
    with pytest.raises(_lib.CtrError):
        CtrEngine(spec_from_hparams("deepfm", H.hp_c1()), 8)
    with pytest.raises(_lib.CtrError):
        _lib.Handle(0)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "deep_learning_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(".py"):
                txt = open(os.path.join(dp, f)).read()
                assert "oracle" not in txt.replace("no oracle", ""), os.path.join(dp, f)


def test_gpu_probes_still_compile(tmp_path):
    """tools/probe_gather4.cu is run by tools/gpu_round.sh on the GPU box: keep it assembling for sm_100a (and keep the
    instruction it is about in the SASS)"""
    import shutil
    import subprocess
    nvcc = "/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else shutil.which("nvcc")
    if not nvcc:
        pytest.skip("no nvcc")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    obj = str(tmp_path / "probe.o")
    r = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-c", os.path.join(root, "tools", "probe_gather4.cu"),
                        "-o", obj], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    sass = subprocess.run([os.path.join(os.path.dirname(nvcc), "cuobjdump"), "-sass", obj], capture_output=True, text=True).stdout
    assert "UTMALDG.2D.GATHER4" in sass


def test_no_undefined_names_in_files_that_only_run_on_the_gpu_box():
    """most of the step drivers, bench.py and the -m gpu tests never execute in the CPU suite: a name used but never
    bound anywhere in its file (a removed import, a typo) would first show on the GPU box.  File-level heuristic."""
    import ast
    import builtins
    import glob
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    files = glob.glob(os.path.join(root, "deep_learning_b200", "**", "*.py"), recursive=True) + \
        glob.glob(os.path.join(root, "tests", "*.py")) + [os.path.join(root, "bench.py"), os.path.join(root, "__graft_entry__.py")]
    assert len(files) > 40
    bad = {}
    for f in files:
        tree = ast.parse(open(f).read())
        bound = set(dir(builtins)) | {"__file__"}
        for n in ast.walk(tree):
            if isinstance(n, (ast.Import, ast.ImportFrom)):
                bound.update((a.asname or a.name).split(".")[0] for a in n.names)
            elif isinstance(n, (ast.FunctionDef, ast.AsyncFunctionDef, ast.Lambda)):
                if not isinstance(n, ast.Lambda):
                    bound.add(n.name)
                a = n.args
                bound.update(x.arg for x in a.args + a.kwonlyargs + a.posonlyargs)
                bound.update(x.arg for x in (a.vararg, a.kwarg) if x)
            elif isinstance(n, ast.ClassDef):
                bound.add(n.name)
            elif isinstance(n, ast.Name) and isinstance(n.ctx, (ast.Store, ast.Del)):
                bound.add(n.id)
            elif isinstance(n, ast.ExceptHandler) and n.name:
                bound.add(n.name)
        used = {n.id for n in ast.walk(tree) if isinstance(n, ast.Name) and isinstance(n.ctx, ast.Load)}
        if used - bound:
            bad[os.path.relpath(f, root)] = sorted(used - bound)
    assert not bad, bad


def test_library_records_the_hash_of_its_sources():
    """the built library carries the SHA-256 of csrc/ + the header + the compiler flags (build.py), readable both from
    the file and through ctr_source_hash()"""
    from deep_learning_b200 import build, _lib
    assert build.library_hash() == build.source_hash()
    assert _lib.load().ctr_source_hash().decode() == build.source_hash()
    assert not build.needs_build()


def test_stale_library_is_refused(monkeypatch):
    """a library built from other sources than the ones beside it does not load (no silent stale binary)"""
    import pytest
    from deep_learning_b200 import build, _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(build, "source_hash", lambda: "0" * 64)
    with pytest.raises(_lib.CtrError, match="built from other sources"):
        _lib.load()
    monkeypatch.setenv("CTR_ALLOW_STALE_LIB", "1")
    assert _lib.load() is not None
