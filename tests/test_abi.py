"""C-ABI surface: the library builds/loads and exports exactly what include/pio_b200.h declares,
and the ctypes prototypes agree with the header on arity.  No compute calls (no GPU needed)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT, PKG_DIR

HEADER = os.path.join(ROOT, "include", "pio_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"typedef struct \{.*?\} \w+;", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(?:int|size_t|long long|const char\*)\s+(pio_\w+)\s*\(([^;]*?)\)\s*;", src, flags=re.S):
        args = m.group(2).strip()
        n = 0 if args in ("", "void") else args.count(",") + 1
        out[m.group(1)] = n
    return out


def _lib_path():
    import importlib.util
    spec = importlib.util.spec_from_file_location("pio_build", os.path.join(PKG_DIR, "build.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    if not os.path.exists(mod.LIB):
        mod.build()
    return mod.LIB


def test_header_declares_something():
    d = _declared()
    assert len(d) >= 35
    assert "pio_cross_attend_fwd" in d and "pio_latent_block_bwd" in d and "pio_image_encode" in d


def test_library_exports_every_declared_symbol():
    handle = ctypes.CDLL(_lib_path())
    for name in _declared():
        assert hasattr(handle, name), f"{name} declared in pio_b200.h but not exported"
    handle.pio_version.restype = ctypes.c_int
    assert handle.pio_version() >= 100


def test_ctypes_prototypes_match_header_arity():
    from perceiver import _native
    d = _declared()
    assert set(_native.EXPORTS) == set(d), set(_native.EXPORTS) ^ set(d)
    for name, (_, args) in _native._SIGNATURES.items():
        assert len(args) == d[name], f"{name}: ctypes has {len(args)} args, header {d[name]}"


def test_struct_layouts_match_header():
    from perceiver import _native
    src = open(HEADER).read()
    for cname, cls in (("pio_attn_weights", _native.AttnWeights), ("pio_mlp_weights", _native.MlpWeights)):
        body = re.search(r"typedef struct \{([^}]*)\} %s;" % cname, src).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = re.findall(r"\*\s*(\w+)", body)
        assert tuple(names) == cls._names, (names, cls._names)
    for cname, cls in (("pio_gemm_desc", _native.GemmDesc), ("pio_attn_desc", _native.AttnDesc)):
        body = re.search(r"typedef struct \{([^}]*)\} %s;" % cname, src).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        names = []
        for stmt in body.split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            decl = re.sub(r"^(const\s+)?\w+\s*\*?", "", stmt, count=1)
            names += [n.strip().lstrip("*").strip() for n in decl.split(",")]
        assert names == [f[0] for f in cls._fields_], (cname, names)


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from perceiver import _native
    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no CPU"):
        _native.lib()
