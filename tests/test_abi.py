"""CPU tests (not gpu): the C-ABI library builds, loads and exports every symbol that
include/la_b200.h declares.  No compute call is made (there is no GPU here)."""
import re
from pathlib import Path

import linalg_b200
from linalg_b200 import ffi

ROOT = Path(__file__).resolve().parent.parent


def _header_symbols():
    text = (ROOT / "include" / "la_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    syms = set(re.findall(r"\b((?:cblas_|openblas_|la_b200_)\w+|[sdz]\w+_|ilaver_)\s*\(", text))
    # expand the LA_B200_DECL_DEV macro
    body = re.search(r"#define LA_B200_DECL_DEV\(X, T\)(.*?)LA_B200_DECL_DEV\(s, float\)", text,
                     re.S).group(1)
    for name in re.findall(r"la_b200_##X##(\w+)", body):
        for x in "sdz":
            syms.add(f"la_b200_{x}{name}")
    return {s for s in syms if "##" not in s and s != "la_b200_"}


def test_library_is_built_and_loads():
    assert linalg_b200.LIB_PATH.exists(), "run python -m linalg_b200.build"
    assert "sm_100a" in linalg_b200.version()


def test_every_declared_symbol_is_exported():
    lib = linalg_b200.lib()
    declared = _header_symbols()
    assert len(declared) > 100
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, f"declared in la_b200.h but not exported: {missing}"
    # and the Python list used by other tests agrees with the header
    assert set(ffi.declared_symbols()) <= declared


def test_product_never_imports_oracle():
    for py in (ROOT / "linalg_b200").rglob("*.py"):
        src = py.read_text()
        assert "import oracle" not in src and "from oracle" not in src, py
