"""The C-ABI library loads and exports every symbol include/cyten_b200.h declares (no GPU needed)."""
import ctypes
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def _declared():
    txt = (ROOT / "include" / "cyten_b200.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(cyb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from cyten_b200 import _lib, build
    build.build()
    lib = ctypes.CDLL(str(_lib.LIB_PATH))
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/cyten_b200.h but not exported"
    assert lib.cyb_version() == 100


def test_python_binding_covers_the_header():
    from cyten_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared()


def test_oracle_covers_the_header():
    from oracle import cabi_numpy
    for n in _declared():
        if n == "cyb_device_info":
            continue
        assert hasattr(cabi_numpy, n), n


def test_product_fails_loudly_without_gpu():
    import pytest
    import torch
    from cyten_b200 import _lib
    from cyten_b200.backend import B200AbelianBackend
    with pytest.raises(_lib.CytenB200Error):
        B200AbelianBackend("cpu")
    if not torch.cuda.is_available():
        with pytest.raises(Exception):
            B200AbelianBackend("cuda:0").lexsort(__import__("numpy").zeros((3, 2), dtype="int64"))
