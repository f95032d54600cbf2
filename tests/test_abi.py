"""CPU test: the C-ABI library loads and exports every symbol include/b200_ggml.h declares (no compute calls)."""
import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols(header: Path) -> set[str]:
    return set(re.findall(r"B200_API\s+[\w\s\*]+?\b(b200_\w+)\s*\(", header.read_text()))


def test_library_exports_every_declared_symbol():
    import koboldcpp_b200.lib as L
    lib = L.load()
    syms = declared_symbols(ROOT / "include" / "b200_ggml.h")
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in b200_ggml.h but not exported"
    assert syms <= set(L._SIGS), syms - set(L._SIGS)  # the Python mirror binds all of them


def test_no_gpu_means_loud_failure_not_fallback():
    import numpy as np
    import pytest

    import koboldcpp_b200.host as H
    import koboldcpp_b200.lib as L
    if L.lib().b200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(L.B200Error):
        H.mul_mat_vec_q(12, np.zeros((1, 144), np.uint8), np.zeros((1, 256), np.float32), 256, 1)


def test_product_does_not_import_the_oracle():
    """The product path (package + csrc) must never reference oracle/ -- only tests, smoke() and bench's CPU legs may."""
    for p in list((ROOT / "koboldcpp_b200").rglob("*.py")) + list((ROOT / "koboldcpp_b200" / "csrc").glob("*")):
        if p.is_file() and p.suffix in (".py", ".cu", ".cuh", ".cpp", ".h"):
            txt = p.read_text()
            assert "ggml_oracle" not in txt and "oracle/" not in txt and "import orc" not in txt, p
