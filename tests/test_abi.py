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


def test_library_exports_backend_ops_and_decoder_symbols():
    """Every B200_API symbol of the other public headers (ggml_backend boundary, graph ops, decoder) is exported too."""
    import koboldcpp_b200.lib as L
    lib = L.load()
    n = 0
    for h in ("b200_ggml_backend.h", "b200_ggml_ops.h", "b200_llama.h"):
        txt = (ROOT / "include" / h).read_text()
        for s in set(re.findall(r"B200_API\s+[\w\s\*]+?\b((?:b200|ggml_backend_b200)_\w+)\s*\(", txt)):
            assert hasattr(lib, s), f"{s} declared in {h} but not exported"
            n += 1
    assert n >= 35, n
    # Route 1: the reference's public CUDA-backend symbol set (ggml/include/ggml-cuda.h:22-44)
    for s in ("ggml_cuda_set_mul_mat_q", "ggml_backend_cuda_init", "ggml_backend_is_cuda", "ggml_backend_cuda_buffer_type",
              "ggml_backend_cuda_split_buffer_type", "ggml_backend_cuda_host_buffer_type", "ggml_backend_cuda_get_device_count",
              "ggml_backend_cuda_get_device_description", "ggml_backend_cuda_get_device_memory", "ggml_backend_cuda_register_host_buffer",
              "ggml_backend_cuda_unregister_host_buffer", "ggml_backend_cuda_reg"):
        assert hasattr(lib, s), s


def test_decoder_refuses_to_run_without_a_gpu():
    import pytest

    import koboldcpp_b200.lib as L
    from koboldcpp_b200.llama import SHAPES, LlamaModel
    if L.lib().b200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(L.B200Error):
        LlamaModel(SHAPES["test-dense"], 256)


def test_no_gpu_means_loud_failure_not_fallback():
    import numpy as np
    import pytest

    import koboldcpp_b200.host as H
    import koboldcpp_b200.lib as L
    if L.lib().b200_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(L.B200Error):
        H.mul_mat_vec_q(12, np.zeros((1, 144), np.uint8), np.zeros((1, 256), np.float32), 256, 1)


def test_static_scratch_entry_point_without_a_gpu():
    """b200_flash_attn_ext_static (no reference counterpart): the stated limit is what the header says, an over-sized shape is
    refused before anything touches a device, and without a GPU a fitting shape fails loudly (no fallback)."""
    import ctypes as C
    import re
    from pathlib import Path

    import pytest

    import koboldcpp_b200.lib as L
    lib = L.lib()
    hdr = (Path(__file__).resolve().parent.parent / "include" / "b200_ggml.h").read_text()
    m = re.search(r"#define\s+B200_FATTN_STATIC_WS_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", hdr)
    assert m and lib.b200_flash_attn_static_workspace_max() == int(m.group(1)) << int(m.group(2)) == 8 << 20

    def args(nh, nhkv, D, nkv, N):
        return L.FattnArgs(256, 256, 256, None, 256, D, N, nh, nhkv, nkv, L.F16, L.F16, 4 * D, 4 * D * N, 2 * D, 2 * D * nkv, 2 * D, 2 * D * nkv, 0, 0.1, 0.0, 0.0)
    big = args(64, 64, 128, 32768, 24)
    assert lib.b200_flash_attn_workspace_bytes(C.byref(big)) > 8 << 20
    assert lib.b200_flash_attn_ext_static(C.byref(big), None) != 0
    assert b"static array holds" in lib.b200_last_error()
    if lib.b200_device_count() > 0:
        pytest.skip("a GPU is present")
    assert lib.b200_flash_attn_ext_static(C.byref(args(32, 8, 128, 4096, 1)), None) != 0


def test_product_does_not_import_the_oracle():
    """The product path (package + csrc) must never reference oracle/ -- only tests, smoke() and bench's CPU legs may."""
    for p in list((ROOT / "koboldcpp_b200").rglob("*.py")) + list((ROOT / "koboldcpp_b200" / "csrc").glob("*")):
        if p.is_file() and p.suffix in (".py", ".cu", ".cuh", ".cpp", ".h"):
            txt = p.read_text()
            assert "ggml_oracle" not in txt and "oracle/" not in txt and "import orc" not in txt, p


def test_abi_restatement_matches_reference_headers(tmp_path):
    """include/b200_ggml_abi.h restates the structs that cross the ggml_backend boundary; where the reference checkout
    exists, every size / offset / enum value is diffed against the real headers (tests/cpp/abi_probe.cpp)."""
    import subprocess

    import pytest
    ref = Path("/root/reference/ggml")
    if not (ref / "include" / "ggml.h").exists():
        pytest.skip("reference checkout not present on this machine")
    src = ROOT / "tests" / "cpp" / "abi_probe.cpp"
    outs = []
    for name, flags in (("ref", ["-DUSE_REF", "-I", str(ref / "include"), "-I", str(ref / "src")]), ("abi", ["-I", str(ROOT / "include")])):
        exe = tmp_path / name
        subprocess.run(["g++", "-std=c++17", "-Wno-invalid-offsetof", *flags, str(src), "-o", str(exe)], check=True)
        outs.append(subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    assert outs[0] == outs[1], [(a, b) for a, b in zip(*outs) if a != b]
