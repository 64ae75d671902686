"""The C-ABI library builds, loads and exports exactly what include/atx_b200.h declares; argument
validation works without a GPU (no compute calls here)."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    from atx import _cabi

    return _cabi.load()


def test_every_declared_symbol_is_exported(lib):
    header = (ROOT / "include" / "atx_b200.h").read_text()
    declared = set(re.findall(r"\b(atx_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/atx_b200.h but not exported"
    from atx import _cabi

    assert declared == set(_cabi.EXPORTS)


def test_header_cites_the_reference_for_each_group():
    header = (ROOT / "include" / "atx_b200.h").read_text()
    for cite in ("llm/serve_llm.py:", "llm/hf_cpu_server.py:", "modeling_llama.py:"):
        assert cite in header


def test_version_and_argument_errors_without_gpu(lib):
    assert lib.atx_version() == 100
    h = ctypes.c_void_p()
    rc = lib.atx_gemm_plan_create(ctypes.byref(h), None, 128, 64, None, 8, 64)
    assert rc == -1 and b"null pointer" in lib.atx_last_error()
    rc = lib.atx_rmsnorm(None, None, None, None, None, 4, 64, 1e-5, None, 0, 0, None)
    assert rc == -1 and b"atx_rmsnorm" in lib.atx_last_error()
    assert lib.atx_paged_attn_decode_workspace_bytes(6, 32, 128, 4) == 6 * 32 * 4 * 130 * 4
    assert lib.atx_paged_attn_decode_workspace_bytes(6, 32, 128, 1) == 0


def test_sass_uses_blackwell_tensor_and_tma_paths():
    import shutil
    import subprocess

    from atx import _cabi

    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", str(_cabi.lib_path())], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass and "UTMALDG" in sass and "LDTM" in sass   # tcgen05.mma / TMA / tcgen05.ld
    assert "sm_100a" in sass


def test_tokenizer_stub_roundtrip_and_template():
    from atx.tokenizer import ByteTokenizer, llama3_chat_string

    t = ByteTokenizer()
    s = "<|begin_of_text|><|start_header_id|>user<|end_header_id|>\n\nhéllo<|eot_id|>"
    ids = t.encode(s)
    assert ids[0] == 128000 and ids[1] == 128006 and ids[-1] == 128009
    assert t.decode(ids) == s and t.decode(ids, skip_special_tokens=True) == "user\n\nhéllo"
    assert t.apply_chat_template([{"role": "user", "content": "x"}]) == llama3_chat_string(
        [{"role": "user", "content": "x"}])
    assert isinstance(t.decode([5000, 127999]), str) and len(t.decode([5000, 127999])) == 2
    small = ByteTokenizer(1024)
    assert max(small.encode(s)) < 1024
