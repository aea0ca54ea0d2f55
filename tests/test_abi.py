"""CPU: the C-ABI library loads, exports every symbol include/b200_vllm.h declares, and its
precondition checks return error codes before any launch (no GPU needed, no compute calls)."""
import ctypes as C
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    from vllm_b200 import _lib, build

    build.build(verbose=False)
    return _lib.load()


def header_symbols() -> set[str]:
    text = (ROOT / "include" / "b200_vllm.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", text))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from vllm_b200 import _lib

    declared = header_symbols()
    assert len(declared) >= 18
    assert declared == set(_lib.SIGNATURES), "ctypes table and header disagree"
    for name in declared:
        assert getattr(lib, name) is not None
    assert b"sm_100a" in lib.b200_version()


def test_struct_layout_matches_header():
    from vllm_b200 import _lib

    text = (ROOT / "include" / "b200_vllm.h").read_text()
    body = text[text.index("typedef struct b200_paged_decode_params"):text.index("} b200_paged_decode_params;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split("{", 1)[1].split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.findall(r"[A-Za-z_][A-Za-z0-9_]*", part)[-1])
    assert names == [f[0] for f in _lib.PagedDecodeParams._fields_]


def test_preconditions_fail_before_launch(lib):
    from vllm_b200 import _lib

    # scaled_mm: K % 16 (scaled_mm_entry.cu:203-216)
    rc = lib.b200_scaled_mm_fp8(None, None, None, None, None, None, 4, 32, 24, 24, 24, 32, 1, 1, _lib.DT_BF16, None, 0, None)
    assert rc == _lib.ERR_INVALID_ARG and b"multiples of 16" in lib.b200_last_error()
    # nvfp4 mm: K % 32 (nvfp4_scaled_mm_kernels.cu:303-310)
    rc = lib.b200_scaled_mm_nvfp4(None, None, None, None, None, None, 4, 32, 48, 32, _lib.DT_BF16, None)
    assert rc == _lib.ERR_INVALID_ARG and b"divisible by 32" in lib.b200_last_error()
    # fp4 quant: n % 16
    rc = lib.b200_scaled_fp4_quant(None, None, None, None, 4, 24, 24, 1, _lib.DT_BF16, None)
    assert rc == _lib.ERR_INVALID_ARG
    # cache write: bad kv dtype / scale stride
    rc = lib.b200_reshape_and_cache_flash(None, None, None, None, None, 1, 8, 128, 16, 1024, 1024, 1, 1, 1,
                                          _lib.DT_BF16, 7, None, None, 0, None)
    assert rc == _lib.ERR_INVALID_ARG and b"kv cache" in lib.b200_last_error()
    # paged decode: block size not a multiple of 16, head size not built
    p = _lib.PagedDecodeParams(num_seqs=1, num_q_heads=8, num_kv_heads=2, head_size=128, block_size=8,
                               q_dtype=_lib.DT_BF16, kv_dtype=_lib.KV_AUTO)
    assert lib.b200_paged_attention_decode(C.byref(p), None) == _lib.ERR_UNSUPPORTED
    p.block_size, p.num_q_heads = 16, 7
    assert lib.b200_paged_attention_decode(C.byref(p), None) == _lib.ERR_INVALID_ARG
    # zero-size calls are no-ops
    assert lib.b200_scaled_mm_fp8(None, None, None, None, None, None, 0, 32, 32, 32, 32, 32, 1, 1, _lib.DT_BF16, None, 0, None) == 0


def test_split_heuristic_and_workspace(lib):
    # deterministic in its arguments only (CUDA-graph safe), >= 1, bounded
    s = lib.b200_paged_decode_pick_splits(64, 8, 4096, 148)
    assert s == lib.b200_paged_decode_pick_splits(64, 8, 4096, 148) and 1 <= s <= 128
    assert lib.b200_paged_decode_pick_splits(4096, 8, 4096, 148) <= 2     # plenty of CTAs already
    assert lib.b200_paged_decode_pick_splits(16, 8, 131072, 148) > 8      # cfg5: long context needs many splits
    assert lib.b200_paged_decode_pick_splits(1, 1, 16, 148) == 1          # one chunk cannot be split
    assert lib.b200_paged_decode_workspace_bytes(64, 32, 128, 1) == 0
    assert lib.b200_paged_decode_workspace_bytes(64, 32, 128, 4) == 64 * 32 * 4 * (128 * 4 + 8)


def test_ops_refuse_cpu_tensors():
    """The product path has no CPU fallback: CPU tensors raise instead of silently computing."""
    from vllm_b200 import ops

    x = torch.zeros(4, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.scaled_fp8_quant(x, use_per_token_if_dynamic=True)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.scaled_fp4_quant(x, torch.ones(1))
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.kv_dtype_code("int4")


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from vllm_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", tmp_path / "nope.so")
    with pytest.raises(RuntimeError, match="no CPU or library fallback"):
        _lib.load()


def test_product_never_imports_oracle():
    for p in (ROOT / "vllm_b200").rglob("*.py"):
        src = p.read_text()
        assert "import oracle" not in src and "from oracle" not in src, f"{p} routes through the oracle"
    for p in (ROOT / "vllm_b200" / "csrc").glob("*"):
        assert "oracle" not in p.read_text()
