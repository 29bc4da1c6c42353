"""CPU-only: the C-ABI library is built, loads, and exports every symbol include/zflacb200.h declares.
No compute calls (there is no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "zflacb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(zflacb200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    syms = declared_symbols()
    for s in ("zflacb200_open", "zflacb200_read", "zflacb200_close", "zflacb200_decode", "zflacb200_batch_open", "zflacb200_batch_read_all", "zflacb200_batch_close"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    from zflac_b200 import _native, build
    build.build_library()
    lib = ctypes.CDLL(_native.LIB_PATH)
    for s in declared_symbols():
        assert hasattr(lib, s), f"{s} declared in include/zflacb200.h but not exported"
    assert sorted(_native.EXPORTS) == declared_symbols()
    assert lib.zflacb200_abi_version() == 2


def test_status_names_match_zig_error_names():
    from zflac_b200 import _native
    lib = _native.load()
    expect = {0: "OK", 1: "InvalidSignature", 2: "InvalidMetadataHeader", 3: "Unimplemented", 4: "InvalidChecksum",
              5: "MissingStreaminfo", 6: "InvalidFrameHeader", 7: "InvalidCodedNumber", 8: "InconsistentParameters",
              9: "InvalidSubframeHeader", 10: "InvalidResidualCodingMethod", 11: "EndOfStream", 12: "OutOfMemory"}
    for k, v in expect.items():
        assert lib.zflacb200_status_name(k).decode() == v
    from oracle import oracle
    for k, v in expect.items():
        assert oracle.STATUS_NAMES[k] == v


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import zflac_b200
    with pytest.raises(zflac_b200.CudaError):
        zflac_b200.decode(b"fLaC" + bytes(64))


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "zflac_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.lower().replace("# oracle", ""), f"{f} mentions the oracle"
