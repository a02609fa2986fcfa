"""CPU: the C-ABI library loads without a GPU and exports every symbol include/b2c.h declares;
argument errors come back as codes + messages (never exceptions, never exit)."""
import ctypes
import os
import re

from ray_lightning_accelerators_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "b2c.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2c_[a-z_0-9]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"libb2c.so does not export {n}"
        assert n in _lib.SIGNATURES, f"_lib.py does not bind {n}"
    assert sorted(_lib.SIGNATURES) == names


def test_version_and_enums_match_header():
    lib = _lib.load()
    text = open(os.path.join(ROOT, "include", "b2c.h")).read()
    assert lib.b2c_version() == int(re.search(r"#define B2C_VERSION (\d+)", text).group(1))
    assert _lib.MAX_WORLD == int(re.search(r"#define B2C_MAX_WORLD (\d+)", text).group(1))
    assert _lib.MAX_BLOCKS == int(re.search(r"#define B2C_MAX_BLOCKS (\d+)", text).group(1))
    assert _lib.EXPORT_BLOB_BYTES == int(re.search(r"#define B2C_EXPORT_BLOB_BYTES (\d+)", text).group(1))
    for name in ("CROSSOVER_BYTES", "MAX_BLOCKS_ONESHOT", "MAX_BLOCKS_TWOSHOT", "THREADS", "TIMEOUT_MS",
                 "TWOSHOT_VARIANT", "MAX_BLOCKS_BCAST", "ZEROCOPY", "ZEROCOPY_MIN_BYTES"):
        v = int(re.search(rf"B2C_PARAM_{name} = (\d+)", text).group(1))
        assert getattr(_lib, "PARAM_" + name) == v


def test_argument_errors_without_a_gpu():
    lib = _lib.load()
    h = ctypes.c_void_p()
    assert lib.b2c_comm_create(0, 0, 0, 1 << 20, ctypes.byref(h)) == _lib.E_INVAL
    assert b"world=0" in lib.b2c_last_error()
    assert lib.b2c_comm_create(5, 2, 0, 1 << 20, ctypes.byref(h)) == _lib.E_INVAL
    assert lib.b2c_comm_create(0, 2, 0, 16, ctypes.byref(h)) == _lib.E_INVAL
    assert lib.b2c_allreduce(None, None, 0, 0, 0, ctypes.c_float(1.0), 0, None) == _lib.E_INVAL
    assert lib.b2c_set_param(None, 0, 1) == _lib.E_INVAL
    assert lib.b2c_comm_destroy(None) == 0


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "ray_lightning_accelerators_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("oracle/", "").lower() or f in (), f"{f} mentions the oracle"
