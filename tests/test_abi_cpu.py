"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol
include/pokec_b200.h declares; without a device it refuses to work (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from recommendation_system_pokec_b200 import engine
from recommendation_system_pokec_b200.build import build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build()
    return engine.load_library()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "pokec_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(pk_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(engine.ABI_SYMBOLS)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/pokec_b200.h but not exported"


def test_abi_version(lib):
    assert lib.pk_abi_version() == 1


def test_world_struct_layout_matches_header():
    """cabi.PkWorld must list the fields of struct pk_world in header order."""
    hdr = open(os.path.join(ROOT, "include", "pokec_world.h")).read()
    body = hdr[hdr.index("typedef struct pk_world {"):hdr.index("} pk_world;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = re.findall(r"\b(\w+);", body)
    from recommendation_system_pokec_b200.cabi import PkWorld
    assert names == [f[0] for f in PkWorld._fields_]


def test_no_device_fails_loudly(lib):
    if lib.pk_device_count() > 0:
        pytest.skip("a CUDA device is present")
    h = C.c_void_p()
    rc = lib.pk_store_create(0, C.byref(h))
    assert rc == -2 and not h.value          # PK_ERR_CUDA
    assert b"no CPU path" in lib.pk_last_error()
    from tests.conftest import make_test_world
    with pytest.raises(engine.PokecError):
        engine.DeviceRecommender(make_test_world(n=50, pairs=50))


def test_null_arguments_are_errors(lib):
    assert lib.pk_store_create(0, None) == -1
    assert lib.pk_pair_similarity(None, C.c_int64(0), None, None, None) == -1
    assert lib.pk_stat(None, 0) == -1
