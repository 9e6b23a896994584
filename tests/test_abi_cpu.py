"""The C-ABI library loads without a GPU and exports every symbol include/artspeech_b200.h declares."""
import ctypes
import os
import re

from artspeech_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "artspeech_b200.h")).read()
    return sorted(set(re.findall(r"ASB_API[^;(]*?\b(asb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    names = _declared()
    assert len(names) >= 13
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(handle, n), f"{n} declared in the header but not exported"
    assert sorted(_lib.SIGNATURES) == names, "ctypes SIGNATURES and the header disagree"
    assert _lib.lib().asb_abi_version() == 1
    assert _lib.lib().asb_launch_count() == 0   # nothing has launched: no compute without a GPU


def test_argument_errors_do_not_need_a_gpu():
    lib = _lib.lib()
    rc = lib.asb_p2cp(None, None, None, -1, 50, 50, 0, None)
    assert rc == _lib.ASB_ERR_INVALID and b"asb_p2cp" in lib.asb_last_error_string()
    assert lib.asb_p2cp(None, None, None, 0, 50, 50, 0, None) == _lib.ASB_OK   # empty input is a no-op
    idx = (ctypes.c_int32 * 6)(0, 1, 2, 3, 99, 5)
    assert lib.asb_tract_variables(None, 4, 11, idx, None, None, None, None) == _lib.ASB_ERR_INVALID


def test_library_carries_the_hash_of_the_sources_beside_it():
    """build() proves it compiled this tree: asb_build_hash() == hash(sources + headers + flags); the
    binding refuses a library built from other sources."""
    from artspeech_b200 import build
    assert build.library_hash() == build.source_hash()
    assert _lib.lib().asb_build_hash().decode() == build.source_hash()
    assert len(build.source_hash()) == 16
