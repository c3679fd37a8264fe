"""libacx.so builds, loads, and exports every symbol include/acx.h declares (no compute calls)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    with open(os.path.join(ROOT, "include", "acx.h")) as f:
        src = re.sub(r"/\*.*?\*/", "", f.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(acx_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as entry

    entry.build()
    from acx import _lib

    L = ctypes.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 6
    for name in names:
        assert hasattr(L, name), f"{name} declared in include/acx.h but not exported"
    # the loader binds exactly the declared set
    assert sorted(_lib.exported_symbols()) == names


def test_version_and_error_text():
    from acx import _lib

    assert _lib.lib().acx_version() >= 100
    assert isinstance(_lib.lib().acx_last_error(), bytes)


def test_bad_arguments_are_rejected_without_a_device():
    from acx import _lib

    L = _lib.lib()
    assert L.acx_move_batch(0, 0, 0, 0, 0, 4, 0, 1, 0) == -1  # m out of range
    assert L.acx_move_batch(0, 0, 0, 0, 0, 4, 36, 1, 0) == -1  # null pointers
    assert b"acx_move_batch" in L.acx_last_error()
