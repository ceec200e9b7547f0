"""CPU: libgba.so loads and exports every entry point include/gba.h declares (no compute)."""
import ctypes
import os
import re

from helpers import ROOT


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "gba.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gba_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_the_boundary():
    syms = declared_symbols()
    for s in ("gba_create", "gba_destroy", "gba_last_error", "gba_ba", "gba_dspo", "gba_ba_begin", "gba_ba_build",
              "gba_ba_system", "gba_ba_solve", "gba_ba_end", "gba_query", "gba_get"):
        assert s in syms


def test_library_exports_every_declared_symbol():
    path = os.path.join(ROOT, "glorie-slam_b200", "lib", "libgba.so")
    if not os.path.exists(path):
        import __graft_entry__

        __graft_entry__.build()
    L = ctypes.CDLL(path)
    for s in declared_symbols():
        assert hasattr(L, s), "libgba.so does not export " + s


def test_no_cpu_fallback_without_device():
    """On a box without a GPU the entry point must fail loudly, not compute on the host."""
    import torch

    if torch.cuda.is_available():
        return
    import _gba

    h = ctypes.c_void_p()
    rc = _gba.lib().gba_create(ctypes.byref(h), 0, 8, 24, 48, 64)
    assert rc != 0 and not h
    assert b"CUDA" in _gba.lib().gba_last_error(None)


def test_product_path_does_not_import_oracle():
    pkg = os.path.join(ROOT, "glorie-slam_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                assert "import oracle" not in txt and "from oracle" not in txt and "liboracle" not in txt, f


def test_python_mirror_has_the_reference_call_surface():
    """droid_backends.py mirrors the reference extension's exports on the path (src/lib/droid.cpp:241-245):
    same names, same positional argument counts."""
    import inspect

    import common
    import droid_backends

    want = {"ba": 16, "frame_distance": 6, "projmap": 5, "depth_filter": 5, "iproj": 3}
    for name, n in want.items():
        fn = getattr(droid_backends, name)
        pos = [p for p in inspect.signature(fn).parameters.values() if not p.name.startswith("_")]
        assert len(pos) == n, (name, len(pos))
    assert len(inspect.signature(droid_backends.ba_interleaved).parameters) == 16
    assert list(inspect.signature(common.align_scale_and_shift).parameters) == ["prediction", "target", "weights"]
