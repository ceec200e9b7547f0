"""ctypes binding of libgba.so (C ABI in include/gba.h) plus a per-device context cache.

There is no fallback: if the library is missing or no CUDA device is usable, importing
callers get a RuntimeError -- the product path never computes on the CPU.
"""
from __future__ import annotations

import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "lib", "libgba.so")

MODE_DROID_CUDA = 0
MODE_GEOM_PY = 1
MODES = {"droid_cuda": MODE_DROID_CUDA, "geom_py": MODE_GEOM_PY}
LAYOUT_INTERLEAVED = 0x100

GET_KX, GET_EPTR, GET_EIDX, GET_EDGE_HV, GET_Q, GET_W, GET_SYSTEM, GET_TILEMASK, GET_GRAM, GET_WATCHDOG = range(1, 11)
GET_CHAIN_TRACE, GET_PERM = 11, 12

_lib = None
_ctx = {}


class Info(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "n_depth_frames", "n_poses", "max_out_degree", "chol_failures", "n_tiles", "n_active_tiles",
        "kernel_launches", "chain_columns", "last_fail_code", "watchdog_trips", "n_segments", "segment_tiles")]


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "libgba.so not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`make -C glorie-slam_b200/csrc`; there is no CPU fallback." % LIB_PATH)
        L = ctypes.CDLL(LIB_PATH)
        vp, ci, cf = ctypes.c_void_p, ctypes.c_int, ctypes.c_float
        L.gba_create.argtypes = [ctypes.POINTER(vp), ci, ci, ci, ci, ci]
        L.gba_destroy.argtypes = [vp]
        L.gba_destroy.restype = None
        L.gba_last_error.argtypes = [vp]
        L.gba_last_error.restype = ctypes.c_char_p
        L.gba_ba.argtypes = [vp] + [vp] * 9 + [ci] * 6 + [cf, cf, ci, ci, ci, vp, vp, vp]
        L.gba_dspo.argtypes = [vp] + [vp] * 12 + [ci] * 4 + [cf, cf, cf, ci, vp]
        L.gba_ba_begin.argtypes = [vp] + [vp] * 9 + [ci] * 5 + [cf, cf, ci, ci, ci, vp, ci, vp, vp, vp, ci, vp]
        L.gba_ba_build.argtypes = [vp]
        L.gba_ba_system.argtypes = [vp, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
        L.gba_ba_solve.argtypes = [vp]
        L.gba_ba_end.argtypes = [vp, vp, vp]
        L.gba_query.argtypes = [vp, ctypes.POINTER(Info)]
        L.gba_get.argtypes = [vp, ci, vp, ctypes.c_size_t]
        L.gba_ba_set_shard.argtypes = [vp, ci, ci, ci]
        L.gba_ba_eliminate.argtypes = [vp]
        L.gba_ba_finish.argtypes = [vp]
        L.gba_ba_system_part.argtypes = [vp, ci, ctypes.POINTER(vp), ctypes.POINTER(ctypes.c_size_t)]
        L.gba_plan.argtypes = [vp, vp, vp, ci, ci, ci, ci, vp]
        for f in ("gba_ba_set_shard", "gba_ba_eliminate", "gba_ba_finish", "gba_ba_system_part", "gba_plan"):
            getattr(L, f).restype = ci
        L.gba_last_depth_frames.argtypes = [vp]
        L.gba_last_depth_frames.restype = ci
        L.gba_edge_blocks_static.argtypes = [vp] * 7 + [ci] * 5 + [vp, vp]
        L.gba_edge_blocks_static.restype = ci
        L.gba_frame_distance.argtypes = [vp] * 5 + [ci, ci, ci, cf, vp, vp]
        L.gba_frame_distance.restype = ci
        L.gba_depth_filter.argtypes = [vp] * 5 + [ci, ci, ci, ci, vp, vp]
        L.gba_depth_filter.restype = ci
        L.gba_align_scale_shift.argtypes = [vp] * 3 + [ci, ci, ci, vp, vp, vp, vp]
        L.gba_align_scale_shift.restype = ci
        L.gba_iproj.argtypes = [vp] * 3 + [ci, ci, ci, vp, vp]
        L.gba_iproj.restype = ci
        L.gba_projmap.argtypes = [vp] * 5 + [ci, ci, ci, vp, vp, vp]
        L.gba_projmap.restype = ci
        L.gba_proximity_scratch_bytes.argtypes = [ci, ci]
        L.gba_proximity_scratch_bytes.restype = ctypes.c_size_t
        L.gba_proximity_select.argtypes = [vp, ci, ci, ci, ci, ci, cf, ci, ci, vp, ctypes.c_size_t, vp, vp, ci, vp, vp]
        L.gba_proximity_select.restype = ci
        L.gba_valid_depth_mask.argtypes = [vp, vp, ci, ci, ci, ci, vp, vp, vp, vp]
        L.gba_valid_depth_mask.restype = ci
        L.gba_mono_filter.argtypes = [vp, vp, vp, vp, ci, ci, ci, cf, vp, vp, ci, ci, vp, vp, vp, vp]
        L.gba_mono_filter.restype = ci
        L.gba_set_timing.argtypes = [vp, ci]
        L.gba_get_timing.argtypes = [vp, ctypes.POINTER(cf), ctypes.POINTER(ci)]
        L.gba_set_timing.restype = ci
        L.gba_get_timing.restype = ci
        for f in ("gba_create", "gba_ba", "gba_dspo", "gba_ba_begin", "gba_ba_build", "gba_ba_system",
                  "gba_ba_solve", "gba_ba_end", "gba_query", "gba_get"):
            getattr(L, f).restype = ci
        _lib = L
    return _lib


class GbaError(RuntimeError):
    pass


def check(ctx, rc):
    if rc != 0:
        msg = lib().gba_last_error(ctx)
        raise GbaError("libgba error %d: %s" % (rc, msg.decode() if msg else "?"))


def _round_up(x, m):
    return ((max(int(x), 1) + m - 1) // m) * m


def context(device_index, n_frames, n_edges, ht, wd):
    """Context for (device, grid size) with room for n_frames / n_edges; grown geometrically.
    Growing frees the old workspaces (cudaFree synchronises the device)."""
    key = (int(device_index), int(ht), int(wd))
    cur = _ctx.get(key)
    if cur is not None and cur[1] >= n_frames and cur[2] >= n_edges:
        return cur[0]
    cap_f = _round_up(max(n_frames, 64, 0 if cur is None else cur[1]), 64)
    cap_e = _round_up(max(n_edges, 256, 0 if cur is None else cur[2]), 256)
    if cur is not None:
        cap_f = max(cap_f, _round_up(n_frames * 3 // 2, 64)) if n_frames > cur[1] else cap_f
        cap_e = max(cap_e, _round_up(n_edges * 3 // 2, 256)) if n_edges > cur[2] else cap_e
        lib().gba_destroy(cur[0])
        del _ctx[key]
    h = ctypes.c_void_p()
    rc = lib().gba_create(ctypes.byref(h), key[0], cap_f, cap_e, key[1], key[2])
    if rc != 0:
        msg = lib().gba_last_error(None)
        raise GbaError("gba_create failed (%d): %s" % (rc, msg.decode() if msg else "?"))
    _ctx[key] = (h, cap_f, cap_e)
    return h


def release_all():
    for h, _, _ in _ctx.values():
        lib().gba_destroy(h)
    _ctx.clear()


STAGES = ("index", "linearize", "pose_blocks", "gram", "schur", "factor", "solve", "backsub", "retract", "dspo")


def set_timing(ctx, enable):
    check(ctx, lib().gba_set_timing(ctx, int(enable)))


def get_timing(ctx):
    """{stage: (milliseconds, kernel launches)} accumulated since the last read (synchronises)."""
    ms = (ctypes.c_float * len(STAGES))()
    ln = (ctypes.c_int * len(STAGES))()
    check(ctx, lib().gba_get_timing(ctx, ms, ln))
    return {s: (float(ms[i]), int(ln[i])) for i, s in enumerate(STAGES)}


def query(ctx):
    info = Info()
    check(ctx, lib().gba_query(ctx, ctypes.byref(info)))
    return info


class _DevBuf:
    """Zero-copy torch view of library-owned device memory (CUDA array interface v2)."""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2}


def system_view(ctx, device, part=0):
    """torch.float64 view of the packed reduced camera system of the last gba_ba_build: part 0 = all of
    it, part 1 = the closure x closure tiles + rhs (the one range the ranks of a sharded solve exchange)."""
    import torch

    ptr, cnt = ctypes.c_void_p(), ctypes.c_size_t()
    check(ctx, lib().gba_ba_system_part(ctx, int(part), ctypes.byref(ptr), ctypes.byref(cnt)))
    return torch.as_tensor(_DevBuf(ptr.value, cnt.value, "<f8"), device=device)
