"""ctypes front-end of the CPU oracle (oracle/ba_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / `--impl reference` legs of bench.py.  The product path under
glorie-slam_b200/ never imports this module.

`ba(...)` mirrors the reference call surface
`droid_backends.ba(poses, disps, intrinsics, disps_sens, targets, weights, eta,
ii, jj, t0, t1, iterations, lm, ep, motion_only, depth_only)`
(reference src/lib/droid.cpp:89-119) on numpy arrays, and returns the updated
state instead of mutating the inputs.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from types import SimpleNamespace

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
MODE_DROID_CUDA = 0
MODE_GEOM_PY = 1

_lib = None


def build(force=False):
    so = os.path.join(HERE, "liboracle.so")
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(
        os.path.getmtime(os.path.join(HERE, f)) for f in ("ba_oracle.c", "ba_oracle_impl.h")
    ):
        subprocess.check_call(["make", "-C", HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.oracle_num_threads.restype = ctypes.c_int
    return _lib


def num_threads():
    return int(lib().oracle_num_threads())


def set_threads(n):
    """OpenMP threads of the oracle (torchrun exports OMP_NUM_THREADS=1 to its workers)"""
    lib().oracle_set_threads(ctypes.c_int(int(n)))
    return num_threads()


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def ba(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, t0, t1, iterations,
       lm, ep, motion_only=False, depth_only=False, mode=MODE_DROID_CUDA, precision="f32", debug=False):
    """Run the oracle BA.  Returns a namespace with poses, disps (updated copies, in the
    requested precision), dx [P,6], dz [K,HW], kx [K], fail count and (debug=True) the first
    iteration's Hs [E,4,6,6], vs [E,2,6], C, w [K,HW], M [6P,6P] (= A - S before the
    after-Schur damping), rhs [6P]."""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_ba_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    E = int(ii.shape[0])
    HW = ht * wd
    P = t1 - t0
    poses_o = np.ascontiguousarray(poses, dtype=real).copy()
    disps_o = np.ascontiguousarray(disps, dtype=real).copy()
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    intr, sens, tg, wg, et = f32(intrinsics), f32(disps_sens), f32(targets), f32(weights), f32(eta)
    ii64 = np.ascontiguousarray(ii, dtype=np.int64)
    jj64 = np.ascontiguousarray(jj, dtype=np.int64)
    dx = np.zeros((P, 6), dtype=np.float64)
    dz = np.zeros((N, HW), dtype=np.float64)
    kx = np.zeros((N,), dtype=np.int64)
    K = ctypes.c_int(0)
    fail = ctypes.c_int(0)
    dbg = SimpleNamespace(Hs=None, vs=None, C=None, w=None, M=None, rhs=None)
    if debug:
        dbg.Hs = np.zeros((max(E, 1), 4, 6, 6), dtype=np.float64)
        dbg.vs = np.zeros((max(E, 1), 2, 6), dtype=np.float64)
        dbg.C = np.zeros((N, HW), dtype=np.float64)
        dbg.w = np.zeros((N, HW), dtype=np.float64)
        dbg.M = np.zeros((6 * P, 6 * P), dtype=np.float64)
        dbg.rhs = np.zeros((6 * P,), dtype=np.float64)
    rc = fn(_p(poses_o), _p(disps_o), _p(intr), _p(sens), _p(tg), _p(wg), _p(et), _p(ii64), _p(jj64),
            ctypes.c_int(N), ctypes.c_int(E), ctypes.c_int(ht), ctypes.c_int(wd), ctypes.c_int(t0),
            ctypes.c_int(t1), ctypes.c_int(iterations), ctypes.c_float(lm), ctypes.c_float(ep),
            ctypes.c_int(int(motion_only)), ctypes.c_int(int(depth_only)), ctypes.c_int(mode),
            _p(dx), _p(dz), _p(kx), ctypes.byref(K), ctypes.byref(fail),
            _p(dbg.Hs), _p(dbg.vs), _p(dbg.C), _p(dbg.w), _p(dbg.M), _p(dbg.rhs))
    if rc != 0:
        raise RuntimeError("oracle_ba failed with code %d" % rc)
    k = K.value
    out = SimpleNamespace(poses=poses_o, disps=disps_o, dx=dx, dz=dz[:k], kx=kx[:k].copy(), fail=fail.value)
    if debug:
        out.Hs, out.vs, out.M, out.rhs = dbg.Hs[:E], dbg.vs[:E], dbg.M, dbg.rhs
        out.C, out.w = dbg.C[:k], dbg.w[:k]
    return out


def dspo(poses, disps, intrinsics, targets, weights, eta, ii, jj, mono_disps, scales, shifts,
         valid_depth_mask, ignore_frames, lm, ep, alpha, interleaved=True, precision="f32"):
    """One BA_with_scale_shift iteration (reference src/geom/ba.py:127-216) on numpy arrays.
    targets/weights: [E,ht,wd,2] when interleaved (the reference layout for this entry point)
    or planar [E,2,ht,wd].  Returns namespace(disps, scales, shifts, dz, dwq, kx-count M, fail)."""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_dspo_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    E = int(ii.shape[0])
    HW = ht * wd
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    poses_r = np.ascontiguousarray(poses, dtype=real)
    disps_o = np.ascontiguousarray(disps, dtype=real).copy()
    sc = np.ascontiguousarray(scales, dtype=real).copy()
    sh = np.ascontiguousarray(shifts, dtype=real).copy()
    intr, tg, wg, et, mono = f32(intrinsics), f32(targets), f32(weights), f32(eta), f32(mono_disps)
    mask = np.ascontiguousarray(valid_depth_mask, dtype=np.uint8)
    ii64 = np.ascontiguousarray(ii, dtype=np.int64)
    jj64 = np.ascontiguousarray(jj, dtype=np.int64)
    dz = np.zeros((N, HW), dtype=np.float64)
    dwq = np.zeros((N, 2), dtype=np.float64)
    fail = ctypes.c_int(0)
    M = fn(_p(poses_r), _p(disps_o), _p(intr), _p(tg), _p(wg), _p(et), _p(ii64), _p(jj64), _p(mono),
           _p(sc), _p(sh), _p(mask), ctypes.c_int(N), ctypes.c_int(E), ctypes.c_int(ht), ctypes.c_int(wd),
           ctypes.c_int(ignore_frames), ctypes.c_float(lm), ctypes.c_float(ep), ctypes.c_float(alpha),
           ctypes.c_int(int(interleaved)), _p(dz), _p(dwq), ctypes.byref(fail))
    return SimpleNamespace(disps=disps_o, scales=sc, shifts=sh, dz=dz[:M], dwq=dwq[:M], M=M, fail=fail.value)


def frame_distance(poses, disps, intrinsics, ii, jj, beta=0.3, precision="f32"):
    """reference droid_backends.frame_distance (src/lib/droid_kernels.cu:518-657,1441-1463) on numpy arrays"""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_frame_distance_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    po = np.ascontiguousarray(poses, dtype=real)
    di = np.ascontiguousarray(disps, dtype=real)
    intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
    ii64 = np.ascontiguousarray(ii, dtype=np.int64)
    jj64 = np.ascontiguousarray(jj, dtype=np.int64)
    out = np.zeros((ii64.shape[0],), dtype=real)
    fn(_p(po), _p(di), _p(intr), _p(ii64), _p(jj64), ctypes.c_int(ii64.shape[0]), ctypes.c_int(ht), ctypes.c_int(wd),
       ctypes.c_float(beta), _p(out))
    return out


def depth_filter(poses, disps, intrinsics, inds, thresh, precision="f32"):
    """reference droid_backends.depth_filter (src/lib/droid_kernels.cu:661-775,1494-1518) -> counter [n,ht,wd] f32"""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_depth_filter_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    po, di = np.ascontiguousarray(poses, dtype=real), np.ascontiguousarray(disps, dtype=real)
    intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
    ix = np.ascontiguousarray(inds, dtype=np.int64)
    th = np.ascontiguousarray(thresh, dtype=np.float32)
    out = np.zeros((ix.shape[0], ht, wd), dtype=np.float32)
    fn(_p(po), _p(di), _p(intr), _p(ix), _p(th), ctypes.c_int(N), ctypes.c_int(ix.shape[0]), ctypes.c_int(ht),
       ctypes.c_int(wd), _p(out))
    return out


def align_scale_and_shift(prediction, target, weights, precision="f32"):
    """reference src/utils/common.py:402-438 -> (scale [n], shift [n], avg_error [n])"""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_align_scale_shift_" + precision)
    fn.restype = ctypes.c_int
    n, ht, wd = prediction.shape
    p_, t_, w_ = (np.ascontiguousarray(a, dtype=real) for a in (prediction, target, weights))
    s, q, e = (np.zeros((n,), dtype=real) for _ in range(3))
    fn(_p(p_), _p(t_), _p(w_), ctypes.c_int(n), ctypes.c_int(ht * wd), _p(s), _p(q), _p(e))
    return s, q, e


def iproj(poses, disps, intrinsics, precision="f32"):
    """reference droid_backends.iproj (src/lib/droid_kernels.cu:779-850,1520-1545) -> points [N,ht,wd,3]"""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_iproj_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    po, di = np.ascontiguousarray(poses, dtype=real), np.ascontiguousarray(disps, dtype=real)
    intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
    out = np.zeros((N, ht, wd, 3), dtype=real)
    fn(_p(po), _p(di), _p(intr), ctypes.c_int(N), ctypes.c_int(ht), ctypes.c_int(wd), _p(out))
    return out


def projmap(poses, disps, intrinsics, ii, jj, precision="f32"):
    """reference droid_backends.projmap (src/lib/droid_kernels.cu:427-516,1466-1492) -> coords [n,ht,wd,3], valid [n,ht,wd,1]"""
    real = np.float32 if precision == "f32" else np.float64
    fn = getattr(lib(), "oracle_projmap_" + precision)
    fn.restype = ctypes.c_int
    N, ht, wd = disps.shape
    po, di = np.ascontiguousarray(poses, dtype=real), np.ascontiguousarray(disps, dtype=real)
    intr = np.ascontiguousarray(intrinsics, dtype=np.float32)
    ii64, jj64 = np.ascontiguousarray(ii, dtype=np.int64), np.ascontiguousarray(jj, dtype=np.int64)
    n = ii64.shape[0]
    coords, valid = np.zeros((n, ht, wd, 3), dtype=real), np.zeros((n, ht, wd, 1), dtype=real)
    fn(_p(po), _p(di), _p(intr), _p(ii64), _p(jj64), ctypes.c_int(n), ctypes.c_int(ht), ctypes.c_int(wd), _p(coords), _p(valid))
    return coords, valid
