"""Task timeline of the factorisation kernel k_chain (development aid).
    GBA_CHAIN_TRACE=1 python profiles/chain_trace.py [config] > profiles/rNN_chain_trace_<config>.txt
Prints per tile column: when its potrf started / ended, when its critical trsm / update slabs ran, and the
overall share of time the spine spent computing vs waiting."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "glorie-slam_b200"))
os.environ.setdefault("GBA_CHAIN_TRACE", "1")
import numpy as np, torch
import _gba, droid_backends, synth
name = sys.argv[1] if len(sys.argv) > 1 else "E"
g = synth.make_config(name)
t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
d = [t(x) for x in (g.poses, g.disps, g.intrinsics, g.disps_sens, g.targets, g.weights, g.eta, g.ii, g.jj)]
for _ in range(3):
    d[0].copy_(t(g.poses)); d[1].copy_(t(g.disps))
    droid_backends.ba(*d, g.t0, g.t1, 1, g.lm, g.ep, False, False)
torch.cuda.synchronize()
ctx = _gba.context(0, g.n_frames, g.n_edges, g.disps.shape[1], g.disps.shape[2])
buf = np.zeros(1 + 4 * 65536, dtype=np.uint64)
_gba.check(ctx, _gba.lib().gba_get(ctx, 11, buf.ctypes.data_as(ctypes.c_void_p), buf.nbytes))
n = int(buf[0]); rec = buf[1:1 + 4 * n].reshape(n, 4)
role = (rec[:, 0] >> np.uint64(56)).astype(int); kk = ((rec[:, 0] >> np.uint64(40)) & np.uint64(0xffff)).astype(int)
a = ((rec[:, 0] >> np.uint64(28)) & np.uint64(0xfff)).astype(int); b = ((rec[:, 0] >> np.uint64(16)) & np.uint64(0xfff)).astype(int)
t0 = rec[:, 1].min()
take, run, done = [(rec[:, i].astype(np.int64) - int(t0)) / 1e3 for i in (1, 2, 3)]
print("config %s: %d tasks, kernel span %.1f us" % (name, n, done.max()))
names = {0: "potrf", 1: "trsm", 2: "upd-slab", 3: "upd-tile"}
for r in range(4):
    m = role == r
    if m.any():
        print("  %-9s n=%5d  exec mean %.2f us  (min %.2f max %.2f)  wait-for-inputs mean %.2f us" % (
            names[r], m.sum(), (done - run)[m].mean(), (done - run)[m].min(), (done - run)[m].max(), (run - take)[m].mean()))
print("column: potrf start..end | critical trsm end | critical update end | last bulk update of the column ends")
for k in sorted(set(kk[role == 0])):
    p = (role == 0) & (kk == k)
    tr = (role == 1) & (kk == k) & (a == 0)
    up = (role == 2) & (kk == k)
    bulk = (role == 3) & (kk == k)
    f = lambda m, arr: ("%8.1f" % arr[m].max()) if m.any() else "       -"
    print("  %3d: %8.1f .. %8.1f | %s | %s | %s" % (k, run[p].min(), done[p].max(), f(tr, done), f(up, done), f(bulk, done)))
