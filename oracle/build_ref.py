"""Build recipe for the *reference's own* native code, used only as a checker.

TEST INFRASTRUCTURE ONLY.  Nothing here is imported by the product path
(`glorie-slam_b200/`); only tests/, __graft_entry__.build() and the
`--impl reference` / cpu_baseline legs of bench.py may touch `oracle/`.

What it builds (sources are compiled where they lie under /root/reference,
never copied; outputs go to the git-ignored `oracle/_ref/`):

* `lietorch_backends`   <- thirdparty/lietorch/lietorch/src/{lietorch.cpp,lietorch_cpu.cpp,lietorch_gpu.cu}
  needed to import the reference's PyTorch BA (`src/geom/ba.py`) on CPU in
  this container (golden-vector generation, tests/golden/make_golden.py).
* `ref_droid_backends`  <- src/lib/{droid.cpp,droid_kernels.cu,correlation_kernels.cu,altcorr_kernel.cu}
  the reference's CUDA `droid_backends` rebuilt for sm_100.  It cannot run in
  this container (no GPU); the prebuilt .so travels to the GPU box where the
  `-m gpu` parity tests use it as the true `droid_backends.ba` oracle.

Both need one compatibility overload that torch>=2 removed
(`detail::scalar_type(DeprecatedTypeProperties)`, used by
thirdparty/lietorch/lietorch/include/dispatch.h:39); it is force-included
from a generated header, the reference files themselves stay untouched.

Usage:  python oracle/build_ref.py [lietorch] [droid] [pyref]
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("GLORIE_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

SHIM = """#pragma once
#include <ATen/ATen.h>
namespace detail {
inline at::ScalarType scalar_type(const at::DeprecatedTypeProperties& t) { return t.scalarType(); }
}
"""


def _shim_path():
    os.makedirs(OUT, exist_ok=True)
    p = os.path.join(OUT, "torch_compat_shim.h")
    if not os.path.exists(p):
        with open(p, "w") as f:
            f.write(SHIM)
    return p


def _load(name, sources, includes, opt):
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils.cpp_extension import load

    shim = _shim_path()
    bdir = os.path.join(OUT, "build_" + name)
    os.makedirs(bdir, exist_ok=True)
    load(
        name=name,
        sources=sources,
        extra_include_paths=includes,
        extra_cflags=[opt, "-include", shim],
        extra_cuda_cflags=[opt, "-include", shim],
        build_directory=bdir,
        is_python_module=False,
        verbose=False,
    )
    so = os.path.join(bdir, name + ".so")
    dst = os.path.join(OUT, name + ".so")
    if os.path.exists(so):
        import shutil

        shutil.copy2(so, dst)
    return dst


def build_lietorch():
    dst = os.path.join(OUT, "lietorch_backends.so")
    if os.path.exists(dst):
        return dst
    lt = os.path.join(REF, "thirdparty/lietorch/lietorch")
    return _load(
        "lietorch_backends",
        [os.path.join(lt, "src", f) for f in ("lietorch.cpp", "lietorch_gpu.cu", "lietorch_cpu.cpp")],
        [os.path.join(lt, "include"), os.path.join(REF, "thirdparty/eigen")],
        "-O2",
    )


def build_droid():
    dst = os.path.join(OUT, "ref_droid_backends.so")
    if os.path.exists(dst):
        return dst
    lib = os.path.join(REF, "src/lib")
    return _load(
        "ref_droid_backends",
        [os.path.join(lib, f) for f in ("droid.cpp", "droid_kernels.cu", "correlation_kernels.cu", "altcorr_kernel.cu")],
        [os.path.join(REF, "thirdparty/eigen")],
        "-O3",
    )


def stage_python():
    """Copies the handful of pure-Python reference files that `src/geom/ba.py` needs to run -- src/geom/
    and the lietorch package front-end -- next to lietorch_backends.so in the git-ignored oracle/_ref/pyref/,
    so that the GPU box (which has no /root/reference) can time the reference's own PyTorch BA on its host
    cores (bench.py cpu_baseline / --impl reference).  Nothing under oracle/_ref/ is tracked or shipped."""
    import shutil

    dst = os.path.join(OUT, "pyref")
    for sub, names in (("src/geom", ("__init__.py", "ba.py", "chol.py", "projective_ops.py")),
                       ("thirdparty/lietorch/lietorch", ("__init__.py", "groups.py", "group_ops.py", "broadcasting.py"))):
        d = os.path.join(dst, "lietorch" if "lietorch" in sub else sub)
        os.makedirs(d, exist_ok=True)
        for n in names:
            shutil.copy2(os.path.join(REF, sub, n), os.path.join(d, n))
    return dst


def available():
    return os.path.isdir(REF)


if __name__ == "__main__":
    if not available():
        print("reference tree not present at", REF, "- nothing to build")
        sys.exit(0)
    what = sys.argv[1:] or ["lietorch", "droid", "pyref"]
    if "lietorch" in what:
        print("lietorch_backends ->", build_lietorch())
    if "droid" in what:
        print("ref_droid_backends ->", build_droid())
    if "pyref" in what:
        print("python reference ->", stage_python())
