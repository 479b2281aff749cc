"""Recipe: build the reference's vendored lietorch (thirdparty/lietorch @ e7df8655) and byte-compile the
reference's Python drivers into oracle/_ref/, so that the reference's OWN `matching.py`, `tracker.py` and
`global_opt.py` can be executed as oracles — here on CPU tensors (lietorch has a CPU backend) to generate
golden fixtures, and on the GPU box on top of this repo's `mast3r_slam_backends` shim.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path imports this.

* `lietorch_backends` (thirdparty/lietorch/setup.py:16-28: lietorch.cpp, lietorch_gpu.cu, lietorch_cpu.cpp +
  vendored eigen) is compiled from the sources where they lie, unmodified, for sm_100a ->
  `oracle/_ref/lietorch_backends.so`.  `lietorch_extras` (:30-44) is not needed by any caller on the path.
* The Python files (lietorch's 4 modules; splatt3r_slam's matching / image / geometry / nonlinear_optimizer /
  tracker / global_opt / config / lietorch_utils) are compiled with `py_compile` to sourceless byte-code files (`.refbc`) under
  `oracle/_ref/refpy/` — binaries of the reference, like the `.so`: no reference SOURCE is copied into the
  repository, and `oracle/_ref/` is git-ignored (it travels to the GPU box, where /root/reference does not exist).
  `tests/refdrivers.py` imports them with stub modules for what the drivers import but the path never touches
  (`splatt3r_slam.splatt3r_utils`, `splatt3r_slam.frame`).
"""
import os
import py_compile
import shutil
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
PYOUT = os.path.join(OUT, "refpy")

LIETORCH_PY = ["__init__.py", "groups.py", "group_ops.py", "broadcasting.py"]
DRIVERS_PY = ["matching.py", "image.py", "geometry.py", "nonlinear_optimizer.py", "tracker.py", "global_opt.py",
              "config.py", "lietorch_utils.py"]


def build_backend(verbose=False):
    so = os.path.join(OUT, "lietorch_backends.so")
    if os.path.exists(so):
        return so
    import torch  # noqa
    from torch.utils.cpp_extension import load
    src = os.path.join(REF, "thirdparty/lietorch")
    bdir = "/tmp/lietorch_build"
    os.makedirs(bdir, exist_ok=True)
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", "4")
    load(name="lietorch_backends",
         sources=[f"{src}/lietorch/src/lietorch.cpp", f"{src}/lietorch/src/lietorch_gpu.cu",
                  f"{src}/lietorch/src/lietorch_cpu.cpp"],
         extra_include_paths=[f"{src}/lietorch/include", f"{src}/eigen"],
         extra_cflags=["-O2"], extra_cuda_cflags=["-O2"], build_directory=bdir, is_python_module=False,
         verbose=verbose)
    os.makedirs(OUT, exist_ok=True)
    shutil.copy(os.path.join(bdir, "lietorch_backends.so"), so)
    return so


def compile_python():
    made = []
    for pkg, srcdir, names in (("lietorch", f"{REF}/thirdparty/lietorch/lietorch", LIETORCH_PY),
                               ("splatt3r_slam", f"{REF}/splatt3r_slam", DRIVERS_PY)):
        d = os.path.join(PYOUT, pkg)
        os.makedirs(d, exist_ok=True)
        for nm in names:
            if pkg == "splatt3r_slam" and nm == "__init__.py":
                continue
            # ".refbc", not ".pyc": the GPU-box snapshot drops *.pyc files; tests/refdrivers.py loads these with
            # importlib's SourcelessFileLoader, which does not care about the extension
            cfile = os.path.join(d, nm[:-3] + ".refbc")
            # dfile: the path recorded in tracebacks — the reference file itself
            py_compile.compile(os.path.join(srcdir, nm), cfile=cfile, dfile=os.path.join(srcdir, nm), doraise=True)
            made.append(cfile)
    return made


def build(verbose=False):
    if not os.path.isdir(REF):
        return None
    return build_backend(verbose), compile_python()


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
