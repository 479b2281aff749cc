"""Recipe: build the REFERENCE 3DGS rasteriser (GPU oracle for splatt3r_slam_b200/csrc/gs_raster.cu) into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path imports this.

The reference vendors `thirdparty/diff-gaussian-rasterization-modified` (Inria's tile rasteriser: ext.cpp,
rasterize_points.cu, cuda_rasterizer/{forward,backward,rasterizer_impl}.cu, glm under third_party/).  This recipe
compiles those sources WHERE THEY LIE under /root/reference for sm_100a with torch's JIT extension builder (the
reference's setup.py is not run) and copies the resulting pybind module to `oracle/_ref/diffgauss_ref.so`
(git-ignored, travels to the GPU box).  No reference source is copied into the repository.
"""
import os
import shutil
import sys

REF = "/root/reference/thirdparty/diff-gaussian-rasterization-modified"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "diffgauss_ref"


def build(verbose=False):
    if not os.path.isdir(REF):
        return None
    so = os.path.join(OUT, NAME + ".so")
    if os.path.exists(so):
        return so
    import torch  # noqa
    from torch.utils.cpp_extension import load

    bdir = "/tmp/gsref_out"
    os.makedirs(bdir, exist_ok=True)
    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", "4")
    load(
        name=NAME,
        sources=[f"{REF}/ext.cpp", f"{REF}/rasterize_points.cu", f"{REF}/cuda_rasterizer/rasterizer_impl.cu",
                 f"{REF}/cuda_rasterizer/forward.cu", f"{REF}/cuda_rasterizer/backward.cu"],
        extra_include_paths=[REF, f"{REF}/third_party/glm"],
        extra_cflags=["-O3"],
        extra_cuda_cflags=["-O3"],
        build_directory=bdir,
        is_python_module=False,
        verbose=verbose,
    )
    os.makedirs(OUT, exist_ok=True)
    shutil.copy(os.path.join(bdir, NAME + ".so"), so)
    return so


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
