"""Recipe: build the REFERENCE backend (GPU oracle) into oracle/_ref/.

TEST INFRASTRUCTURE ONLY.  Nothing in the product path imports this.

The reference's `splatt3r_slam/backend` extension does not compile unmodified against
torch 2.11 (4 host-side API drifts).  This recipe stages the three
reference sources + gn.h from /root/reference into a scratch dir under /tmp, applies
exactly those 4 host-line edits with string replacement (device code untouched), adds
`oracle/ref_shim.cu` (which #includes the staged gn_kernels.cu to expose the raw
per-edge Hessian kernels as torch.ops.refshim.*), and builds for sm_100a.  Output:
`oracle/_ref/mast3r_slam_backends_ref.so` (git-ignored, travels to the GPU box).

No reference source is copied into the repository.
"""
import os
import shutil
import sys

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
NAME = "mast3r_slam_backends_ref"


def build(verbose=False):
    if not os.path.isdir(REF):
        return None
    so = os.path.join(OUT, NAME + ".so")
    if os.path.exists(so):
        return so
    import torch  # noqa
    from torch.utils.cpp_extension import load

    stage = "/tmp/refbuild_stage"
    bdir = "/tmp/refbuild_out"
    for d in (stage, bdir):
        os.makedirs(d, exist_ok=True)
    src = os.path.join(REF, "splatt3r_slam/backend")

    def patch(path_in, path_out, subs):
        s = open(path_in).read()
        for a, b, cnt in subs:
            assert s.count(a) == cnt, (path_in, a, s.count(a))
            s = s.replace(a, b)
        open(path_out, "w").write(s)

    patch(f"{src}/src/matching_kernels.cu", f"{stage}/matching_kernels.cu",
          [("D11.type()", "D11.scalar_type()", 1)])
    patch(f"{src}/src/gn_kernels.cu", f"{stage}/gn_kernels.cu",
          [("torch::linalg::linalg_norm(dx, std::optional<c10::Scalar>(), {}, false, {})",
            "at::linalg_norm(dx, std::optional<c10::Scalar>(), {}, false, std::nullopt)", 3)])
    shutil.copy(f"{src}/src/gn.cpp", f"{stage}/gn.cpp")
    shutil.copy(f"{src}/include/gn.h", f"{stage}/gn.h")
    shutil.copy(os.path.join(HERE, "ref_shim.cu"), f"{stage}/ref_shim.cu")

    os.environ["TORCH_CUDA_ARCH_LIST"] = "10.0a"
    os.environ.setdefault("MAX_JOBS", "4")
    load(
        name=NAME,
        # ref_shim.cu #includes gn_kernels.cu, so gn_kernels.cu is not listed itself
        sources=[f"{stage}/gn.cpp", f"{stage}/ref_shim.cu", f"{stage}/matching_kernels.cu"],
        extra_include_paths=[stage, f"{REF}/thirdparty/eigen"],
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
