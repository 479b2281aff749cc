"""SURVEY.md 8(f) rank 4 — can the reference's `main.py --no-viz` loop be brought up headless in this image?

This probe imports what `main.py` imports, in order, from /root/reference (read-only), with the two stand-ins the
task allows: `splatt3r_slam.visualization` replaced by a module that only defines the `WindowMsg` dataclass
(visualization.py:48-55: no graphics library is imported, loaded or imitated) and `pyrealsense2` by an empty module;
`lietorch` comes from oracle/_ref (oracle/build_lietorch.py) and `mast3r_slam_backends` is this repo's shim.  It
records, per import, success or the exact missing package, and writes the log to profiles/r02_f4_import_log.txt.
Nothing is installed, nothing outside the repository is written.

    python tools/f4_probe.py
"""
import dataclasses
import importlib
import os
import sys
import traceback
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

lines = []


def log(s):
    print(s)
    lines.append(s)


def main():
    if not os.path.isdir(REF):
        log("no /root/reference on this machine: nothing to probe")
        return
    import refdrivers
    if refdrivers.available():  # lietorch from the reference's own sources
        import importlib.util
        spec = importlib.util.spec_from_file_location("lietorch_backends", os.path.join(refdrivers.REF_DIR, "lietorch_backends.so"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        sys.modules["lietorch_backends"] = mod
        sys.path.insert(0, os.path.join(REF, "thirdparty", "lietorch"))
        log("lietorch: built from thirdparty/lietorch by oracle/build_lietorch.py -> available")
    else:
        log("lietorch: oracle/_ref/lietorch_backends.so missing (run oracle/build_lietorch.py)")
    sys.path.insert(0, REF)
    # stand-ins the task allows
    viz = types.ModuleType("splatt3r_slam.visualization")

    @dataclasses.dataclass
    class WindowMsg:  # visualization.py:48-55
        is_terminated: bool = False
        is_paused: bool = False
        next: bool = False
        C_conf_threshold: float = 1.5
    viz.WindowMsg = WindowMsg
    viz.run_visualization = lambda *a, **k: None
    sys.modules["pyrealsense2"] = types.ModuleType("pyrealsense2")
    import splatt3r_slam_b200.mast3r_slam_backends as shim
    sys.modules["mast3r_slam_backends"] = shim
    log("mast3r_slam_backends: this repo's shim (libb200slam.so loaded)")

    # every top-level package main.py and the splatt3r_slam modules import, in main.py's order
    third_party = ["yaml", "tqdm", "cv2", "PIL", "einops", "numpy", "torch", "lietorch", "natsort", "plyfile", "matplotlib",
                   "evo", "lightning", "lpips", "omegaconf", "huggingface_hub", "roma", "trimesh", "scipy", "asmk", "faiss",
                   "diff_gaussian_rasterization", "in3d", "moderngl", "moderngl_window", "imgui"]
    log("\n-- third-party packages the loop needs (find_spec) --")
    missing = []
    for m in third_party:
        try:
            ok = importlib.util.find_spec(m) is not None
        except Exception:  # noqa: BLE001
            ok = False
        log(f"  {'ok     ' if ok else 'MISSING'}  {m}")
        if not ok:
            missing.append(m)
    log("\n-- reference modules, imported for real --")
    order = ["splatt3r_slam.config", "splatt3r_slam.lietorch_utils", "splatt3r_slam.geometry", "splatt3r_slam.image",
             "splatt3r_slam.nonlinear_optimizer", "splatt3r_slam.matching", "splatt3r_slam.splatt3r_utils",
             "splatt3r_slam.frame", "splatt3r_slam.tracker", "splatt3r_slam.global_opt", "splatt3r_slam.retrieval_database",
             "splatt3r_slam.dataloader", "splatt3r_slam.evaluate", "splatt3r_slam.multiprocess_utils"]
    sys.modules["splatt3r_slam.visualization"] = viz
    for m in order:
        try:
            importlib.import_module(m)
            log(f"  ok       {m}")
        except Exception as exc:  # noqa: BLE001
            last = traceback.extract_tb(exc.__traceback__)[-1]
            log(f"  FAILED   {m}: {type(exc).__name__}: {exc}   [{os.path.relpath(last.filename, REF) if last.filename.startswith(REF) else last.filename}:{last.lineno}]")
    log("\n-- verdict --")
    graphics = {"in3d", "moderngl", "moderngl_window", "imgui"}
    blockers = [m for m in missing if m not in graphics]
    log("graphics-only packages (not needed with --no-viz and the WindowMsg stand-in): "
        + ", ".join(sorted(graphics & set(missing))))
    log("packages that block a headless run and cannot be installed here (no network, not in the wheelhouse): "
        + (", ".join(blockers) if blockers else "none"))
    out = os.path.join(ROOT, "profiles", "r02_f4_import_log.txt")
    open(out, "w").write("\n".join(lines) + "\n")
    print("written", out)


if __name__ == "__main__":
    main()
