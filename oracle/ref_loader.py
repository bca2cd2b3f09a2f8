"""TEST INFRASTRUCTURE — loader for the UNMODIFIED reference (never shipped, never on the product path).

Imports ``/root/reference/src/replay_no_ros.py`` in-process with stub ``matplotlib`` / ``dt_apriltags`` /
``cv2`` modules (the reference imports them at top level, ``replay_no_ros.py:5-10``; none is needed by the
EKF arithmetic), and AST-extracts the pure functions of ``src/EKF-SLAM.py`` (``:12-13, :29-117``), whose
module body cannot be imported.

Only ``oracle/gen_golden.py`` and the ``reference``-marked CPU tests use this; ``/root/reference`` does
not exist on the GPU box, so nothing under ``-m gpu``, ``smoke()`` or ``bench.py`` may call it.
"""
from __future__ import annotations

import ast
import importlib.util
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("EKF_REFERENCE_ROOT", "/root/reference")
# build outputs of oracle/build_ref.py: the reference's own code byte-compiled (travels to the GPU box)
REF_BUILD_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def source_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "src", "replay_no_ros.py"))


def compiled_available() -> bool:
    """oracle/_ref holds bytecode of THIS interpreter version (a .pyc of another version cannot be loaded)."""
    pyc = os.path.join(REF_BUILD_DIR, "replay_no_ros.bytecode")
    if not (os.path.isfile(pyc) and os.path.isfile(os.path.join(REF_BUILD_DIR, "ekf_slam_functions.bytecode"))):
        return False
    with open(pyc, "rb") as fh:
        return fh.read(4) == importlib.util.MAGIC_NUMBER


def reference_available() -> bool:
    return source_available() or compiled_available()


def reference_kind() -> str:
    """"source": /root/reference is here; "compiled": oracle/_ref bytecode; "absent"."""
    return "source" if source_available() else ("compiled" if compiled_available() else "absent")


def _stub(name: str, **attrs) -> types.ModuleType:
    mod = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(mod, k, v)
    return mod


class _Anything:
    """Placeholder for matplotlib / detector classes the hot path never touches."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return self

    def __getattr__(self, item):
        return _Anything()


def _install_stubs() -> list:
    added = []

    def put(name, mod):
        if name not in sys.modules:
            sys.modules[name] = mod
            added.append(name)

    try:
        import matplotlib  # noqa: F401
    except Exception:
        mpl = _stub("matplotlib", rcParams={})
        plt = _stub("matplotlib.pyplot", pause=lambda *a, **k: None, ion=lambda: None,
                    subplots=lambda *a, **k: (_Anything(), _Anything()))
        patches = _stub("matplotlib.patches", Polygon=_Anything, PathPatch=_Anything, Ellipse=_Anything)
        path = _stub("matplotlib.path", Path=_Anything)
        mpl.pyplot, mpl.patches, mpl.path = plt, patches, path
        put("matplotlib", mpl)
        put("matplotlib.pyplot", plt)
        put("matplotlib.patches", patches)
        put("matplotlib.path", path)
    try:
        import dt_apriltags  # noqa: F401
    except Exception:
        put("dt_apriltags", _stub("dt_apriltags", Detector=_Anything))
    try:
        import cv2  # noqa: F401
    except Exception:
        put("cv2", _stub("cv2", imread=lambda *a, **k: None, IMREAD_GRAYSCALE=0))
    return added


_cache = {}


def load_replay_module():
    """The reference's ``replay_no_ros`` module object, loaded unmodified."""
    if "replay" in _cache:
        return _cache["replay"]
    if not reference_available():
        raise FileNotFoundError(f"reference not found under {REFERENCE_ROOT} and no oracle/_ref build")
    _install_stubs()
    if source_available():
        path = os.path.join(REFERENCE_ROOT, "src", "replay_no_ros.py")
        spec = importlib.util.spec_from_file_location("_ref_replay_no_ros", path)
    else:
        from importlib.machinery import SourcelessFileLoader
        path = os.path.join(REF_BUILD_DIR, "replay_no_ros.bytecode")
        spec = importlib.util.spec_from_loader("_ref_replay_no_ros", SourcelessFileLoader("_ref_replay_no_ros", path))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    _cache["replay"] = mod
    return mod


def load_ekf3_functions():
    """``predict`` / ``update`` / ``delta_phi`` / ``displacement`` + noise constants of ``src/EKF-SLAM.py``.

    The file is a broken script (import-time Detector, dead main loop); only the function definitions and
    the two noise assignments (``:12-13``) are executed.
    """
    if "ekf3" in _cache:
        return _cache["ekf3"]
    import numpy as np
    from typing import Tuple

    ns = {"np": np, "Tuple": Tuple}
    if source_available():
        path = os.path.join(REFERENCE_ROOT, "src", "EKF-SLAM.py")
        with open(path, "r") as fh:
            tree = ast.parse(fh.read(), filename=path)
        keep = []
        for node in tree.body:
            if isinstance(node, ast.FunctionDef) and node.name in ("predict", "update", "delta_phi", "displacement"):
                keep.append(node)
            elif isinstance(node, ast.Assign) and any(
                isinstance(t, ast.Name) and t.id in ("motion_noise", "observation_noise") for t in node.targets
            ):
                keep.append(node)
        exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    else:
        import marshal
        with open(os.path.join(REF_BUILD_DIR, "ekf_slam_functions.bytecode"), "rb") as fh:
            fh.read(16)
            exec(marshal.load(fh), ns)
    out = types.SimpleNamespace(**{k: ns[k] for k in
                                   ("predict", "update", "delta_phi", "displacement",
                                    "motion_noise", "observation_noise")})
    _cache["ekf3"] = out
    return out


def run_reference_replay(events_dir: str, detections_by_frame: dict):
    """Drive the reference's ``replay()`` UNCHANGED (``replay_no_ros.py:66-248``) on a decoded directory.

    Module globals that touch images / plots are rebound; ``EKF_pose_estimation`` is wrapped
    only to record what it was called with and what it returned. Returns the list of per-step records.
    """
    import builtins
    import copy

    import numpy as np

    mod = load_replay_module()
    real = mod.EKF_pose_estimation
    log = []

    def spy(ang, lin, mean, cov, dt, pairs, tag_index):
        n_det = sum(len(p[1]) for p in pairs)
        mean_out, cov_out, tags = real(ang, lin, mean, cov, dt, pairs, tag_index)
        log.append(dict(ang=float(ang), lin=float(lin), n_det=n_det,
                        mean=np.array(mean_out, dtype=np.float64, copy=True),
                        cov=np.array(cov_out, dtype=np.float64, copy=True),
                        tag_index=copy.deepcopy(dict(tag_index)),
                        order=[int(k) for k in tags.keys()],
                        tags={int(k): [float(x) for x in v] for k, v in tags.items()}))
        return mean_out, cov_out, tags

    saved = {k: getattr(mod, k) for k in
             ("EKF_pose_estimation", "load_grayscale", "detect_tags", "visualize_bounding_boxes", "plot_path")}
    saved_print = builtins.print
    try:
        mod.EKF_pose_estimation = spy
        mod.load_grayscale = os.path.basename
        mod.detect_tags = lambda img, cam: detections_by_frame.get(img, [])
        mod.visualize_bounding_boxes = lambda *a, **k: None
        mod.plot_path = lambda *a, **k: None
        mod.plt.pause = lambda *a, **k: None
        mod.print = lambda *a, **k: None  # swallow ">>>> i" (:198) without touching builtins
        mod.replay(events_dir)
    finally:
        for k, v in saved.items():
            setattr(mod, k, v)
        if hasattr(mod, "print"):
            del mod.print
        builtins.print = saved_print
    return log
