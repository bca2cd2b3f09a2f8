"""TEST INFRASTRUCTURE — recipe that compiles the reference's own hot-path code into ``oracle/_ref/``.

    python oracle/build_ref.py            # needs /root/reference; writes only under oracle/_ref/ (git-ignored)

The reference is pure Python, so "compiling it from the sources where they lie" means byte-compiling:

  * ``/root/reference/src/replay_no_ros.py``  ->  ``oracle/_ref/replay_no_ros.bytecode`` (the whole module, unmodified:
    ``EKF_pose_estimation``, ``delta_phi``, ``displacement``, ``replay`` at their original line numbers);
  * ``/root/reference/src/EKF-SLAM.py``       ->  ``oracle/_ref/ekf_slam_functions.bytecode`` — only the function definitions
    ``predict / update / delta_phi / displacement`` and the two noise assignments (``:12-13``): the module body of
    that file cannot be imported, so the same AST selection ``ref_loader.load_ekf3_functions``
    applies at import time is applied here before compiling.

(The files are named *.bytecode, not *.pyc: snapshot tools commonly drop *.pyc.) No reference SOURCE is copied into
the repository; ``oracle/_ref/`` holds build outputs only, stays out of git
(.gitignore) and travels to the GPU box with the snapshot like the built ``.so``. There ``oracle/ref_loader.py``
loads the bytecode when ``/root/reference`` is absent, so the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` time the reference's OWN ``EKF_pose_estimation`` (``kind: "reference"``) and a ``-m gpu`` test drives the
reference's unchanged ``replay()`` over the CUDA core.
"""
from __future__ import annotations

import ast
import importlib._bootstrap_external as _be
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "_ref")
REFERENCE_ROOT = os.environ.get("EKF_REFERENCE_ROOT", "/root/reference")


def _write_pyc(code, path, source_path):
    st = os.stat(source_path)
    data = _be._code_to_timestamp_pyc(code, int(st.st_mtime), st.st_size & 0xFFFFFFFF)
    with open(path, "wb") as fh:
        fh.write(data)


def build(verbose: bool = True) -> bool:
    src_replay = os.path.join(REFERENCE_ROOT, "src", "replay_no_ros.py")
    src_ekf3 = os.path.join(REFERENCE_ROOT, "src", "EKF-SLAM.py")
    if not (os.path.isfile(src_replay) and os.path.isfile(src_ekf3)):
        if verbose:
            print(f"build_ref: {REFERENCE_ROOT} not present, nothing built")
        return False
    os.makedirs(OUT, exist_ok=True)
    with open(src_replay, "r") as fh:
        code = compile(fh.read(), src_replay, "exec", dont_inherit=True)
    _write_pyc(code, os.path.join(OUT, "replay_no_ros.bytecode"), src_replay)
    with open(src_ekf3, "r") as fh:
        tree = ast.parse(fh.read(), filename=src_ekf3)
    keep = [node for node in tree.body
            if (isinstance(node, ast.FunctionDef) and node.name in ("predict", "update", "delta_phi", "displacement"))
            or (isinstance(node, ast.Assign) and any(isinstance(t, ast.Name) and t.id in ("motion_noise", "observation_noise")
                                                     for t in node.targets))]
    code3 = compile(ast.Module(body=keep, type_ignores=[]), src_ekf3, "exec", dont_inherit=True)
    _write_pyc(code3, os.path.join(OUT, "ekf_slam_functions.bytecode"), src_ekf3)
    with open(os.path.join(OUT, "BUILD_INFO.txt"), "w") as fh:
        fh.write(f"python {sys.version.split()[0]} magic {importlib.util.MAGIC_NUMBER.hex()}\n"
                 f"replay_no_ros.bytecode <- {src_replay}\nekf_slam_functions.bytecode <- {src_ekf3} (function defs + noise constants)\n")
    if verbose:
        print(f"build_ref: wrote {OUT}/replay_no_ros.bytecode, ekf_slam_functions.bytecode")
    return True


if __name__ == "__main__":
    build()
