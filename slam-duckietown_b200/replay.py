"""Headless replay of a decoded bag directory through the CUDA filter core (configuration 1).

Two drivers:

* ``drive_reference_replay(module, dir, detections)`` — runs the reference's own ``replay()``
  (``src/replay_no_ros.py:66-248``) UNCHANGED, with its module globals ``EKF_pose_estimation``,
  ``delta_phi``, ``displacement`` rebound to ``ekf_bindings`` and the image / plot helpers stubbed
  (SURVEY.md §8b). Used wherever the reference file is present.
* ``replay_events(lines, detections)`` — the same event loop restated for machines without the reference
  (the GPU box): identical windowing (``DELTA_TIME`` phase anchored at the first event, at most one EKF step
  per event line, ``False`` sentinels for unset ticks so tick value 0 reads as unset, ``:74-81,:103-118,
  :190-238``).

``detections`` maps an image file name to a list of detection objects with ``tag_id``, ``pose_t`` (3,1),
``pose_err`` (the fields the hot path consumes) — the decoded-detections sidecar that replaces running
dt_apriltags on the PNG frames.
"""
from __future__ import annotations

import os
from types import SimpleNamespace

import numpy as np

from . import ekf_bindings as eb

DELTA_TIME = 0.7            # replay_no_ros.py:17
WHEEL_RADIUS = 0.0318       # :71
BASELINE = 0.1              # :72
RESOLUTION = 135            # :80


def make_detection(tag_id, pose_t, pose_err=0.0):
    """Duck type of a dt_apriltags detection (fields used at replay_no_ros.py:128,:286-301)."""
    return SimpleNamespace(tag_id=int(tag_id), pose_t=np.asarray(pose_t, dtype=np.float64).reshape(3, 1),
                           pose_err=float(pose_err), pose_R=np.eye(3), center=np.zeros(2),
                           corners=np.zeros((4, 2)))


def replay_events(lines, detections, on_step=None, ekf=None):
    """Restated event loop. Returns the list of per-step records (mean, cov, TAG_INDEX copy, update order)."""
    ekf = ekf or eb.EKF_pose_estimation
    mean = np.array([0.0, 0.0, 0.0])
    cov = np.eye(3) * eb._flag("MOTION_MODEL_VARIANCE")                    # :70
    prev_l = prev_r = cur_l = cur_r = False
    prev_ts = False
    tag_index = {}
    window = []
    records = []
    for line in lines:
        ts, event, data, *rest = line.strip().split(",")
        if event == "left_wheel":
            v = int(data)
            if prev_l == False:  # noqa: E712  (0 == False: tick 0 reads as unset, as in the reference)
                prev_l = v
            cur_l = v
        elif event == "right_wheel":
            v = int(data)
            if prev_r == False:  # noqa: E712
                prev_r = v
            cur_r = v
        elif event == "image":
            window.append((ts, detections.get(os.path.basename(data), [])))
        elif event in ("ground_truth", "landmarks", "camera_intrinsis"):
            pass
        else:
            raise ValueError(f"Unknown event {event!r}")                   # :185-188 exits
        ts = float(ts)
        if prev_ts == False:  # noqa: E712
            prev_ts = ts
        if ts - prev_ts > DELTA_TIME:                                     # :196
            prev_ts = prev_ts + DELTA_TIME
            dl = eb.delta_phi(cur_l, prev_l, RESOLUTION)
            dr = eb.delta_phi(cur_r, prev_r, RESOLUTION)
            prev_l, prev_r = cur_l, cur_r
            ang, lin = eb.displacement(WHEEL_RADIUS, BASELINE, dl, dr)
            mean, cov, tags = ekf(ang, lin, mean, cov, DELTA_TIME, window, tag_index)
            window.clear()
            rec = dict(ang=ang, lin=lin, mean=mean.copy(), cov=cov.copy(), tag_index=dict(tag_index),
                       order=[int(k) for k in tags.keys()])
            records.append(rec)
            if on_step:
                on_step(rec)
    return records


def drive_reference_replay(module, events_dir, detections):
    """Run ``module.replay(events_dir)`` — the reference's function, unmodified — against the CUDA core."""
    records = []

    def ekf(ang, lin, mean, cov, dt, pairs, tag_index):
        m, c, tags = eb.EKF_pose_estimation(ang, lin, mean, cov, dt, pairs, tag_index)
        records.append(dict(ang=float(ang), lin=float(lin), mean=m.copy(), cov=c.copy(),
                            tag_index=dict(tag_index), order=[int(k) for k in tags.keys()]))
        return m, c, tags

    names = ("EKF_pose_estimation", "delta_phi", "displacement", "load_grayscale", "detect_tags",
             "visualize_bounding_boxes", "plot_path")
    saved = {k: getattr(module, k) for k in names}
    saved_pause = module.plt.pause
    eb.use_flags_from(module)
    try:
        module.EKF_pose_estimation = ekf
        module.delta_phi = eb.delta_phi
        module.displacement = eb.displacement
        module.load_grayscale = os.path.basename
        module.detect_tags = lambda img, cam: detections.get(img, [])
        module.visualize_bounding_boxes = lambda *a, **k: None
        module.plot_path = lambda *a, **k: None
        module.plt.pause = lambda *a, **k: None
        module.print = lambda *a, **k: None
        module.replay(events_dir)
    finally:
        for k, v in saved.items():
            setattr(module, k, v)
        module.plt.pause = saved_pause
        if "print" in module.__dict__:
            del module.print
        eb.use_flags_from(None)
    return records
