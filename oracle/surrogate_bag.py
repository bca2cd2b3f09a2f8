"""TEST INFRASTRUCTURE — surrogate decoded bag for configuration 1.

The bundled ``quackgpt_small_town_joystick.bag`` is not in the container (``.MISSING_LARGE_BLOBS``) and
``dt_apriltags`` / ``rosbag`` are not installed, so the decoded wheel ticks and AprilTag detections are
SYNTHESISED from the bundled Vicon ground truth (``bags/*.csv``) in the ``events.csv`` wire format that
``scripts/decode_bag_file.py:62-105,:337-347`` writes and ``replay_no_ros.py:96-188`` parses:

    "{ts},{event},{data}\n"   event in {camera_intrinsis, landmarks, ground_truth, left_wheel, right_wheel, image}

Run only where /root/reference exists (by oracle/gen_golden.py); the result is committed under tests/golden/.
"""
from __future__ import annotations

import os

import numpy as np

WHEEL_RADIUS, BASELINE, RESOLUTION = 0.0318, 0.1, 135      # replay_no_ros.py:71-72,:80
GOD_SECRET_KEY = [80, 44, 41, 26, 110, 106, 101, 31, 65, 23, 232, 54]   # tag ids, replay_no_ros.py:26


def _read_vicon(path, skip=5):
    rows = []
    with open(path, "r") as fh:
        for i, line in enumerate(fh):
            if i < skip:
                continue
            parts = line.rstrip("\n").split(",")
            if len(parts) < 3 or parts[0] == "":
                continue
            rows.append([float(p) if p.strip() != "" else np.nan for p in parts])
    width = max(len(r) for r in rows)
    out = np.full((len(rows), width), np.nan)
    for i, r in enumerate(rows):
        out[i, :len(r)] = r
    return out


def build(bags_dir: str, seed: int = 0):
    """Returns (event_lines, detections) where detections maps frame file name -> list of
    (tag_id, pose_t(3,), pose_err)."""
    pose = _read_vicon(os.path.join(bags_dir, "quackgpt_small_town_joystick.csv"))
    traj = _read_vicon(os.path.join(bags_dir, "quackgpt_small_town_joystick_trajectories.csv"))
    rz = np.unwrap(pose[:, 4])
    tx, ty = pose[:, 5] / 1000.0, pose[:, 6] / 1000.0
    # express everything in the frame of the first robot pose (the filter starts at the origin)
    c0, s0 = np.cos(rz[0]), np.sin(rz[0])
    gx = c0 * (tx - tx[0]) + s0 * (ty - ty[0])
    gy = -s0 * (tx - tx[0]) + c0 * (ty - ty[0])
    gth = rz - rz[0]
    n_mark = (traj.shape[1] - 2) // 3
    lms = []
    for k in range(n_mark):
        x, y = traj[:, 2 + 3 * k] / 1000.0, traj[:, 3 + 3 * k] / 1000.0
        if np.all(np.isnan(x)):
            continue
        mx, my = (np.nanmin(x) + np.nanmax(x)) / 2, (np.nanmin(y) + np.nanmax(y)) / 2   # decode_bag_file.py:199-229
        lms.append((c0 * (mx - tx[0]) + s0 * (my - ty[0]), -s0 * (mx - tx[0]) + c0 * (my - ty[0])))
    lms = np.array(lms)
    tag_ids = [GOD_SECRET_KEY[k] if k < len(GOD_SECRET_KEY) else 300 + (k - len(GOD_SECRET_KEY))
               for k in range(len(lms))]

    # wheel ticks from inverse differential-drive kinematics
    dth = np.diff(gth)
    ds = np.diff(gx) * np.cos(gth[:-1]) + np.diff(gy) * np.sin(gth[:-1])
    phi_l = np.concatenate([[0.0], np.cumsum((ds - dth * BASELINE / 2) / WHEEL_RADIUS)])
    phi_r = np.concatenate([[0.0], np.cumsum((ds + dth * BASELINE / 2) / WHEEL_RADIUS)])
    ticks_l = np.floor(phi_l * RESOLUTION / (2 * np.pi)).astype(np.int64) + 1000   # 0 is the "unset" sentinel
    ticks_r = np.floor(phi_r * RESOLUTION / (2 * np.pi)).astype(np.int64) + 1000

    rng = np.random.default_rng(seed)
    t = pose[:, 0] / 100.0
    events = []   # (time, order, text)
    dets = {}
    k_mat = [340.0, 0.0, 328.0, 0.0, 336.0, 257.0, 0.0, 0.0, 1.0]
    intr = [[0.0] * 5, k_mat, [1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0], [0.0] * 12]
    events.append((t[0] - 1.0, "camera_intrinsis", repr(intr)))       # decode_bag_file.py:103
    events.append((t[0], "landmarks", repr([(float(a), float(b)) for a, b in lms])))
    for f in range(len(t)):
        if f % 3 == 0:
            events.append((t[f], "left_wheel", str(int(ticks_l[f]))))
            events.append((t[f], "right_wheel", str(int(ticks_r[f]))))
        if f % 10 == 0:
            events.append((t[f], "ground_truth", f"{gx[f]}, {gy[f]}"))
        if f % 7 == 0:
            name = "frame%06i.png" % f
            lst = []
            for k, (lx, ly) in enumerate(lms):
                dx, dy = lx - gx[f], ly - gy[f]
                xr = np.cos(gth[f]) * dx + np.sin(gth[f]) * dy
                yr = -np.sin(gth[f]) * dx + np.cos(gth[f]) * dy
                r, b = np.hypot(xr, yr), np.arctan2(yr, xr)
                if xr > 0.1 and abs(b) < 1.0 and r < 2.0:
                    xn, yn = xr + rng.normal(0, 0.02), yr + rng.normal(0, 0.02)
                    lst.append((int(tag_ids[k]), np.array([-yn, 0.0, xn]), float(abs(rng.normal(0, 1e-6)))))
            dets[name] = lst
            events.append((t[f], "image", name))
    order = sorted(range(len(events)), key=lambda i: events[i][0])     # stable, like decode_bag_file.py:337
    lines = ["%r,%s,%s\n" % (float(events[i][0]), events[i][1], events[i][2]) for i in order]
    return lines, dets, lms, tag_ids
