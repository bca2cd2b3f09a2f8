"""Decoded-bag wire format: ``events.csv`` reader / writer and the decoded-detections sidecar.

``scripts/decode_bag_file.py:62-105,:337-347`` writes one line per event, sorted by timestamp::

    "{timestamp},{event},{data}\\n"

with ``event`` in ``image`` (data = frame file name), ``left_wheel`` / ``right_wheel`` (data = int tick),
``ground_truth`` (data = "x,y"), ``landmarks`` (data = repr of a list of [x, y]) and ``camera_intrinsis``
(sic; data = repr of [D, K, R, P]); ``src/replay_no_ros.py:96-188`` parses exactly that. The PNG frames themselves
are replaced by a sidecar ``detections.npz`` holding what dt_apriltags would return for every frame — the only
fields the hot path consumes (``tag_id``, ``pose_t``, ``pose_err``; ``replay_no_ros.py:128,:286-301``).

``windows_from_events`` turns an event stream into the per-step inputs of the batched Monte Carlo core: encoder
ticks at every ``DELTA_TIME`` boundary and the window's detections in order (same windowing as ``replay()``,
``:190-238``), so that a real decoded bag can drive ``BatchedEKF.step_ticks`` + ``BatchedEKF.associate`` directly.
"""
from __future__ import annotations

import ast
import gzip
import os
from dataclasses import dataclass, field
from typing import Iterable

import numpy as np

from .replay import DELTA_TIME, make_detection

EVENT_KINDS = ("image", "left_wheel", "right_wheel", "ground_truth", "landmarks", "camera_intrinsis")


@dataclass
class Event:
    ts: float
    kind: str
    data: object          # str (image), int (ticks), (x, y) (ground_truth), list (landmarks / camera_intrinsis)
    ts_text: str = ""     # the timestamp exactly as written (replay() hands this string on to the EKF, :128)


def _open_text(path, mode="rt"):
    return gzip.open(path, mode) if str(path).endswith(".gz") else open(path, mode)


def parse_event_line(line: str) -> Event:
    """One line of events.csv (replay_no_ros.py:97-188). Unknown events raise, where the reference exits."""
    first = line.find(",")
    second = line.find(",", first + 1)
    if first < 0 or second < 0:
        raise ValueError(f"malformed event line {line!r}")
    ts_text, kind, payload = line[:first], line[first + 1:second], line[second + 1:].rstrip("\n")
    if kind in ("left_wheel", "right_wheel"):
        data = int(payload)
    elif kind == "image":
        data = payload.strip()
    elif kind == "ground_truth":
        x, y, *_ = payload.split(",")
        data = (float(x), float(y))
    elif kind in ("landmarks", "camera_intrinsis"):
        data = ast.literal_eval(payload)          # the reference uses eval() here (:141,:167)
    else:
        raise ValueError(f"Unknown event {kind!r}")
    return Event(float(ts_text), kind, data, ts_text)


def read_events(path) -> list[Event]:
    with _open_text(path) as fh:
        return [parse_event_line(line) for line in fh if line.strip()]


def format_event(ev: Event) -> str:
    """The writer's formatting (decode_bag_file.py:337-347): non-string data goes through repr()."""
    if ev.kind == "ground_truth":
        data = f"{ev.data[0]!r}, {ev.data[1]!r}"          # "x, y" with the decoder's blank (decode_bag_file.py:178)
    elif isinstance(ev.data, str):
        data = ev.data
    else:
        data = repr(ev.data).replace("\n", "")
    ts = ev.ts_text or repr(ev.ts)
    return f"{ts},{ev.kind},{data}\n"


def write_events(path, events: Iterable[Event]) -> None:
    """Events sorted on the timestamp (stable), as decode_bag_file.py:335 does."""
    with _open_text(path, "wt") as fh:
        for ev in sorted(events, key=lambda e: e.ts):
            fh.write(format_event(ev))


# ---------------------------------------------------------------------------------------------------------
# decoded-detections sidecar
# ---------------------------------------------------------------------------------------------------------
def save_detections(path, detections: dict) -> None:
    """detections: frame file name -> list of objects with tag_id, pose_t (3,1), pose_err. Flat arrays + offsets."""
    names = sorted(detections)
    counts = np.array([len(detections[n]) for n in names], dtype=np.int64)
    flat = [d for n in names for d in detections[n]]
    np.savez_compressed(path, frames=np.array(names), offsets=np.concatenate([[0], np.cumsum(counts)]),
                        tag_id=np.array([d.tag_id for d in flat], dtype=np.int64),
                        pose_t=np.array([np.asarray(d.pose_t, dtype=np.float64).reshape(3) for d in flat]).reshape(-1, 3),
                        pose_err=np.array([d.pose_err for d in flat], dtype=np.float64))


def load_detections(path) -> dict:
    z = np.load(path, allow_pickle=False)
    out = {}
    off = z["offsets"]
    for i, name in enumerate(z["frames"] if "frames" in z.files else z["names"]):
        out[str(name)] = [make_detection(int(z["tag_id"][k]), z["pose_t"][k], float(z["pose_err"][k]))
                          for k in range(int(off[i]), int(off[i + 1]))]
    return out


# ---------------------------------------------------------------------------------------------------------
# event stream -> per-step inputs of the batched core
# ---------------------------------------------------------------------------------------------------------
@dataclass
class StepWindow:
    ticks_l: int
    ticks_r: int
    prev_l: int
    prev_r: int
    tag_id: np.ndarray            # (D,) raw tag ids of the window's detections, in window order
    tx: np.ndarray                # (D,) pose_t[0]
    tz: np.ndarray                # (D,) pose_t[2]
    ground_truth: tuple | None    # last (x, y) seen up to this step, if the stream carries any
    ts: float = 0.0
    frames: list = field(default_factory=list)


def windows_from_events(events: Iterable[Event], detections: dict) -> list[StepWindow]:
    """The windowing of replay() (:190-238): phase anchored at the first event, an EKF step whenever an event lies
    more than DELTA_TIME past the anchor (at most one per event), tick value 0 read as "unset" (:105,:114)."""
    prev_l = prev_r = cur_l = cur_r = False
    prev_ts = False
    gt = None
    tag, tx, tz, frames = [], [], [], []
    out = []
    for ev in events:
        if ev.kind == "left_wheel":
            if prev_l == False:  # noqa: E712
                prev_l = ev.data
            cur_l = ev.data
        elif ev.kind == "right_wheel":
            if prev_r == False:  # noqa: E712
                prev_r = ev.data
            cur_r = ev.data
        elif ev.kind == "image":
            name = os.path.basename(ev.data)
            frames.append(name)
            for d in detections.get(name, []):
                tag.append(int(d.tag_id))
                tx.append(float(d.pose_t[0][0]))
                tz.append(float(d.pose_t[2][0]))
        elif ev.kind == "ground_truth":
            gt = ev.data
        if prev_ts == False:  # noqa: E712
            prev_ts = ev.ts
        if ev.ts - prev_ts > DELTA_TIME:
            prev_ts = prev_ts + DELTA_TIME
            out.append(StepWindow(int(cur_l), int(cur_r), int(prev_l), int(prev_r), np.array(tag, dtype=np.int64),
                                  np.array(tx, dtype=np.float64), np.array(tz, dtype=np.float64), gt, ev.ts,
                                  frames))
            prev_l, prev_r = cur_l, cur_r
            tag, tx, tz, frames = [], [], [], []
    return out


# ---------------------------------------------------------------------------------------------------------
# ground-truth alignment (scripts/decode_bag_file.py:111-253 `get_ground_truth`)
# ---------------------------------------------------------------------------------------------------------
def capture_window_from_xcp(path) -> tuple[float, float]:
    """(start, end) of the Vicon capture as POSIX seconds from the system file's ``Camera/Capture`` element, with the
    decoder's fixed UTC-5 correction (decode_bag_file.py:128-133)."""
    import xml.etree.ElementTree as ET
    from datetime import datetime, timedelta
    cap = ET.parse(path).getroot().find("Camera/Capture")
    shift = timedelta(hours=5).total_seconds()
    return (datetime.fromisoformat(cap.get("START_TIME")).timestamp() + shift,
            datetime.fromisoformat(cap.get("END_TIME")).timestamp() + shift)


def _vicon_rows(path, skip=5):
    with open(path, "r") as fh:
        for _ in range(skip):
            fh.readline()
        return [line.split(",") for line in fh.read().strip().splitlines()]


def _into_start_frame(ax, ay, x, y, theta):
    """The decoder's ``rotate_around`` (decode_bag_file.py:107-109): (x, y) turned by +theta about (ax, ay)."""
    from math import cos, sin
    return ax + (x - ax) * cos(theta) - (y - ay) * sin(theta), ay + (x - ax) * sin(theta) + (y - ay) * cos(theta)


def ground_truth_events(first_timestamp: float, robot_csv, trajectories_csv, start_capture_time: float,
                        end_capture_time: float, delay: float = 0.0) -> list[Event]:
    """Vicon exports -> the ``landmarks`` event (at ``first_timestamp``) and one ``ground_truth`` event per valid robot
    frame, in the frame of the robot pose closest in time to the bag's first event, metres. Semantics of the decoder,
    including its quirks:

    * frame -> time: ``frame / framerate + start + delay`` with ``framerate = round(all rows / capture length)``
      (:151-157); only rows with exactly 8 fields are poses (:150);
    * the start pose is the one closest in time to ``first_timestamp``; more than 2 s away is an error (:160-167,
      raised here where the decoder exits);
    * every position is turned about the start position by +RZ of the start pose, shifted to it and divided by 1000
      (:170-181) — positions only, no heading;
    * landmarks: per marker the midpoint of the min / max of its x and y over the capture (:199-229), a coordinate of
      exactly 0 counting as missing (``!= False``); markers never seen are skipped here (the decoder would fail on
      them); the de-duplication test compares the raw millimetre midpoint with the already NORMALISED landmarks
      (:233-238), so it only ever fires by coincidence — kept as is.
    """
    rows = _vicon_rows(robot_csv)
    framerate = round(len(rows) / (end_capture_time - start_capture_time))
    poses = [tuple(float(v) for v in r) for r in rows if len(r) == 8]
    timed = [(int(p[0]) / framerate + start_capture_time + delay,) + p[1:] for p in poses]
    start = min(timed, key=lambda p: abs(p[0] - first_timestamp))
    if abs(start[0] - first_timestamp) > 2:
        raise ValueError("no ground-truth pose within 2 s of the bag's first event "
                         f"(closest differs by {start[0] - first_timestamp:.3f} s): pass a delay")
    th0, x0, y0 = start[4], start[5], start[6]
    out = []
    for p in timed:
        x, y = _into_start_frame(x0, y0, p[5], p[6], th0)
        out.append(Event(p[0], "ground_truth", ((x - x0) / 1000, (y - y0) / 1000)))
    marks = [r for r in _vicon_rows(trajectories_csv) if len(r) > 4]
    vals = [[float(v) if v != "" else 0.0 for v in r] for r in marks]     # 0.0 doubles as "missing", as in the decoder
    n_marks = (len(vals[0]) + 1) // 3 if vals else 0                      # len(time-stamped row) // 3 (:197)
    landmarks = []
    for k in range(n_marks):
        xs = [v[3 * k + 2] for v in vals if len(v) > 3 * k + 2 and v[3 * k + 2] != 0.0]
        ys = [v[3 * k + 3] for v in vals if len(v) > 3 * k + 3 and v[3 * k + 3] != 0.0]
        if not xs or not ys:
            continue
        mx, my = (max(xs) + min(xs)) / 2, (max(ys) + min(ys)) / 2
        if any((mx - lx) ** 2 + (my - ly) ** 2 < 10 ** 2 for lx, ly in landmarks):
            continue
        x, y = _into_start_frame(x0, y0, mx, my, th0)
        landmarks.append(((x - x0) / 1000, (y - y0) / 1000))
    out.insert(0, Event(first_timestamp, "landmarks", landmarks))
    return out
