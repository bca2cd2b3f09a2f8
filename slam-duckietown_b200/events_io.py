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
        data = f"{ev.data[0]!r},{ev.data[1]!r}"
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
