"""events.csv reader / writer, detections sidecar and the windowing that feeds the batched core (host logic, CPU)."""
import os

import numpy as np

from conftest import GOLDEN


def test_events_roundtrip_and_windows(surrogate, tmp_path):
    from slam_duckietown_b200 import capi, events_io as io
    lines, dets, ref = surrogate
    evs = [io.parse_event_line(l) for l in lines]
    assert {e.kind for e in evs} <= set(io.EVENT_KINDS) and len(evs) == len(lines)
    # tick / image lines are reproduced byte for byte; every line re-parses to the same event
    for l, e in zip(lines, evs):
        out = io.format_event(e)
        if e.kind in ("left_wheel", "right_wheel", "image"):
            assert out == l
        e2 = io.parse_event_line(out)
        assert (e2.ts, e2.kind) == (e.ts, e.kind) and e2.data == e.data
    p = tmp_path / "events.csv.gz"
    io.write_events(p, evs)
    back = io.read_events(p)
    assert [(e.ts, e.kind) for e in back] == [(e.ts, e.kind) for e in sorted(evs, key=lambda e: e.ts)]
    # sidecar round trip
    side = tmp_path / "detections.npz"
    io.save_detections(side, dets)
    d2 = io.load_detections(side)
    assert sorted(d2) == sorted(dets)
    for k in dets:
        assert [x.tag_id for x in d2[k]] == [x.tag_id for x in dets[k]]
        for a, b in zip(d2[k], dets[k]):
            assert np.array_equal(np.asarray(a.pose_t).reshape(3), np.asarray(b.pose_t).reshape(3))
    # the committed fixture loads through the same reader
    d3 = io.load_detections(os.path.join(GOLDEN, "surrogate", "detections.npz"))
    assert sorted(d3) == sorted(dets)
    # windowing == the reference replay's steps: same count, same odometry from the ticks
    wins = io.windows_from_events(evs, dets)
    assert len(wins) == len(ref["n"]) == 70
    lib = capi.lib()
    import ctypes as C
    for i, w in enumerate(wins):
        dl = lib.ekf_delta_phi(w.ticks_l, w.prev_l, 135)
        dr = lib.ekf_delta_phi(w.ticks_r, w.prev_r, 135)
        ang, lin = C.c_double(), C.c_double()
        assert lib.ekf_displacement(0.0318, 0.1, dl, dr, C.byref(ang), C.byref(lin)) == 0
        assert ang.value == ref["ang"][i] and lin.value == ref["lin"][i]
    assert sum(len(w.tag_id) for w in wins) > 0 and any(w.ground_truth is not None for w in wins)


def test_unknown_event_raises():
    import pytest
    from slam_duckietown_b200 import events_io as io
    with pytest.raises(ValueError):
        io.parse_event_line("1.0,bogus,3\n")


def test_compact_tag_table_first_appearance_and_ignored():
    """The Monte Carlo driver's tag id -> compact index map: first appearance over the stream, ignored tags left
    out (they reach the device as -1)."""
    from slam_duckietown_b200 import events_io as io, monte_carlo as mc
    w = lambda tags: io.StepWindow(0, 0, 0, 0, np.array(tags, dtype=np.int64), np.zeros(len(tags)), np.zeros(len(tags)), None)
    wins = [w([44, 80, 44]), w([]), w([26, 80, 300])]
    assert mc.compact_tag_table(wins) == {44: 0, 80: 1, 26: 2, 300: 3}
    assert mc.compact_tag_table(wins, ignore_tags=(80,)) == {44: 0, 26: 1, 300: 2}
