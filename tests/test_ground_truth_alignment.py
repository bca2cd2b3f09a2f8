"""events_io.ground_truth_events (the decoder's Vicon alignment, scripts/decode_bag_file.py:111-253, as a product
function): against the reference's OWN get_ground_truth on the bundled capture (this container only), and — on
synthetic Vicon exports, everywhere — the properties the alignment must have."""
import ast
import os

import numpy as np
import pytest

from slam_duckietown_b200 import events_io as io

REF = os.environ.get("EKF_REFERENCE_ROOT", "/root/reference")


def _reference_get_ground_truth():
    """get_ground_truth / rotate_around AST-extracted from the reference's decoder (its module body imports rosbag)."""
    path = os.path.join(REF, "scripts", "decode_bag_file.py")
    with open(path) as fh:
        tree = ast.parse(fh.read(), filename=path)
    keep = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in ("get_ground_truth", "rotate_around")]
    ns = {"os": os, "plot_path": lambda *a, **k: None}
    exec(compile(ast.Module(body=keep, type_ignores=[]), path, "exec"), ns)
    return ns["get_ground_truth"]


@pytest.mark.reference
@pytest.mark.skipif(not os.path.isfile(os.path.join(REF, "bags", "quackgpt_small_town_joystick.xcp")),
                    reason="needs the reference's bundled capture")
@pytest.mark.parametrize("offset,delay", [(3.21, 0.0), (0.0, 0.0), (20.0, 0.4)])
def test_alignment_equals_reference_decoder_on_bundled_capture(offset, delay):
    pre = os.path.join(REF, "bags", "quackgpt_small_town_joystick")
    start, end = io.capture_window_from_xcp(pre + ".xcp")
    first = start + offset
    want = _reference_get_ground_truth()(first, pre, delay)
    got = io.ground_truth_events(first, pre + ".csv", pre + "_trajectories.csv", start, end, delay)
    assert len(got) == len(want) > 4000
    assert (got[0].ts, got[0].kind) == want[0][:2] and repr(got[0].data) == want[0][2]      # the landmarks event
    for g, w in zip(got[1:], want[1:]):
        assert g.ts == w[0] and g.kind == w[1] == "ground_truth"
        assert f"{g.data[0]}, {g.data[1]}" == w[2]                                           # bit-exact text
    line = io.format_event(got[1])
    assert line == f"{want[1][0]},{want[1][1]},{want[1][2]}\n"                               # decoder's wire format
    back = io.parse_event_line(line)
    assert back.data == got[1].data


def _write_vicon(tmp_path, n_frames=400, rate=100, theta0=0.3):
    t = np.arange(1, n_frames + 1)
    x = 1000.0 + 2.0 * t
    y = -400.0 + 0.5 * t
    rz = np.full(n_frames, theta0)
    robot = tmp_path / "run.csv"
    with open(robot, "w") as fh:
        fh.write("Objects\n100\n,,Global Angle r:r,,,,,,\nFrame,Sub Frame,RX,RY,RZ,TX,TY,TZ\n,,rad,rad,rad,mm,mm,mm\n")
        for i in range(n_frames):
            if i == 17:
                fh.write(f"{t[i]},0\n")                                    # a dropped frame: not 8 fields
                continue
            fh.write(f"{t[i]},0,0.0,0.0,{rz[i]},{x[i]},{y[i]},140.0\n")
    marks = np.array([[1500.0, -300.0], [900.0, 250.0], [2100.0, 40.0]])
    traj = tmp_path / "run_trajectories.csv"
    with open(traj, "w") as fh:
        fh.write("Trajectories\n100\n,,m0,,,m1,,,m2,,\nFrame,Sub Frame,X,Y,Z,X,Y,Z,X,Y,Z\n,,mm,mm,mm,mm,mm,mm,mm,mm,mm\n")
        for i in range(n_frames):
            cells = []
            for k, (mx, my) in enumerate(marks):
                if k == 1 and i % 3 == 0:
                    cells += ["", "", ""]                                   # marker 1 occluded on every third frame
                else:
                    cells += [f"{mx + (i % 5) - 2.0}", f"{my + (i % 3) - 1.0}", "10.0"]
            fh.write(f"{t[i]},0," + ",".join(cells) + "\n")
    return str(robot), str(traj), marks, (x, y), theta0, rate


def test_alignment_properties_on_synthetic_capture(tmp_path):
    robot, traj, marks, (x, y), th0, rate = _write_vicon(tmp_path)
    start, end = 1000.0, 1000.0 + 400 / rate
    first = start + 1.005                                                   # closest frame: 100 or 101
    ev = io.ground_truth_events(first, robot, traj, start, end)
    assert ev[0].kind == "landmarks" and ev[0].ts == first and len(ev) == 1 + 399
    gts = ev[1:]
    i0 = int(np.argmin([abs(e.ts - first) for e in gts]))
    assert gts[i0].data == (0.0, 0.0)                                       # the start pose is the origin
    # every position: turned by +theta0 about the start position, shifted, millimetres -> metres
    f0 = round((gts[i0].ts - start) * rate)
    fk = round((gts[-1].ts - start) * rate)
    dx, dy = x[fk - 1] - x[f0 - 1], y[fk - 1] - y[f0 - 1]
    want = ((dx * np.cos(th0) - dy * np.sin(th0)) / 1000, (dx * np.sin(th0) + dy * np.cos(th0)) / 1000)
    assert np.allclose(gts[-1].data, want, rtol=0, atol=1e-12)
    # landmarks: midpoint of min / max per marker, same transform
    assert len(ev[0].data) == 3
    mx, my = marks[1][0], marks[1][1] + 0.5        # y jitter of the visible frames of marker 1 is {0, +1}
    ddx, ddy = mx - x[f0 - 1], my - y[f0 - 1]
    want_l = ((ddx * np.cos(th0) - ddy * np.sin(th0)) / 1000, (ddx * np.sin(th0) + ddy * np.cos(th0)) / 1000)
    assert np.allclose(ev[0].data[1], want_l, rtol=0, atol=1e-12)
    # the events round-trip through the wire format and feed windows_from_events as ground truth
    lines = [io.format_event(e) for e in ev]
    back = [io.parse_event_line(l) for l in lines]
    assert back[5].data == ev[5].data and back[0].data == [list(p) for p in ev[0].data] or back[0].data == ev[0].data
    with pytest.raises(ValueError):
        io.ground_truth_events(start + 500.0, robot, traj, start, end)      # > 2 s from every pose
    assert io.ground_truth_events(start + 5.9, robot, traj, start, end, delay=2.0)[1].ts == start + 0.01 + 2.0
