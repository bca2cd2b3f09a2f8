"""Batched Monte Carlo replay of ONE decoded bag (BASELINE.json configs[1] fed from events.csv instead of a
synthetic trajectory): B filters see the same encoder ticks and the same AprilTag detections, each with its own
process / measurement noise parameters and its own perturbation of the detections, and everything after the
event parsing runs on the device:

    windows_from_events (host, once; all detections uploaded in one copy)  ->  per step:  associate (a3/a4 per
    filter, the per-member detection noise generated inside the kernel)  ->  step_ticks_common (a1/a2 from the shared
    encoder ticks + fused a5-a11)  ->  accumulate_ate (against the stream's ground_truth events)

This is the README's ablation study (README.md:421-492: sweeps of MOTION_MODEL_VARIANCE / MEASUREMENT_MODEL_VARIANCE
judged by eye on plots) as numbers: an absolute-trajectory RMSE per ensemble member. With zero perturbation and the
reference's noise flags every member reproduces the scalar replay (tests/test_gpu_parity.py).
"""
from __future__ import annotations

import numpy as np

from . import ekf_bindings as eb
from .replay import BASELINE, RESOLUTION, WHEEL_RADIUS


def compact_tag_table(windows, ignore_tags=()):
    """tag id -> compact index in first-appearance order over the whole stream; ignored tags are left out."""
    table = {}
    for w in windows:
        for t in w.tag_id:
            t = int(t)
            if t not in ignore_tags and t not in table:
                table[t] = len(table)
    return table


def monte_carlo_replay(windows, batch, n_landmarks, m, motion_var, meas_var, det_sigma=0.0, seed=0,
                       ignore_tags=None, device=0, joseph=False, on_step=None, first_filter=0):
    """Returns dict(rmse (B,), pose (B,3), n_landmarks_seen (B,), faults, overflows, steps, tag_table (tag id ->
    compact index), tag_slot (n_tags, B) per-member slot of each compact tag, filter=BatchedEKF, last_step_args = the
    arguments of the last ``step_ticks_common`` call).

    motion_var / meas_var: scalars or (B,) arrays (MOTION_MODEL_VARIANCE / MEASUREMENT_MODEL_VARIANCE per member);
    det_sigma: std-dev (m) of the per-member Gaussian perturbation of each detection's pose_t[0] / pose_t[2], drawn
    ON the device by the counter-based generator (ekf_batch_detection_noise): a pure function of (seed, first_filter +
    b, window index, detection index), so a member's inputs are the same whatever the sharding and can be regenerated
    on the host (oracle/mc_noise.py:detection_noise) for per-member parity checks.
    first_filter: global index of this handle's filter 0 (rank r of a filter-sharded ensemble passes its range start)."""
    import torch
    ignore = tuple(eb._flag("IGNORE_TAGS")) if ignore_tags is None else tuple(ignore_tags)
    f = eb.BatchedEKF(batch, n_landmarks, device=device)
    dev = f.dev
    f.set_noise(np.broadcast_to(np.asarray(motion_var, dtype=np.float64), (batch,)).copy(),
                np.broadcast_to(np.asarray(meas_var, dtype=np.float64), (batch,)).copy())
    f.set_flags(joseph=joseph)
    f.init(float(np.asarray(motion_var, dtype=np.float64).reshape(-1)[0]))     # replay_no_ros.py:70
    table = compact_tag_table(windows, ignore)
    tag_slot, n_seen = f.new_slot_table(max(len(table), 1))
    sse = torch.zeros(batch, dtype=torch.float64, device=dev)
    scored = 0
    last_args = None
    # every window's detections go to the device in ONE copy (compact tag index, pose_t[0], pose_t[2] of all windows
    # back to back); a window is then three views into those arrays
    counts = [len(w.tag_id) for w in windows]
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    if offs[-1]:
        all_tag = np.concatenate([[table.get(int(t), -1) for t in w.tag_id] for w in windows if len(w.tag_id)])
        packed = np.concatenate([all_tag.astype(np.float64), np.concatenate([w.tx for w in windows]),
                                 np.concatenate([w.tz for w in windows])])
    else:
        packed = np.zeros(3)
    dpk = torch.from_numpy(packed).to(dev)
    total = int(offs[-1])
    d_tag = dpk[:total].to(torch.int32)
    d_tx, d_tz = dpk[total:2 * total], dpk[2 * total:3 * total]
    out_bufs = (torch.empty((batch, m), dtype=torch.int32, device=dev),
                torch.empty((batch, m), dtype=torch.float64, device=dev),
                torch.empty((batch, m), dtype=torch.float64, device=dev))
    for step, w in enumerate(windows):
        lo, hi = int(offs[step]), int(offs[step + 1])
        gen = (seed, step, first_filter, det_sigma) if (det_sigma > 0.0 and hi > lo) else None
        slot, rng, brg = f.associate(d_tag[lo:hi], d_tx[lo:hi], d_tz[lo:hi], tag_slot, n_seen, m, generated=gen,
                                     out=out_bufs)
        last_args = (w.ticks_l, w.ticks_r, w.prev_l, w.prev_r, WHEEL_RADIUS, BASELINE, RESOLUTION, slot, rng, brg)
        f.step_ticks_common(*last_args)
        if w.ground_truth is not None:
            f.accumulate_ate(w.ground_truth[0], w.ground_truth[1], sse)
            scored += 1
        if on_step:
            on_step(f, w)
    rmse = torch.sqrt(sse / max(scored, 1)).cpu().numpy()
    return dict(rmse=rmse, pose=f.mu[:, :3].cpu().numpy(), n_landmarks_seen=n_seen.cpu().numpy(),
                faults=f.fault_count(), overflows=f.associate_overflow_count(), steps=len(windows),
                scored_steps=scored, tag_table=table, tag_slot=tag_slot.cpu().numpy(), filter=f,
                last_step_args=last_args)
