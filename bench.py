#!/usr/bin/env python
"""bench.py — EKF filter-steps/s on the batched Monte Carlo configuration of BASELINE.json
(configs[1]: 65,536 independent filters x 20 landmarks, 43-dim state, M = 3 updates per step, 1 B200),
plus ms/update of the 4096-landmark large-map filter as a secondary figure on the same JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W]          # own arm (CUDA core through the C-ABI)
    python bench.py --impl reference [...]                        # the reference's CPU path (oracle port)

A "step" is one fused filter-step of every filter on the GPU (one kernel launch). For N > 1 every rank runs
its own 65,536 filters (filters never interact: weak scaling, no data-path collective).
"""
import os
import sys

os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")      # the CPU baseline parallelises over filters
os.environ.setdefault("OMP_NUM_THREADS", "1")

import argparse  # noqa: E402
import json  # noqa: E402
import subprocess  # noqa: E402
import threading  # noqa: E402
import time  # noqa: E402

import numpy as np  # noqa: E402

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BATCH, NLM, M = 65536, 20, 3
N_STATE = 3 + 2 * NLM
# algorithmic bytes per filter-step: one read + one write of mu and P, the per-filter
# odometry, M x (range, bearing, slot) and the two noise parameters
BYTES_PER_FILTER_STEP = 16 * (N_STATE * N_STATE + N_STATE) + 8 * (2 + 2 * M) + 4 * M + 16     # 30 364
LM_LANDMARKS = 4096
LM_N = 3 + 2 * LM_LANDMARKS
LM_BYTES_PER_UPDATE = 16 * LM_N * LM_N + 96 * LM_N                                            # 1 075 315 120
METRIC, UNIT = "EKF filter-steps/s (B=65536, 20 lm)", "filter-steps/s"
WORKLOAD = "batched Monte Carlo replay: 65536 filters x 20 landmarks (n=43), predict + 3 updates per step"


def set_workload(name, world):
    """--workload M: BASELINE.json configs[3], 1,048,576 filters x 8 landmarks (n=19, 2 updates per step), the
    filters split contiguously across the ranks (strong scaling). Default B is the contract line's workload."""
    global BATCH, NLM, M, N_STATE, BYTES_PER_FILTER_STEP, METRIC, WORKLOAD, SCALING
    if name == "M":
        total = 1 << 20
        BATCH, NLM, M = total // world, 8, 2
        N_STATE = 3 + 2 * NLM
        BYTES_PER_FILTER_STEP = 16 * (N_STATE * N_STATE + N_STATE) + 8 * (2 + 2 * M) + 4 * M + 16   # 6 152
        METRIC = "EKF filter-steps/s (B=1048576, 8 lm)"
        WORKLOAD = (f"batched Monte Carlo replay: 1048576 filters x 8 landmarks (n=19) split over {world} GPU(s), "
                    "predict + 2 updates per step")
        SCALING = "strong"


SCALING = "weak"


def kernel_label():
    """(ncu_traffic key, description) of the kernel the C-ABI dispatches this workload to (ekf_capi.cu: capacities
    <= 8 landmarks -> group kernel, else the W-in-registers warp kernel), unless EKF_BATCH_KERNEL forces a variant."""
    forced = os.environ.get("EKF_BATCH_KERNEL", "")
    if forced:
        return f"{forced}<{NLM},{M}>", f"EKF_BATCH_KERNEL={forced} variant <{NLM},{M}>"
    if NLM <= 8:
        return (f"batch_step_group_kernel<{NLM},{M},8>",
                f"batch_step_group_kernel<{NLM},{M},8> (8 lanes per filter, four filters per warp)")
    return (f"batch_step_regw_kernel<{NLM},{M}>",
            f"batch_step_regw_kernel<{NLM},{M},1> (one warp per filter, W in registers)")


def make_config(joseph=False):
    """The `config` object both arms print (same workload, same keys)."""
    return {"workload": WORKLOAD, "filters_per_gpu": BATCH, "landmarks": NLM, "updates_per_step": M,
            "cov_form": "joseph" if joseph else "plain (reference, replay_no_ros.py:480)",
            "l2": f"inputs larger than L2 ({BATCH * N_STATE * ((N_STATE + 3) & ~3) * 8 / 1e9:.2f} GB of covariance per "
                  "GPU streamed every step)",
            "sharding": "filters split across ranks, no data-path collective"}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def ncu_traffic(kernel_key):
    """dram bytes per launch of the dominant kernel from the committed ncu summary (profiles/), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
            return json.load(fh).get(kernel_key)
    except Exception:
        return None


class ClockSampler:
    """Samples SM clocks / throttle reasons while the timed region runs: NVML in a thread every ~2 ms (the timed
    region of the default run lasts tens of milliseconds), nvidia-smi -lms 100 as the fallback."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20, "sw_power_cap": 0x4}

    def __init__(self, index):
        self.index, self.rows, self.proc, self.thread, self.stop_flag = index, [], None, None, False
        self.sm, self.mx, self.reasons, self.source = [], 0.0, set(), None

    def _nvml_loop(self, nv, h):
        while not self.stop_flag:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))
                for name, bit in self.BITS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and vis.split(",")[self.index].isdigit() else self.index
            h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM))
            self.source = "nvml"
            self.thread = threading.Thread(target=self._nvml_loop, args=(nv, h), daemon=True)
            self.thread.start()
            return
        except Exception:
            self.source = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        self.stop_flag = True
        if self.thread:
            self.thread.join(timeout=1)
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = list(self.sm), self.mx, set(self.reasons)
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm), "source": self.source}


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the reference's dense numpy cost profile (oracle port, form="dense"), one filter at a time,
# parallel over filters on the host cores. This is the ONLY place bench.py executes oracle/ code, and only
# as the measured baseline — never on the product path.
# ---------------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    seed, n_filters, n_steps, form = args
    from oracle import ekf_oracle as eo
    from slam_duckietown_b200 import workloads as wl
    w = wl.make_batched(n_filters, NLM, M, steps=-(-NLM // M) + n_steps, perturb_seed=seed)
    states = []
    for b in range(n_filters):          # untimed: initialise all landmarks (the 7 augmenting steps)
        mean = np.zeros(N_STATE)
        cov = np.zeros((N_STATE, N_STATE))
        cov[:3, :3] = np.eye(3) * 0.1
        nk = 0
        fl = eo.Flags(motion_var=float(w.motion_var[b]), meas_var=float(w.meas_var[b]))
        for t in range(w.warm_steps):
            obs = [(int(s), float(r), float(q)) for s, r, q in zip(w.slot[t, b], w.rng[t, b], w.brg[t, b]) if s >= 0]
            mean, cov, nk = eo.step_from_observations(mean, cov, w.ang[t, b], w.lin[t, b], obs, fl, nk)
        states.append((mean, cov, nk, fl))
    t0 = time.perf_counter()
    for b in range(n_filters):          # timed: steady-state steps, predict + M updates each
        mean, cov, nk, fl = states[b]
        for t in range(w.warm_steps, w.warm_steps + n_steps):
            obs = [(int(s), float(r), float(q)) for s, r, q in zip(w.slot[t, b], w.rng[t, b], w.brg[t, b])]
            mean, cov, nk = eo.step_from_observations(mean, cov, w.ang[t, b], w.lin[t, b], obs, fl, nk, form=form)
    return time.perf_counter() - t0, n_filters * n_steps


def cpu_filter_steps_per_s(filters_per_core=48, steps=8, form="dense", cores=None):
    """Whole-host throughput of the reference's CPU path on a bounded sample of the workload."""
    import multiprocessing as mp
    cores = cores or os.cpu_count() or 1
    ctx = mp.get_context("fork")
    with ctx.Pool(cores) as pool:
        t0 = time.perf_counter()
        res = pool.map(_cpu_worker, [(100 + i, filters_per_core, steps, form) for i in range(cores)])
        wall = time.perf_counter() - t0
    work = sum(r[1] for r in res)
    busy = max(r[0] for r in res)
    return work / busy, cores, work, wall


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = []
    total = args.warmup + args.steps
    for i in range(total):
        rate, cores, work, _ = cpu_filter_steps_per_s(filters_per_core=32, steps=60, form="dense", cores=cores)
        if i >= args.warmup:
            per_step.append((rate, work))
    value = float(np.mean([r for r, _ in per_step]))
    sample = (f"{per_step[0][1]} filter-steps per bench step (32 filters x 60 steady-state steps per core) of the "
              f"{WORKLOAD}; dense numpy restatement of replay_no_ros.py:269-482 (oracle port; the Python reference "
              "itself cannot travel to the GPU box), one process per core, OPENBLAS_NUM_THREADS=1")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * per_step[0][1] / value, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": make_config(args.joseph),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# own arm
# ---------------------------------------------------------------------------------------------------------
def _dist_setup(world):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    return rank, local


def _barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def _max_over_ranks(x, world):
    import torch
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def bench_large_map(updates=60, warm=6):
    """Secondary figure: ms per range-bearing update of the 4096-landmark filter (P = 537 MB > L2)."""
    import torch
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    w = wl.make_large_map(LM_LANDMARKS, updates=updates + warm)
    g = eb.LargeMapEKF(LM_LANDMARKS)
    g.set_noise(0.1, 0.7)
    dev = g.mu.device
    a = torch.from_numpy(w.factor).to(dev)
    g.P[:, :g.n] = a @ a.T / 64.0
    g.P[:, :g.n].diagonal().add_(0.1)
    g.mu[:g.n] = torch.from_numpy(w.mu0).to(dev)
    del a
    ups = [o for o in w.ops if o[0] == "update"]
    for o in ups[:warm]:
        g.update(o[1], o[2], o[3])
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in ups[warm:]]
    for (e0, e1), o in zip(ev, ups[warm:]):
        e0.record()
        g.update(o[1], o[2], o[3])
        e1.record()
    torch.cuda.synchronize()
    ms = np.array([e0.elapsed_time(e1) for e0, e1 in ev])
    # fused form: the 3 updates of a step in ONE pass over P (ekf_lm_update_multi)
    trip = [ups[i:i + 3] for i in range(0, len(ups) - 2, 3)]
    for t3 in trip[:2]:
        g.update_multi([o[1] for o in t3], [o[2] for o in t3], [o[3] for o in t3])
    torch.cuda.synchronize()
    ev3 = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in trip[2:]]
    for (e0, e1), t3 in zip(ev3, trip[2:]):
        e0.record()
        g.update_multi([o[1] for o in t3], [o[2] for o in t3], [o[3] for o in t3])
        e1.record()
    torch.cuda.synchronize()
    ms3 = np.array([e0.elapsed_time(e1) for e0, e1 in ev3])
    # large-map predict (a6-a7 in rank form: rows / columns 0-2 only, ~96 n bytes, latency-bound): microseconds
    preds = [o for o in w.ops if o[0] == "predict"][:20] or [("predict", 0.05, 0.03)] * 20
    g.predict(preds[0][1], preds[0][2])
    torch.cuda.synchronize()
    evp = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in preds]
    for (e0, e1), o in zip(evp, preds):
        e0.record()
        g.predict(o[1], o[2])
        e1.record()
    torch.cuda.synchronize()
    msp = np.array([e0.elapsed_time(e1) for e0, e1 in evp])
    finite = bool(torch.isfinite(g.mu).all())
    g.close()
    peak, which = measured_peak_gbs()
    ach = LM_BYTES_PER_UPDATE / (float(np.mean(ms)) * 1e-3) / 1e9
    return {"metric": "ms/update @4096 lm", "ms_per_update": float(np.mean(ms)), "ms_min": float(ms.min()),
            "updates": len(ms), "finite": finite, "predict_us": float(np.median(msp)) * 1e3,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "peak_source": which, "unit": "GB/s",
                         "frac": ach / peak, "frac_of_8TBs": ach / 8000.0, "bytes_per_update": LM_BYTES_PER_UPDATE,
                         "traffic": ncu_traffic("lm_rank_update_kernel"),
                         "note": "gather + streaming kernels of one update timed together"},
            "fused_step": {"updates_per_step": 3, "ms_per_step": float(np.mean(ms3)),
                           "ms_per_update_equiv": float(np.mean(ms3)) / 3.0, "steps": len(ms3),
                           "hbm_gbs": (LM_BYTES_PER_UPDATE - 96 * (3 + 2 * LM_LANDMARKS)
                                       + 3 * 96 * (3 + 2 * LM_LANDMARKS)) / (float(np.mean(ms3)) * 1e-3) / 1e9,
                           "note": "ekf_lm_update_multi: 3 sequential updates with one read + write of P "
                                   "(16 n^2 + 3 x 96 n bytes per step), bitwise equal to 3 ekf_lm_update calls"}}


def bench_config_m(device=0, steps=20, warmup=3):
    """Secondary figure on the contract line: BASELINE.json configs[3] on ONE GPU — 1,048,576 filters x 8 landmarks
    (n = 19), predict + 2 updates per step, device-resident, CUDA events around every launch (`--workload M` is the
    full bench of this configuration, with its end-to-end figure and its split over N GPUs)."""
    import torch
    from slam_duckietown_b200 import ekf_bindings as eb, workloads as wl
    batch, nlm, m = 1 << 20, 8, 2
    n = 3 + 2 * nlm
    bytes_per = 16 * (n * n + n) + 8 * (2 + 2 * m) + 4 * m + 16                                  # 6 152
    distinct = 8
    w = wl.make_batched(batch, nlm, m, steps=-(-nlm // m) + distinct, perturb_seed=2)
    f = eb.BatchedEKF(batch, nlm, device=device)
    f.set_noise(w.motion_var, w.meas_var)
    f.init(0.1)

    def up(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(f.dev)
    for t in range(w.warm_steps):
        f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
    steady = [(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
              for t in range(w.warm_steps, w.warm_steps + distinct)]
    for i in range(warmup):
        f.step(*steady[i % distinct])
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for i in range(steps):
        ev[i][0].record()
        f.step(*steady[(warmup + i) % distinct])
        ev[i][1].record()
    torch.cuda.synchronize()
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    ok = bool(torch.isfinite(f.mu).all()) and f.fault_count() == 0
    peak, which = measured_peak_gbs()
    ach = bytes_per * batch / (ms * 1e-3) / 1e9
    key = f"batch_step_group_kernel<{nlm},{m},8>"
    return {"metric": "EKF filter-steps/s (B=1048576, 8 lm)", "value": batch / (ms * 1e-3), "unit": UNIT,
            "ms_per_step": ms, "steps": steps, "finite_and_faultless": ok,
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "peak_source": which, "unit": "GB/s",
                         "frac": ach / peak, "frac_of_8TBs": ach / 8000.0, "bytes_per_filter_step": bytes_per,
                         "kernel": key + " (8 lanes per filter, four filters per warp)", "traffic": ncu_traffic(key)}}


def run_own_arm(args, world):
    import torch
    from slam_duckietown_b200 import capi, ekf_bindings as eb, workloads as wl
    capi.lib()
    rank, local = _dist_setup(world)
    K, W = args.steps, max(args.warmup, 3)
    distinct = min(K + W, 24)                   # distinct steady-state input sets, cycled
    w = wl.make_batched(BATCH, NLM, M, steps=-(-NLM // M) + distinct, perturb_seed=2 + rank)
    f = eb.BatchedEKF(BATCH, NLM, device=local)
    f.set_noise(w.motion_var, w.meas_var)
    f.set_flags(joseph=args.joseph)
    f.init(0.1)
    dev = f.dev

    def up(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for t in range(w.warm_steps):               # untimed: every landmark gets initialised
        f.step(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
    steady = [(up(w.ang[t]), up(w.lin[t]), up(w.slot[t]), up(w.rng[t]), up(w.brg[t]))
              for t in range(w.warm_steps, w.warm_steps + distinct)]
    pinned = [tuple(torch.from_numpy(np.ascontiguousarray(a[t])).pin_memory()
                    for a in (w.ang, w.lin, w.slot, w.rng, w.brg))
              for t in range(w.warm_steps, w.warm_steps + distinct)]
    pose = torch.empty((BATCH, 3), dtype=torch.float64).pin_memory()
    assert f.fault_count() == 0 and int(f.n_known.min()) == NLM

    # ---- device-resident timing: K launches, CUDA events on the launching stream ----------------------
    for i in range(W):
        f.step(*steady[i % distinct])
    _barrier(world)
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    t_all0, t_all1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_all0.record()
    for i in range(K):
        ev[i][0].record()
        f.step(*steady[(W + i) % distinct])
        ev[i][1].record()
    t_all1.record()
    _barrier(world)
    total_ms = _max_over_ranks(t_all0.elapsed_time(t_all1), world)
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    ok = bool(torch.isfinite(f.mu).all()) and f.fault_count() == 0

    # ---- end to end: host buffers in, pose out, through the C-ABI host entry point -----------------------
    # every step: H2D of that step's pinned inputs, the fused step, D2H of every filter's pose; the pipelined
    # entry point overlaps step i's copies with step i-1's kernel (three streams, two staging sets)
    poses = [pose, torch.empty_like(pose).pin_memory()]
    for i in range(W):
        f.step_host_async(*pinned[i % distinct], poses[i & 1])
    f.host_sync()
    _barrier(world)
    t0 = time.perf_counter()
    for i in range(K):
        f.step_host_async(*pinned[(W + i) % distinct], poses[i & 1])
    f.host_sync()
    _barrier(world)
    e2e_s = _max_over_ranks(time.perf_counter() - t0, world)
    # the synchronous call (one step at a time, nothing overlapped), for reference
    t0 = time.perf_counter()
    for i in range(K):
        f.step_host(*pinned[(W + i) % distinct], pose)
    _barrier(world)
    e2e_sync_s = _max_over_ranks(time.perf_counter() - t0, world)
    ok = ok and bool(np.isfinite(poses[0].numpy()).all()) and bool(np.isfinite(poses[1].numpy()).all())
    clk = clocks.stop() if rank == 0 else None
    h2d = BATCH * (8 + 8 + M * (4 + 8 + 8))
    d2h = BATCH * 3 * 8

    if rank != 0:
        return
    peak, which = measured_peak_gbs()
    ach = BYTES_PER_FILTER_STEP * BATCH / (kern_ms * 1e-3) / 1e9
    value = BATCH * world * K / (total_ms * 1e-3)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": SCALING, "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": make_config(args.joseph),
            "roofline": {"bound": "hbm", "achieved": ach, "peak": peak, "peak_source": which, "unit": "GB/s",
                         "frac": ach / peak, "frac_of_8TBs": ach / 8000.0,
                         "bytes_per_filter_step": BYTES_PER_FILTER_STEP, "kernel": kernel_label()[1],
                         "kernel_ms": kern_ms, "traffic": ncu_traffic(kernel_label()[0])},
            "e2e": {"value": BATCH * world * K / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h,
                    "unpipelined_value": BATCH * world * K / e2e_sync_s,
                    "note": "ekf_batch_step_host_async + ekf_batch_host_sync: every step copies its pinned host "
                            "odometry+observations in, runs the fused step and copies the robot pose (x,y,theta) of "
                            "every filter out; copies of step i overlap the kernel of step i-1; filter state stays "
                            "resident in HBM. unpipelined_value = the synchronous ekf_batch_step_host"},
            "gpu_launches": K, "clocks": clk, "finite_and_faultless": ok}
    if world == 1:
        if not args.no_cpu_baseline:
            rate, cores, work, wall = cpu_filter_steps_per_s(filters_per_core=192, steps=200, form="dense")
            line["cpu_baseline"] = {
                "value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                "sample": f"{work} filter-steps ({cores} processes x 192 filters x 200 steady-state steps) of the same "
                          "workload through the dense numpy restatement of replay_no_ros.py:269-482 "
                          f"(oracle port), OPENBLAS_NUM_THREADS=1, {wall:.1f} s wall"}
        if not args.no_large_map:
            line["large_map"] = bench_large_map()
            if args.workload == "B":
                del f, steady                      # 1 GB of covariance back before the 3.4 GB configuration
                torch.cuda.empty_cache()
                line["small_filters"] = bench_config_m(local)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--workload", default="B", choices=["B", "M"],
                    help="B: 65,536 x 20 lm per GPU (contract line); M: 1,048,576 x 8 lm split over the GPUs")
    ap.add_argument("--joseph", action="store_true", help="one-pass Joseph-form covariance update")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-large-map", action="store_true")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    set_workload(args.workload, world)
    if args.impl == "reference":
        run_reference_arm(args, int(os.environ.get("RANK", "0")), world)
        return
    run_own_arm(args, world)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
