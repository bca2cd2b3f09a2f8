"""Drop-in Python call surface of the reference's EKF hot path, backed by libekf_b200.so (sm_100a CUDA).

This is the module ``src/ekf_bindings.py`` of the reference was meant to be (it is an empty file there): the
five module-level functions keep the reference's names, positional order and return order, so ``replay()``
(``src/replay_no_ros.py:66-248``) drives them unchanged once its globals are rebound (see ``replay.py``):

    delta_phi(ticks, prev_ticks, resolution) -> float                      replay_no_ros.py:250-266
    displacement(R, baseline, dphi_left, dphi_right) -> (angular, linear)  replay_no_ros.py:484-497
    EKF_pose_estimation(ang, lin, mean, cov, delta_t, pairs, TAG_INDEX)
        -> (mean, covariance, tags_positions)                              replay_no_ros.py:269-482
    predict(state, covariance, control, dt) -> (state, covariance)         EKF-SLAM.py:29-56
    update(state, covariance, observation, landmark_pos) -> (state, cov)   EKF-SLAM.py:59-84

plus the handle classes that keep state resident in HBM: ``BatchedEKF`` (B filters, fused step kernel) and
``LargeMapEKF`` (one filter, streamed covariance, optional row-block shard).

PyTorch is used only to own device memory and streams. There is NO CPU fallback: every compute call raises
if the CUDA library or a device is missing.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import capi

# --- module constants the reference reads at call time (replay_no_ros.py:15-36) ---------------------------
MOTION_MODEL_VARIANCE = 0.1
MEASUREMENT_MODEL_VARIANCE = 0.7
ENABLE_MEASUREMENT_MODEL = True
ENABLE_CIRCULAR_INTERPOLATION = True
DISABLE_MOTION_MODEL = False
IGNORE_TAGS = []
JOSEPH_FORM = False          # False = plain (I-KH)P like the reference (:480); True = one-pass Joseph

_flag_source = None          # module whose globals override the constants above (set by replay.py)


def use_flags_from(module) -> None:
    """Read the flag constants from ``module``'s globals at call time, as the reference does."""
    global _flag_source
    _flag_source = module


def _flag(name):
    if _flag_source is not None and hasattr(_flag_source, name):
        return getattr(_flag_source, name)
    return globals()[name]


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise capi.EkfLibraryError("no CUDA device: the EKF core has no CPU fallback")
    return torch


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream_ptr(torch):
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---------------------------------------------------------------------------------------------------------
# a1 / a2
# ---------------------------------------------------------------------------------------------------------
def delta_phi(ticks: int, prev_ticks: int, resolution: int) -> float:
    return float(capi.lib().ekf_delta_phi(int(ticks), int(prev_ticks), int(resolution)))


def displacement(R: float, baseline: float, delta_phi_left: float, delta_phi_right: float):
    ang, lin = C.c_double(), C.c_double()
    capi.call("ekf_displacement", float(R), float(baseline), float(delta_phi_left), float(delta_phi_right),
              C.byref(ang), C.byref(lin))
    return ang.value, lin.value


# ---------------------------------------------------------------------------------------------------------
# a3 / a4 — host bookkeeping for the single-filter call surface (integers, dict order: bit-exact by
# construction). The batched path has its own device kernel for perturbed detections.
# ---------------------------------------------------------------------------------------------------------
def associate_window(mean, timestamp_detectedTags_pair_list, TAG_INDEX, ignore_tags=()):
    """Gate, slot assignment (mutates TAG_INDEX), window average, range/bearing and global guess
    (replay_no_ros.py:280-337). Returns dict slot -> [gx, gy, err, tag_id, range, bearing] in
    first-seen-in-window order."""
    buckets = {}
    for _ts, dets in timestamp_detectedTags_pair_list:
        for tag in dets:
            if tag.tag_id in ignore_tags:
                continue
            tx, tz = float(tag.pose_t[0][0]), float(tag.pose_t[2][0])
            if tz * tz + tx * tx > 2.25:
                continue
            if tag.tag_id not in TAG_INDEX:
                TAG_INDEX[tag.tag_id] = len(TAG_INDEX)
            acc = buckets.setdefault(TAG_INDEX[tag.tag_id], [0.0, 0.0, 0.0, 0])
            acc[0] += tx                 # sequential sums == np.mean along the window axis
            acc[1] += tz
            acc[2] += float(tag.pose_err)
            acc[3] += 1
    inv = {slot: tid for tid, slot in TAG_INDEX.items()}
    px, py, pth = float(mean[0]), float(mean[1]), float(mean[2])
    out = {}
    for slot, (sx, sz, se, k) in buckets.items():
        x_rel, y_rel = sz / k, -(sx / k)
        rng = math.sqrt(x_rel * x_rel + y_rel * y_rel)
        brg = math.atan2(y_rel, x_rel)
        out[slot] = [px + rng * math.cos(brg + pth), py + rng * math.sin(brg + pth), se / k, inv[slot], rng, brg]
    return out


# ---------------------------------------------------------------------------------------------------------
# large-map / single-filter handle
# ---------------------------------------------------------------------------------------------------------
class LargeMapEKF:
    """One EKF-SLAM filter with its state resident on one GPU (or one row block of it).

    ``capacity`` landmarks are pre-allocated (n = 3 + 2*capacity, leading dimension padded to 16 doubles);
    landmarks not yet augmented have zero rows/columns, which leaves every result identical to the
    reference's growing arrays.
    """

    def __init__(self, capacity: int, device: int = 0, row_range=None):
        torch = _torch()
        self.torch = torch
        self.capacity = int(capacity)
        self.device = int(device)
        n, ld = C.c_int(), C.c_int64()
        capi.call("ekf_lm_layout", self.capacity, C.byref(n), C.byref(ld))
        self.n, self.ld = n.value, ld.value
        self.row_begin, self.row_end = (0, self.n) if row_range is None else map(int, row_range)
        dev = torch.device("cuda", self.device)
        self.mu = torch.zeros(self.ld, dtype=torch.float64, device=dev)
        self.P = torch.zeros((self.row_end - self.row_begin, self.ld), dtype=torch.float64, device=dev)
        h = C.c_void_p()
        capi.call("ekf_lm_create", self.capacity, self.device, C.byref(h))
        self._h = h
        capi.call("ekf_lm_bind_state", h, _ptr(self.mu), _ptr(self.P), self.ld, self.row_begin, self.row_end)
        self.rows5 = None

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().ekf_lm_destroy(self._h)
            self._h = None

    def __del__(self):
        if capi is not None:        # module globals are already gone at interpreter exit: the process frees the rest
            self.close()

    # -- configuration ----------------------------------------------------------------------------------
    def set_noise(self, motion_var: float, meas_var: float):
        capi.call("ekf_lm_set_noise", self._h, float(motion_var), float(meas_var))

    def set_flags(self, circular=True, measurement=True, disable_motion=False, joseph=False):
        capi.call("ekf_lm_set_flags", self._h, int(circular), int(measurement), int(disable_motion),
                  capi.EKF_COV_JOSEPH if joseph else capi.EKF_COV_PLAIN)

    # -- state transfer (reference layout <-> device layout) ---------------------------------------------
    def load(self, mean: np.ndarray, cov: np.ndarray):
        """Upload a reference-layout state of size m <= n (rows of this block only): ONE C-ABI call, the rest of the
        capacity is zeroed on the device."""
        mean = np.ascontiguousarray(mean, dtype=np.float64)
        cov = np.ascontiguousarray(cov, dtype=np.float64)
        m = len(mean)
        if m > self.n or cov.shape != (m, m):
            raise ValueError("state larger than capacity or inconsistent shapes")
        capi.call("ekf_lm_upload", self._h, m, mean.ctypes.data, cov.ctypes.data, _stream_ptr(self.torch))
        self.torch.cuda.current_stream().synchronize()        # the host arrays may be released by the caller

    def state(self, m: int | None = None):
        """Download (mean[:m], cov block[:, :m]) as NEW numpy float64 arrays: ONE C-ABI call."""
        m = self.n if m is None else int(m)
        hi = min(self.row_end, m)
        mu = np.empty(m)
        blk = np.empty((max(hi - self.row_begin, 0), m))
        capi.call("ekf_lm_download", self._h, m, mu.ctypes.data, blk.ctypes.data, _stream_ptr(self.torch))
        return mu, blk

    # -- steps ------------------------------------------------------------------------------------------
    def augment(self, slot: int, gx: float, gy: float):
        capi.call("ekf_lm_augment", self._h, int(slot), float(gx), float(gy), _stream_ptr(self.torch))

    def predict(self, ang: float, lin: float):
        capi.call("ekf_lm_predict", self._h, float(ang), float(lin), _stream_ptr(self.torch))

    def update(self, slot: int, rng: float, brg: float):
        capi.call("ekf_lm_update", self._h, int(slot), float(rng), float(brg), _stream_ptr(self.torch))

    def update_multi(self, slots, rngs, brgs):
        """The step's sequential updates (replay_no_ros.py:436-480, given order) in ONE pass over P; bitwise equal
        to calling update() once per observation."""
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        rg = np.ascontiguousarray(rngs, dtype=np.float64)
        bg = np.ascontiguousarray(brgs, dtype=np.float64)
        assert sl.shape == rg.shape == bg.shape and sl.ndim == 1
        capi.call("ekf_lm_update_multi", self._h, int(sl.size), sl.ctypes.data, rg.ctypes.data, bg.ctypes.data,
                  _stream_ptr(self.torch))

    def export_rows_multi(self, slots, rows=None):
        """Fused step on a row-block shard: this shard's share of rows {0,1,2} U {j_u, j_u+1} into a
        (3 + 2m, ld + 4) device buffer (created when not given)."""
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        if rows is None:
            rows = self.torch.zeros((3 + 2 * sl.size, self.ld + 4), dtype=self.torch.float64, device=self.mu.device)
        capi.call("ekf_lm_export_rows_multi", self._h, int(sl.size), sl.ctypes.data, _ptr(rows),
                  _stream_ptr(self.torch))
        return rows

    def update_multi_from_rows(self, slots, rngs, brgs, rows):
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        rg = np.ascontiguousarray(rngs, dtype=np.float64)
        bg = np.ascontiguousarray(brgs, dtype=np.float64)
        capi.call("ekf_lm_update_multi_from_rows", self._h, int(sl.size), sl.ctypes.data, rg.ctypes.data,
                  bg.ctypes.data, _ptr(rows), _stream_ptr(self.torch))

    def export_rows(self, slot: int):
        if self.rows5 is None:
            self.rows5 = self.torch.zeros((5, self.ld + 4), dtype=self.torch.float64, device=self.mu.device)
        capi.call("ekf_lm_export_rows", self._h, int(slot), _ptr(self.rows5), _stream_ptr(self.torch))
        return self.rows5

    def update_from_rows(self, slot: int, rng: float, brg: float, rows5=None):
        rows5 = self.rows5 if rows5 is None else rows5
        capi.call("ekf_lm_update_from_rows", self._h, int(slot), float(rng), float(brg), _ptr(rows5),
                  _stream_ptr(self.torch))

    # ---- row exchange over NVLink peer memory (row-block shards, one process per GPU) ----
    P2P_HANDLE_BYTES = 64

    def p2p_init(self, world: int, rank: int) -> bytes:
        """Allocate this shard's exchange area; returns its cudaIpc handle (to be gathered from every rank)."""
        buf = (C.c_ubyte * self.P2P_HANDLE_BYTES)()
        capi.call("ekf_lm_p2p_init", self._h, int(world), int(rank), buf, self.P2P_HANDLE_BYTES)
        return bytes(buf)

    def p2p_area(self) -> int:
        v = C.c_void_p(0)
        capi.call("ekf_lm_p2p_area", self._h, C.byref(v))
        return int(v.value or 0)

    def p2p_connect(self, handles=None, local_areas=None):
        """handles: list of ``world`` handle byte strings (other processes); local_areas: list of device pointers
        (ints, 0 = use the handle) for shards living in this process."""
        hb = None
        if handles is not None:
            blob = b"".join(bytes(h).ljust(self.P2P_HANDLE_BYTES, b"\0") for h in handles)
            hb = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        la = None
        if local_areas is not None:
            la = (C.c_void_p * len(local_areas))(*[C.c_void_p(int(a) or None) for a in local_areas])
        capi.call("ekf_lm_p2p_connect", self._h, hb, self.P2P_HANDLE_BYTES, la)

    def p2p_update(self, slot: int, rng: float, brg: float, owner: int):
        capi.call("ekf_lm_p2p_update", self._h, int(slot), float(rng), float(brg), int(owner),
                  _stream_ptr(self.torch))

    def p2p_update_multi(self, slots, rngs, brgs, owners):
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        rg = np.ascontiguousarray(rngs, dtype=np.float64)
        bg = np.ascontiguousarray(brgs, dtype=np.float64)
        ow = np.ascontiguousarray(owners, dtype=np.int32)
        capi.call("ekf_lm_p2p_update_multi", self._h, int(sl.size), sl.ctypes.data, rg.ctypes.data, bg.ctypes.data,
                  ow.ctypes.data, _stream_ptr(self.torch))

    def p2p_invalidate(self):
        """P was written from outside the library (torch): the robot rows pre-pushed by the last update are stale."""
        capi.call("ekf_lm_p2p_invalidate", self._h)

    def p2p_error(self) -> int:
        v = C.c_int64(0)
        capi.call("ekf_lm_p2p_error", self._h, C.byref(v))
        return int(v.value)

    def sync(self):
        capi.call("ekf_lm_sync", self._h)


_scalar_cache = {}
_scalar_last = {"filter": None}     # what the last EKF_pose_estimation call returned (device-resident fast path)


def _scalar_filter(n_landmarks: int) -> LargeMapEKF:
    cap = max(4, 1 << max(0, (n_landmarks - 1)).bit_length())   # round capacity up to a power of two
    f = _scalar_cache.get(cap)
    if f is None:
        f = _scalar_cache[cap] = LargeMapEKF(cap)
    return f


def _still_resident(f, mean_in, cov_in) -> bool:
    """True when (mean_in, cov_in) are the very arrays the previous call returned from filter ``f`` and still hold
    what was returned: the state is then already on the device and the upload (2 x 537 MB of PCIe per call at 4096
    landmarks) is skipped. Identity of the two array objects, plus a content check of the whole mean and of a strided
    sample of the covariance against copies kept from the previous call (an in-place edit by the caller falls back
    to the upload)."""
    last = _scalar_last
    if last["filter"] is not f or mean_in is not last.get("mean") or cov_in is not last.get("cov"):
        return False
    if not np.array_equal(mean_in, last["mean_copy"]):
        return False
    return np.array_equal(cov_in.reshape(-1)[last["sample_idx"]], last["cov_sample"])


def EKF_pose_estimation(angular_displacement, linear_displacement, motion_model_mean, motion_model_covariance,
                        delta_t, timestamp_detectedTags_pair_list, TAG_INDEX):
    """Drop-in for replay_no_ros.py:269-482: NEW (mean, covariance) arrays come back, TAG_INDEX is mutated in
    place, ``tags_positions`` is the reference's dict of 6-element lists. ``delta_t`` is unused (as there).
    When the caller hands back exactly what the previous call returned (what ``replay()`` does, :229), the state is
    still resident on the device and is not uploaded again."""
    if isinstance(motion_model_mean, np.ndarray) and motion_model_mean.dtype == np.float64:
        mean_in = motion_model_mean
    else:
        mean_in = np.asarray(motion_model_mean, dtype=np.float64)
    if isinstance(motion_model_covariance, np.ndarray) and motion_model_covariance.dtype == np.float64:
        cov_in = motion_model_covariance
    else:
        cov_in = np.asarray(motion_model_covariance, dtype=np.float64)
    tags = associate_window(mean_in, timestamp_detectedTags_pair_list, TAG_INDEX, tuple(_flag("IGNORE_TAGS")))
    n_old = len(mean_in)
    n_new = max(n_old, 3 + 2 * len(TAG_INDEX))
    f = _scalar_filter((n_new - 3) // 2)
    f.set_noise(_flag("MOTION_MODEL_VARIANCE"), _flag("MEASUREMENT_MODEL_VARIANCE"))
    f.set_flags(_flag("ENABLE_CIRCULAR_INTERPOLATION"), _flag("ENABLE_MEASUREMENT_MODEL"),
                _flag("DISABLE_MOTION_MODEL"), JOSEPH_FORM)
    if not _still_resident(f, mean_in, cov_in):
        f.load(mean_in, cov_in)
    for i in range(n_old, n_new, 2):                      # :355-360 (KeyError like the reference at :359)
        slot = (i - 3) // 2
        f.augment(slot, tags[slot][0], tags[slot][1])
    f.predict(float(angular_displacement), float(linear_displacement))
    if tags:                                              # first-seen-in-window order (:436), one pass over P
        f.update_multi(list(tags.keys()), [rec[4] for rec in tags.values()], [rec[5] for rec in tags.values()])
    mean, cov = f.state(n_new)
    idx = np.linspace(0, cov.size - 1, num=min(cov.size, 257), dtype=np.int64)
    _scalar_last.update(filter=f, mean=mean, cov=cov, mean_copy=mean.copy(), sample_idx=idx,
                        cov_sample=cov.reshape(-1)[idx].copy())
    return mean, cov, tags


# ---------------------------------------------------------------------------------------------------------
# batched handle
# ---------------------------------------------------------------------------------------------------------
class BatchedEKF:
    """B independent EKF-SLAM filters of fixed landmark capacity, state resident in HBM (SoA records)."""

    def __init__(self, batch: int, n_landmarks: int, device: int = 0):
        torch = _torch()
        self.torch = torch
        self.batch, self.n_landmarks, self.device = int(batch), int(n_landmarks), int(device)
        n, ld, ps, ms = C.c_int(), C.c_int(), C.c_int64(), C.c_int64()
        capi.call("ekf_batch_layout", self.n_landmarks, C.byref(n), C.byref(ld), C.byref(ps), C.byref(ms))
        self.n, self.ld, self.p_stride, self.mu_stride = n.value, ld.value, ps.value, ms.value
        dev = torch.device("cuda", self.device)
        self.dev = dev
        self.mu = torch.zeros((self.batch, self.mu_stride), dtype=torch.float64, device=dev)
        self.P = torch.zeros((self.batch, self.p_stride), dtype=torch.float64, device=dev)
        self.n_known = torch.zeros(self.batch, dtype=torch.int32, device=dev)
        self.motion_var = torch.full((self.batch,), MOTION_MODEL_VARIANCE, dtype=torch.float64, device=dev)
        self.meas_var = torch.full((self.batch,), MEASUREMENT_MODEL_VARIANCE, dtype=torch.float64, device=dev)
        h = C.c_void_p()
        capi.call("ekf_batch_create", self.n_landmarks, self.batch, self.device, C.byref(h))
        self._h = h
        capi.call("ekf_batch_bind_state", h, _ptr(self.mu), _ptr(self.P), _ptr(self.n_known))
        capi.call("ekf_batch_set_noise", h, _ptr(self.motion_var), _ptr(self.meas_var))

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().ekf_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        if capi is not None:        # module globals are already gone at interpreter exit: the process frees the rest
            self.close()

    def set_noise(self, motion_var, meas_var):
        t = self.torch
        self.motion_var.copy_(t.as_tensor(np.broadcast_to(motion_var, (self.batch,)).copy(), dtype=t.float64))
        self.meas_var.copy_(t.as_tensor(np.broadcast_to(meas_var, (self.batch,)).copy(), dtype=t.float64))

    def set_flags(self, circular=True, measurement=True, disable_motion=False, joseph=False):
        capi.call("ekf_batch_set_flags", self._h, int(circular), int(measurement), int(disable_motion),
                  capi.EKF_COV_JOSEPH if joseph else capi.EKF_COV_PLAIN)

    def init(self, p0_diag: float = MOTION_MODEL_VARIANCE):
        """mean = 0, covariance = p0 * I3 (replay_no_ros.py:69-70), no landmark known."""
        capi.call("ekf_batch_init", self._h, float(p0_diag), _stream_ptr(self.torch))

    def step(self, ang, lin, slot=None, rng=None, brg=None):
        """One fused filter-step; arguments are DEVICE tensors: ang/lin (B,), slot int32 / rng / brg (B, m)."""
        m = 0 if slot is None else int(slot.shape[1])
        z = C.c_void_p(0)
        capi.call("ekf_batch_step", self._h, _ptr(ang), _ptr(lin), _ptr(slot) if m else z,
                  _ptr(rng) if m else z, _ptr(brg) if m else z, m, _stream_ptr(self.torch))

    def step_ticks(self, ticks_l, ticks_r, prev_l, prev_r, wheel_radius, baseline, resolution, slot=None,
                   rng=None, brg=None):
        m = 0 if slot is None else int(slot.shape[1])
        z = C.c_void_p(0)
        capi.call("ekf_batch_step_ticks", self._h, _ptr(ticks_l), _ptr(ticks_r), _ptr(prev_l), _ptr(prev_r),
                  float(wheel_radius), float(baseline), int(resolution), _ptr(slot) if m else z,
                  _ptr(rng) if m else z, _ptr(brg) if m else z, m, _stream_ptr(self.torch))

    def step_ticks_common(self, ticks_l, ticks_r, prev_l, prev_r, wheel_radius, baseline, resolution, slot=None,
                          rng=None, brg=None):
        """``step_ticks`` when every filter sees the same encoder ticks (four Python ints, no per-filter tensors)."""
        m = 0 if slot is None else int(slot.shape[1])
        z = C.c_void_p(0)
        capi.call("ekf_batch_step_ticks_common", self._h, int(ticks_l), int(ticks_r), int(prev_l), int(prev_r),
                  float(wheel_radius), float(baseline), int(resolution), _ptr(slot) if m else z,
                  _ptr(rng) if m else z, _ptr(brg) if m else z, m, _stream_ptr(self.torch))

    def new_slot_table(self, n_tags: int):
        """Per-filter TAG_INDEX on the device: (tag_slot int32 [n_tags, B] filled with -1, n_seen int32 [B])."""
        t = self.torch
        return (t.full((int(n_tags), self.batch), -1, dtype=t.int32, device=self.dev),
                t.zeros(self.batch, dtype=t.int32, device=self.dev))

    def associate(self, det_tag, det_tx, det_tz, tag_slot, n_seen, m, noise_tx=None, noise_tz=None, generated=None,
                  out=None):
        """a3/a4 on the device for every filter (replay_no_ros.py:280-337): the window's detections — compact tag
        index int32 (D,), pose_t[0] and pose_t[2] (D,), optional per-filter perturbations (D, B) — become the
        (slot, rng, brg) tensors of shape (B, m) that ``step`` takes; ``tag_slot`` / ``n_seen`` are updated in
        place. All arguments are DEVICE tensors. ``generated=(seed, step, first_filter, sigma)``: the perturbations
        are generated inside the kernel (the numbers ``detection_noise`` would write) instead of read from arrays.
        ``out=(slot, rng, brg)`` reuses output tensors."""
        t = self.torch
        d = int(det_tag.shape[0])
        if out is None:
            slot = t.empty((self.batch, m), dtype=t.int32, device=self.dev)
            rng = t.empty((self.batch, m), dtype=t.float64, device=self.dev)
            brg = t.empty((self.batch, m), dtype=t.float64, device=self.dev)
        else:
            slot, rng, brg = out
        z = C.c_void_p(0)
        if generated is not None:
            seed, step, first, sigma = generated
            capi.call("ekf_batch_associate_generated", self._h, d, _ptr(det_tag) if d else z, _ptr(det_tx) if d else z,
                      _ptr(det_tz) if d else z, int(seed), int(step), int(first), float(sigma),
                      int(tag_slot.shape[0]), _ptr(tag_slot), _ptr(n_seen), _ptr(slot), _ptr(rng), _ptr(brg), int(m),
                      _stream_ptr(t))
            return slot, rng, brg
        capi.call("ekf_batch_associate", self._h, d, _ptr(det_tag) if d else z, _ptr(det_tx) if d else z,
                  _ptr(det_tz) if d else z, _ptr(noise_tx) if noise_tx is not None else z,
                  _ptr(noise_tz) if noise_tz is not None else z, int(tag_slot.shape[0]), _ptr(tag_slot),
                  _ptr(n_seen), _ptr(slot), _ptr(rng), _ptr(brg), int(m), _stream_ptr(t))
        return slot, rng, brg

    def associate_overflow_count(self) -> int:
        v = C.c_int64(0)
        capi.call("ekf_batch_associate_overflow_count", self._h, C.byref(v))
        return int(v.value)

    def accumulate_ate(self, gt_x: float, gt_y: float, sse):
        """sse[b] += squared position error of filter b against the ground-truth point (device tensor (B,))."""
        capi.call("ekf_batch_accumulate_ate", self._h, float(gt_x), float(gt_y), _ptr(sse), _stream_ptr(self.torch))

    def moments(self, out=None):
        """(sum_b mu_b, sum_b mu_b mu_b^T) over this handle's filters as one device tensor of n + n*n doubles
        (fixed summation order). Feed to ``ensemble.allreduce_moments`` for the ensemble mean / covariance."""
        n = self.n
        if out is None:
            out = self.torch.empty(n + n * n, dtype=self.torch.float64, device=self.dev)
        capi.call("ekf_batch_moments", self._h, _ptr(out), _stream_ptr(self.torch))
        return out

    def pose_moments(self, out=None):
        """(sum_b pose_b, sum_b pose_b pose_b^T) of the robot pose over this handle's filters: 12 doubles on the device."""
        if out is None:
            out = self.torch.empty(12, dtype=self.torch.float64, device=self.dev)
        capi.call("ekf_batch_pose_moments", self._h, _ptr(out), _stream_ptr(self.torch))
        return out

    # ---- Monte Carlo perturbations generated on the device (counter-based: seed, global filter index, step) ----
    @staticmethod
    def mc_record(ang, lin, slots=(), rngs=(), brgs=(), out=None):
        """Pack one step's COMMON inputs: doubles [ang, lin, rng[0..m), brg[0..m)] followed by int32 slot[0..m).
        Returns a numpy float64 array (or fills ``out``, e.g. a pinned torch tensor's numpy view)."""
        m = len(slots)
        n = 2 + 2 * m + (m + 1) // 2
        rec = np.zeros(n) if out is None else out
        assert rec.dtype == np.float64 and rec.size >= n
        rec[0], rec[1] = float(ang), float(lin)
        rec[2:2 + m] = np.asarray(rngs, dtype=np.float64)
        rec[2 + m:2 + 2 * m] = np.asarray(brgs, dtype=np.float64)
        rec[2 + 2 * m:n].view(np.int32)[:m] = np.asarray(slots, dtype=np.int32)
        return rec

    def perturb(self, seed, step, record_dev, m, sigma4, first_filter=0):
        """Per-filter (ang, lin, slot, rng, brg) DEVICE tensors for ``step`` from a device-resident common record."""
        t = self.torch
        ang = t.empty(self.batch, dtype=t.float64, device=self.dev)
        lin = t.empty(self.batch, dtype=t.float64, device=self.dev)
        slot = t.empty((self.batch, m), dtype=t.int32, device=self.dev)
        rng = t.empty((self.batch, m), dtype=t.float64, device=self.dev)
        brg = t.empty((self.batch, m), dtype=t.float64, device=self.dev)
        sg = (C.c_double * 4)(*[float(x) for x in sigma4])
        z = C.c_void_p(0)
        capi.call("ekf_batch_perturb", self._h, int(seed), int(step), int(first_filter), int(m), _ptr(record_dev), sg,
                  _ptr(ang), _ptr(lin), _ptr(slot) if m else z, _ptr(rng) if m else z, _ptr(brg) if m else z,
                  _stream_ptr(t))
        return ang, lin, slot, rng, brg

    def step_generated(self, seed, step, record_dev, m, sigma4, first_filter=0):
        """One fused filter-step whose per-member inputs are drawn inside the kernel from the device-resident common
        record (bitwise the step of ``perturb`` + ``step``, without the per-filter arrays)."""
        sg = (C.c_double * 4)(*[float(x) for x in sigma4])
        capi.call("ekf_batch_step_generated", self._h, int(seed), int(step), int(first_filter), int(m),
                  _ptr(record_dev), sg, _stream_ptr(self.torch))

    def detection_noise(self, seed, step, n_det, sigma, first_filter=0):
        """(noise_tx, noise_tz) device tensors (n_det, B) for ``associate``."""
        t = self.torch
        ntx = t.empty((int(n_det), self.batch), dtype=t.float64, device=self.dev)
        ntz = t.empty((int(n_det), self.batch), dtype=t.float64, device=self.dev)
        if n_det:
            capi.call("ekf_batch_detection_noise", self._h, int(seed), int(step), int(first_filter), int(n_det),
                      float(sigma), _ptr(ntx), _ptr(ntz), _stream_ptr(t))
        return ntx, ntz

    def mc_step_host_async(self, seed, step, record_host, m, sigma4, first_filter=0, gt=None, sse=None,
                           pose_moments_out=None):
        """One pipelined Monte Carlo step from a HOST record (``mc_record``; ideally pinned): the perturbations are
        generated on the device, the fused step runs, ``sse`` (device, (B,)) accumulates the squared position error
        against ``gt`` when given, and the 12 pose moments land in ``pose_moments_out`` (host, pinned) after
        ``host_sync()``."""
        def hp(a):
            return C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        sg = (C.c_double * 4)(*[float(x) for x in sigma4])
        z = C.c_void_p(0)
        score = gt is not None and sse is not None
        capi.call("ekf_batch_mc_step_host_async", self._h, int(seed), int(step), int(first_filter), int(m),
                  hp(record_host), sg, int(score), float(gt[0]) if score else 0.0, float(gt[1]) if score else 0.0,
                  _ptr(sse) if score else z, hp(pose_moments_out) if pose_moments_out is not None else z,
                  _stream_ptr(self.torch))

    def step_host(self, ang, lin, slot, rng, brg, pose_out):
        """End-to-end step with HOST (ideally pinned) torch tensors / numpy arrays: H2D, step, D2H of the
        pose (B, 3) into ``pose_out``; synchronous."""
        def hp(a):
            return C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        m = 0 if slot is None else int(slot.shape[1])
        z = C.c_void_p(0)
        capi.call("ekf_batch_step_host", self._h, hp(ang), hp(lin), hp(slot) if m else z, hp(rng) if m else z,
                  hp(brg) if m else z, m, hp(pose_out), _stream_ptr(self.torch))

    def step_host_async(self, ang, lin, slot, rng, brg, pose_out):
        """Pipelined ``step_host``: enqueue only. Keep the (pinned) host buffers alive and do not read
        ``pose_out`` until ``host_sync()``; consecutive calls overlap H2D, the fused step and D2H."""
        def hp(a):
            return C.c_void_p(a.data_ptr() if hasattr(a, "data_ptr") else a.ctypes.data)
        m = 0 if slot is None else int(slot.shape[1])
        z = C.c_void_p(0)
        capi.call("ekf_batch_step_host_async", self._h, hp(ang), hp(lin), hp(slot) if m else z,
                  hp(rng) if m else z, hp(brg) if m else z, m, hp(pose_out), _stream_ptr(self.torch))

    def host_sync(self):
        capi.call("ekf_batch_host_sync", self._h)

    def load(self, mean: np.ndarray, cov: np.ndarray, first: int = 0):
        """Upload reference-layout states: mean (k, n), cov (k, n, n) -> filters [first, first+k)."""
        t = self.torch
        mean = np.ascontiguousarray(mean, dtype=np.float64)
        cov = np.ascontiguousarray(cov, dtype=np.float64)
        k = mean.shape[0]
        assert mean.shape == (k, self.n) and cov.shape == (k, self.n, self.n)
        dm, dc = t.from_numpy(mean).to(self.dev), t.from_numpy(cov).to(self.dev)
        capi.call("ekf_batch_pack", self._h, _ptr(dm), _ptr(dc), int(first), int(k), _stream_ptr(t))
        t.cuda.current_stream().synchronize()

    def state(self, first: int = 0, count: int | None = None):
        """Download reference-layout (mean (k, n), cov (k, n, n)) of filters [first, first+count)."""
        t = self.torch
        count = self.batch - first if count is None else count
        dm = t.empty((count, self.n), dtype=t.float64, device=self.dev)
        dc = t.empty((count, self.n, self.n), dtype=t.float64, device=self.dev)
        capi.call("ekf_batch_unpack", self._h, _ptr(dm), _ptr(dc), int(first), int(count), _stream_ptr(t))
        return dm.cpu().numpy(), dc.cpu().numpy()

    def fault_count(self) -> int:
        v = C.c_int64()
        capi.call("ekf_batch_fault_count", self._h, C.byref(v))
        return v.value

    def nonfinite_count(self) -> int:
        """Filters whose mu or P holds a non-finite value right now (a full scan on the device)."""
        v = C.c_int64()
        capi.call("ekf_batch_nonfinite_count", self._h, C.byref(v), _stream_ptr(self.torch))
        return v.value


# ---------------------------------------------------------------------------------------------------------
# 3-state surface (src/EKF-SLAM.py)
# ---------------------------------------------------------------------------------------------------------
def predict3_batched(state, cov, v, omega, dt):
    """In place on device tensors: state (3, B), cov (9, B), v / omega / dt (B,)."""
    torch = _torch()
    capi.call("ekf3_predict", int(state.shape[1]), _ptr(state), _ptr(cov), _ptr(v), _ptr(omega), _ptr(dt),
              _stream_ptr(torch))


def update3_batched(state, cov, rng, brg, lx, ly):
    torch = _torch()
    capi.call("ekf3_update", int(state.shape[1]), _ptr(state), _ptr(cov), _ptr(rng), _ptr(brg), _ptr(lx),
              _ptr(ly), _stream_ptr(torch))


def _dev1(torch, x):
    return torch.tensor([float(x)], dtype=torch.float64, device="cuda")


def predict(state, covariance, control, dt):
    """Drop-in for EKF-SLAM.py:29-56 (returns new arrays)."""
    torch = _torch()
    st = torch.tensor(np.asarray(state, dtype=np.float64).reshape(3, 1), device="cuda")
    cv = torch.tensor(np.asarray(covariance, dtype=np.float64).reshape(9, 1), device="cuda")
    predict3_batched(st, cv, _dev1(torch, control[0]), _dev1(torch, control[1]), _dev1(torch, dt))
    return st.cpu().numpy().reshape(3), cv.cpu().numpy().reshape(3, 3)


def update(state, covariance, observation, landmark_pos):
    """Drop-in for EKF-SLAM.py:59-84 (returns new arrays)."""
    torch = _torch()
    st = torch.tensor(np.asarray(state, dtype=np.float64).reshape(3, 1), device="cuda")
    cv = torch.tensor(np.asarray(covariance, dtype=np.float64).reshape(9, 1), device="cuda")
    update3_batched(st, cv, _dev1(torch, observation[0]), _dev1(torch, observation[1]),
                    _dev1(torch, landmark_pos[0]), _dev1(torch, landmark_pos[1]))
    return st.cpu().numpy().reshape(3), cv.cpu().numpy().reshape(3, 3)
