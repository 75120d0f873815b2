#!/usr/bin/env python
"""bench.py — ORB extract+match frames/s (640x480, 1000 kp) with the REFERENCE'S OWN matcher on both arms:
brute-force Hamming 2-NN + ratio 0.8 (core/feature/orb_matcher.cpp:22-36), one rank per B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path (8-level pyramid, FAST+NMS, retainBest, Harris, IC angle, 7x7 blur,
rBRIEF, Hamming 2-NN + ratio test of every frame against its predecessor) over `--chunks` consecutive
chunks of `--batch` frames of a synthetic streaming sequence (default 8 x 128 = 1024 frames per step).

  value  frames/s with the chunk already resident in HBM (vxorb_resident_run), CUDA events on the library's
         own streams, max over ranks.  Passes are pipelined over the handle's pipeline slots exactly as the
         streaming path pipelines chunks; nothing crosses PCIe.  The working set of a chunk (pyramid + blurred
         pyramid + lists, ~2.3 MB/frame) is larger than the 126 MB L2.
  e2e    the same metric through the public streaming call (vxorb_track_submit/_wait) with pinned HOST frames
         in and keypoints, descriptors and matches copied back to the host inside the timed region.
  roofline / stages: per-stage CUDA-event times of a separate, serial instrumented pass over the same chunk,
         each against its algorithmic bytes and MEASURED_PEAKS.json.
  windowed: the same two numbers with BASELINE config 2's windowed (r = 32 px) matcher, as extra keys.
  per_call / c3 / c4: BASELINE configs 1, 3 and 4 measured in the same run (rank 0, N = 1 only).
  parity_checked: after the timed region, frames of the resident batch (keypoints, descriptors, match rows)
         are compared with the CPU oracle bit for bit; a mismatch aborts the run.

`--impl reference` times the reference's own CPU path (cv2 = the OpenCV the reference's ORBExtractor /
ORBMatcher wrap, orb_extractor.cpp:13, orb_matcher.cpp:22-36) on the host cores, same workload and matcher;
if cv2 is missing it falls back to the C oracle port.  No file under /root/reference is read at run time.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

W, H, NKP, RADIUS = 640, 480, 1000, 32.0
METRIC = "ORB extract+match frames/s (640x480, 1k kp)"
WORKLOAD = "streaming 640x480 TUM-shape synthetic sequence, 1000 kp/frame, 8 levels x1.2, frame-to-frame " \
           "brute-force Hamming 2-NN + ratio 0.8 (the reference's matcher, orb_matcher.cpp:22-36)"


def make_config(batch, chunks):
    """Identical on both arms (the driver compares it)."""
    return {"workload": WORKLOAD, "matcher": "bruteforce", "frames_per_step_per_gpu": batch * chunks,
            "chunk_frames": batch, "chunks_per_step": chunks,
            "sharding": "one independent sequence per GPU, no collective",
            "l2_policy": f"inputs larger than L2: {batch} frames x ~2.3 MB working set per chunk"}


# ---- algorithmic bytes per frame (SURVEY.md 8d staged-minimum model), computed from the geometry ----

def level_sizes(w=W, h=H, n_levels=8, sf=1.2):
    sf = float(np.float32(sf))
    out = []
    for l in range(n_levels):
        inv = np.float32(1.0) / np.float32(sf ** l)
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(n_kp=NKP, w=W, h=H):
    px = [a * b for a, b in level_sizes(w, h)]
    s = sum(px)
    return {
        "pyramid": (s - px[-1]) + (s - px[0]),   # read every source level + write levels 1..7
        "fast": s,                                # one read of every level pixel
        "select": 0,                              # list work, no image traffic in the model
        "blur": 2 * s,                            # read + write every level
        "describe": n_kp * 512 + n_kp * (32 + 24),  # 512 gathers + descriptor + keypoint record
        "match": 0,                               # integer-pipe bound: popc ops, not bytes
    }


EXTRACT_STAGES = ("pyramid", "fast", "select", "blur", "describe")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- clocks sampled during the timed region (pynvml; nvidia-smi fields of the profiling recipe) ----

class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._dev = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._dev, pynvml.NVML_CLOCK_SM))
        except Exception as e:  # no NVML: report that instead of inventing clocks
            self._nv = None
            self.error = repr(e)

    def _run(self):
        nv = self._nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._dev, nv.NVML_CLOCK_SM)))
                r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._dev))
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.004)

    def start(self):
        if self._nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join()
        if not self._nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "error": self.error}
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ---- reference arm: the reference's own CPU path (OpenCV through cv2), all host threads ----------

def cpu_reference(frames, seconds_target, threads, windowed):
    """Times cv::ORB::detectAndCompute + BFMatcher.knnMatch(k=2) + ratio (orb_extractor.cpp:13,
    orb_matcher.cpp:22-36) over consecutive frames of the sequence: `threads` workers, each with
    OpenCV's internal pool set to 1 thread, take disjoint sub-sequences (frame-level parallelism is
    how the CPU path uses many cores: cv::ORB itself is serial).  Returns (frames/s, frames, s, kind)."""
    import concurrent.futures as cf
    try:
        import cv2
        cv2.setNumThreads(1)
        kind = "reference"

        def run(sub):
            orb = cv2.ORB_create(NKP, 1.2, 8)
            bf = cv2.BFMatcher(cv2.NORM_HAMMING)
            prev = None
            nm = 0
            for f in sub:
                kps, desc = orb.detectAndCompute(f, None)
                if prev is not None and desc is not None and prev[1] is not None:
                    mask = None
                    if windowed:
                        a = np.array([k.pt for k in prev[0]], np.float32)
                        b = np.array([k.pt for k in kps], np.float32)
                        mask = ((np.abs(a[:, None, 0] - b[None, :, 0]) <= RADIUS) &
                                (np.abs(a[:, None, 1] - b[None, :, 1]) <= RADIUS)).astype(np.uint8)
                    for row in bf.knnMatch(prev[1], desc, 2, mask=mask):
                        if len(row) >= 2 and row[0].distance < 0.8 * row[1].distance:
                            nm += 1
                prev = (kps, desc)
            return nm
    except ImportError:
        import oracle  # CPU port of the same algorithm (test infrastructure; only timed here)
        kind = "port"

        def run(sub):
            prev = None
            nm = 0
            for f in sub:
                kp, desc = oracle.orb_extract(f)
                if prev is not None:
                    qxy = np.stack([prev[0]["x"], prev[0]["y"]], 1) if windowed else None
                    txy = np.stack([kp["x"], kp["y"]], 1) if windowed else None
                    nm += len(oracle.match(prev[1], desc, 0.8, qxy, txy, RADIUS if windowed else 0.0))
                prev = (kp, desc)
            return nm

    # calibrate on one worker, then size the sample for ~seconds_target of wall time
    t0 = time.perf_counter()
    run(frames[:3])
    per_frame = (time.perf_counter() - t0) / 3
    per_worker = int(min(max(4, seconds_target / max(per_frame, 1e-4)), 400))
    subs = [[frames[(w * 7 + i) % len(frames)] for i in range(per_worker)] for w in range(threads)]
    with cf.ThreadPoolExecutor(threads) as ex:
        list(ex.map(run, [s[:2] for s in subs]))  # warm every worker
        t0 = time.perf_counter()
        list(ex.map(run, subs))
        dt = time.perf_counter() - t0
    n = per_worker * threads
    return n / dt, n, dt, kind


def host_threads():
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    if rank != 0:
        return None
    from visionx_slam_b200.synth import synth_sequence
    frames = list(synth_sequence(W, H, 24, seed=2026))
    thr = host_threads()
    total_t, total_n, kind = 0.0, 0, "reference"
    budget = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for step in range(args.warmup + args.steps):
        fps, n, dt, kind = cpu_reference(frames, budget, thr, windowed=False)
        if step >= args.warmup:
            total_t += dt
            total_n += n
    value = total_n / total_t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total_t / max(1, args.steps),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": make_config(args.batch, args.chunks),
        "reference_path": "cv2 (OpenCV) ORB_create(1000,1.2,8).detectAndCompute + BFMatcher(NORM_HAMMING).knnMatch(k=2) + "
                          "ratio 0.8, brute force as orb_matcher.cpp:22-36; each step is a bounded sample of the workload",
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": thr, "kind": kind,
                         "sample": f"{total_n} frames of the synthetic 640x480 sequence in {total_t:.1f} s, "
                                   f"{thr} worker threads x 1 OpenCV thread"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    return json.dumps(line)


# ---- our arm -------------------------------------------------------------------------------------

def check_parity(vx, res, seq, radius, frames_to_check):
    """The fetched resident result against the CPU oracle (test infrastructure used as the checker only):
    keypoints, descriptors and the match rows of frame f against its predecessor (frame 0: the carried last
    frame of the previous pass).  Raises on any difference."""
    import oracle
    B = len(seq)
    need = sorted({f for f in frames_to_check} | {(f - 1) % B for f in frames_to_check})
    ext = {f: oracle.orb_extract(seq[f]) for f in need}
    for f in frames_to_check:
        kp, desc, m = res[f]
        okp, odesc = ext[f]
        pk, pd = ext[(f - 1) % B]
        for name in ("x", "y", "response", "angle", "octave"):
            if not np.array_equal(kp[name], okp[name]):
                raise SystemExit(f"bench.py: PARITY FAILURE frame {f} keypoint field {name}")
        if not np.array_equal(desc, odesc):
            raise SystemExit(f"bench.py: PARITY FAILURE frame {f} descriptors")
        win = radius > 0
        om = oracle.match(pd, odesc, 0.8, np.stack([pk["x"], pk["y"]], 1) if win else None,
                          np.stack([okp["x"], okp["y"]], 1) if win else None, radius)
        if not (np.array_equal(m["query_idx"], om["query_idx"]) and np.array_equal(m["train_idx"], om["train_idx"]) and
                np.array_equal(m["distance"], om["distance"])):
            raise SystemExit(f"bench.py: PARITY FAILURE frame {f} match rows (radius {radius})")
    return {"frames": list(frames_to_check), "radius": radius, "fields": "keypoints, descriptors, match rows",
            "oracle": "oracle/orb_oracle.c (pinned by the cv2 golden vectors)"}


def median_ms(fn, n, warm=10):
    for _ in range(warm):
        fn()
    t = np.empty(n)
    for i in range(n):
        t0 = time.perf_counter()
        fn()
        t[i] = time.perf_counter() - t0
    return float(np.median(t) * 1e3), float(np.min(t) * 1e3)


def cv2_per_call(a, b, n_features, train=None, n=30):
    """cv2 single-thread medians for the same inputs (detectAndCompute; knnMatch + ratio)."""
    try:
        import cv2
    except ImportError:
        return None
    cv2.setNumThreads(1)
    orb = cv2.ORB_create(n_features, 1.2, 8)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING)
    _, da = orb.detectAndCompute(a, None)
    _, db = orb.detectAndCompute(b, None)
    t = db if train is None else train

    def m():
        return sum(1 for r in bf.knnMatch(da, t, 2) if len(r) >= 2 and r[0].distance < 0.8 * r[1].distance)
    ext_ms, _ = median_ms(lambda: orb.detectAndCompute(a, None), n, warm=3)
    match_ms, _ = median_ms(m, max(3, n // 6) if train is not None else n, warm=1)
    return {"detectAndCompute_ms": ext_ms, "knnMatch_ratio_ms": match_ms, "threads": 1, "cv2": cv2.__version__}


def bench_per_call(vx, peak):
    """BASELINE config 1: what Tracking::ProcessFrame pays per frame through the drop-in entry points
    (tracking.cpp:42 Extract, :289 Match): one 640x480 pair, host buffers in and out, synchronous calls."""
    from visionx_slam_b200.synth import synth_scene
    a = synth_scene(W, H, 11)
    b = np.ascontiguousarray(np.roll(a, (3, 5), axis=(0, 1)))
    out = {"config": "config1: one synthetic 640x480 pair (B = roll(A,(3,5))), 1000 kp, host buffers, synchronous calls"}
    with vx.VxOrb(max_batch=1) as h1:
        cap = h1.kp_cap
        kp = np.zeros(cap, vx.KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mt = np.zeros(cap, vx.MATCH_DTYPE)
        na = h1.extract_into(a, kp, desc)
        da = desc[:na].copy()
        nb = h1.extract_into(b, kp, desc)
        db = desc[:nb].copy()
        out["extract_ms"], out["extract_ms_min"] = median_ms(lambda: h1.extract_into(a, kp, desc), 300)
        out["match_ms"], out["match_ms_min"] = median_ms(lambda: h1.match_into(da, db, mt), 300)
        # device-resident descriptors: Extract keeps its result on the GPU under an id, Match takes two ids
        ida = h1.extract_into(a, kp, desc, want_id=True)[1]
        idb = h1.extract_into(b, kp, desc, want_id=True)[1]
        nm_ids = h1.match_ids_into(ida, idb, mt)
        nm_host = h1.match_into(da, db, mt)
        assert nm_ids == nm_host, "match by ids differs from match by host descriptors"
        out["match_ids_ms"], out["match_ids_ms_min"] = median_ms(lambda: h1.match_ids_into(ida, idb, mt), 300)

        state = {"last": idb}

        def frame():  # what Tracking::ProcessFrame does per frame: Extract(curr), Match(last, curr); curr becomes last
            i = h1.extract_into(b, kp, desc, want_id=True)[1]
            h1.match_ids_into(state["last"], i, mt)
            state["last"] = i
        out["extract_plus_match_ms"], _ = median_ms(frame, 300)
        out["keypoints"], out["matches"] = int(na), int(nm_host)
        ab = sum(v for k, v in algorithmic_bytes(na).items() if k in EXTRACT_STAGES)
        out["extract_frac_of_hbm_peak"] = ab / (out["extract_ms"] * 1e-3) / 1e9 / peak
    out["cv2_single_thread"] = cv2_per_call(a, b, NKP)
    return out


def bench_config(vx, peak, popc_peak, w, h, n_kp, batch, radius, steps, label, bytes_model, stereo_box=None):
    """Resident-batch extraction + matching throughput of another BASELINE geometry (temporal frame-to-frame matching,
    or, with stereo_box, chunks of (left, right) pairs: left -> right box-masked + left_(k-1) -> left_k windowed)."""
    from visionx_slam_b200.synth import synth_sequence, synth_stereo_sequence
    seq = synth_stereo_sequence(w, h, batch // 2, seed=404) if stereo_box is not None else synth_sequence(w, h, batch, seed=404)
    out = {"config": label, "batch": batch}
    with vx.VxOrb(n_features=n_kp, max_width=w, max_height=h, max_batch=batch, max_train=max(16384, 2 * n_kp)) as hc:
        run = (lambda: hc.resident_run_stereo(stereo_box, radius)) if stereo_box is not None else (lambda: hc.resident_run(radius))
        hc.resident_upload(seq)
        for _ in range(4):
            run()
        hc.sync()
        ms = 1e30
        for _ in range(2):  # best of two timed regions: these side configurations run only a few milliseconds each
            hc.timer_begin()
            for _ in range(steps):
                run()
            ms = min(ms, hc.timer_end())
        out["frames_per_s_resident"] = batch * steps / (ms * 1e-3)
        hc.stage_times_reset(True)
        for _ in range(steps):
            run()
        st = hc.stage_times()
        hc.stage_times_reset(False)
        res = hc.resident_fetch()
        nk = float(np.mean([len(r[0]) for r in res]))
        ext_ms = sum(st[s][0] for s in EXTRACT_STAGES) / steps
        out["keypoints_per_frame"] = nk
        out["extract_ms_per_frame_serial"] = ext_ms / batch
        out["algorithmic_bytes_per_frame"] = bytes_model
        out["extraction_frac_of_hbm_peak"] = bytes_model * batch / (ext_ms * 1e-3) / 1e9 / peak
        out["stages_ms_per_batch"] = {s: st[s][0] / steps for s in st}
        out["match"] = {"radius": radius, "ms_per_batch": st["match"][0] / steps}
    return out, seq, res


def bench_c3(vx, peak, popc_peak):
    """BASELINE config 3: KITTI-shape 1241x376 stereo pairs, 2000 kp per view, through the stereo chunk path
    (vxorb_resident_run_stereo = the kernels of vxorb_track_submit_stereo): every view extracted, left -> right
    through the box mask 0 <= xl - xr <= 128, |yl - yr| <= 2, and left_(k-1) -> left_k through a 64 px window."""
    box = np.float32([0.0, 128.0, -2.0, 2.0])
    pairs = 48
    out, seq, res = bench_config(vx, peak, popc_peak, 1241, 376, 2000, 2 * pairs, 64.0, 5,
                                 "config3: KITTI-shape 1241x376 stereo pairs, 2000 kp/view, left-right box-masked + temporal "
                                 f"windowed (r=64) matching, {pairs} pairs ({2 * pairs} frames) per resident chunk", 7853539,
                                 stereo_box=box)
    out["stereo_pairs_per_s_resident"] = out["frames_per_s_resident"] / 2
    out["matches_per_pair"] = {"left_right": float(np.mean([len(res[k][2]) for k in range(pairs)])),
                               "temporal": float(np.mean([len(res[pairs + k][2]) for k in range(1, pairs)]))}
    # parity of the benchmarked chunk: pair 3's two views and both of its match legs against the oracle
    import oracle
    ext = {f: oracle.orb_extract(seq[f], 2000) for f in (4, 6, 7)}
    xy = {f: np.stack([ext[f][0]["x"], ext[f][0]["y"]], 1) for f in ext}
    ok = all(np.array_equal(res[f][0], ext[f][0]) and np.array_equal(res[f][1], ext[f][1]) for f in ext)
    ok = ok and np.array_equal(res[3][2], oracle.match_boxed(ext[6][1], ext[7][1], xy[6], xy[7], box))
    ok = ok and np.array_equal(res[pairs + 3][2], oracle.match(ext[4][1], ext[6][1], 0.8, xy[4], xy[6], 64.0))
    if not ok:
        raise SystemExit("bench.py: PARITY FAILURE in the config-3 stereo chunk")
    out["parity_checked"] = True
    # one pair through the host-buffer calls (what a per-frame caller pays)
    left, right = seq[6], seq[7]
    with vx.VxOrb(n_features=2000, max_width=1241, max_height=376, max_batch=2, max_train=4096) as hs:
        (kl, dl), (kr, dr) = hs.extract_batch([left, right])
        qxy, txy = np.stack([kl["x"], kl["y"]], 1), np.stack([kr["x"], kr["y"]], 1)
        ms, ms_min = median_ms(lambda: hs.match_boxed(dl, dr, qxy, txy, box), 100)
        pair_ms, _ = median_ms(lambda: hs.extract_batch([left, right]), 50)
        out["per_call"] = {"left_right_boxed_match_ms": ms, "left_right_boxed_match_ms_min": ms_min,
                           "extract_stereo_pair_ms": pair_ms, "path": "host buffers, synchronous calls"}
    return out


def bench_c4(vx, peak, popc_peak):
    """BASELINE config 4: 1920x1080, 4000 kp, brute-force match of the frame against a 16 384-row local map."""
    out, seq, res = bench_config(vx, peak, popc_peak, 1920, 1080, 4000, 32, 0.0, 6,
                                 "config4: 1920x1080, 4000 kp/frame, batch 32 resident (8 frames leave the launches "
                                 "too short: 18.6 k against 21.7 k frames/s); frame vs 16 384-descriptor map brute force",
                                 32133669)
    rng = np.random.default_rng(0)
    train = rng.integers(0, 256, (16384, 32), dtype=np.uint8)
    q = res[1][1]
    k = min(len(q), 2000)
    train[:k] = q[:k] ^ (rng.integers(0, 256, (k, 32), dtype=np.uint8) & rng.integers(0, 256, (k, 32), dtype=np.uint8)
                         & rng.integers(0, 256, (k, 32), dtype=np.uint8))  # the frame's own rows, ~1/8 of the bits flipped
    with vx.VxOrb(n_features=4000, max_width=1920, max_height=1080, max_batch=1, max_train=16384) as h4:
        cap = h4.kp_cap
        kp = np.zeros(cap, vx.KEYPOINT_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        mt = np.zeros(cap, vx.MATCH_DTYPE)
        n = h4.extract_into(seq[1], kp, desc)
        out["extract_call_ms"], _ = median_ms(lambda: h4.extract_into(seq[1], kp, desc), 60)
        qd = desc[:n].copy()
        nm = h4.match_into(qd, train, mt)
        out["map_match_call_ms"], _ = median_ms(lambda: h4.match_into(qd, train, mt), 60)
        # the map stays on the device between frames (it changes by a few rows per keyframe): resident set + frame id
        sid = h4.set_upload(train)
        fid = h4.extract_into(seq[1], kp, desc, want_id=True)[1]
        assert h4.match_ids_into(fid, sid, mt) == nm
        out["map_match_resident_call_ms"], _ = median_ms(lambda: h4.match_ids_into(fid, sid, mt), 100)
        h4.stage_times_reset(True)
        reps = 20
        for _ in range(reps):
            h4.match_ids_into(fid, sid, mt)
        ms = h4.stage_times()["match"][0] / reps
        h4.stage_times_reset(False)
        pairs = float(n) * 16384.0
        out["map_match"] = {"nq": int(n), "nt": 16384, "matches": int(nm), "device_ms": ms,
                            "popc_issued_per_s": pairs * 4 / (ms * 1e-3), "popc_algorithmic_per_s": pairs * 8 / (ms * 1e-3),
                            "frac_of_popc_peak_issued_4_per_pair": pairs * 4 / (ms * 1e-3) / popc_peak,
                            "frac_of_popc_peak_algorithmic_8_per_pair": pairs * 8 / (ms * 1e-3) / popc_peak,
                            "popc_peak_per_s_measured": popc_peak,
                            "note": "Harley-Seal distance: the kernel issues 4 POPC per (query, train) pair where the "
                                    "SURVEY 8d model counts 8"}
    return out


def fabric_probe(local_rank, world, h2d_bytes, d2h_bytes, host_barrier, iters=48):
    """What this host can feed this rank's GPU while every other rank does the same: plain pinned cudaMemcpyAsync
    loops through torch (no library code), the chunk's host->device bytes on one stream with its device->host
    bytes on another.  Returns this rank's host->device GB/s.  The end-to-end figure cannot exceed it."""
    import torch
    torch.cuda.set_device(local_rank)
    k = 3  # copies in flight per direction, as the streaming path keeps three chunks in flight
    host = [torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory() for _ in range(k)]
    dev = [torch.empty(h2d_bytes, dtype=torch.uint8, device="cuda") for _ in range(k)]
    back_d = [torch.empty(d2h_bytes, dtype=torch.uint8, device="cuda") for _ in range(k)]
    back_h = [torch.empty(d2h_bytes, dtype=torch.uint8).pin_memory() for _ in range(k)]
    s_up = [torch.cuda.Stream() for _ in range(k)]
    s_dn = [torch.cuda.Stream() for _ in range(k)]

    def run(n):
        torch.cuda.synchronize()
        host_barrier(world)
        t0 = time.perf_counter()
        for i in range(n):
            with torch.cuda.stream(s_up[i % k]):
                dev[i % k].copy_(host[i % k], non_blocking=True)
            with torch.cuda.stream(s_dn[i % k]):
                back_h[i % k].copy_(back_d[i % k], non_blocking=True)
        torch.cuda.synchronize()
        return h2d_bytes * n / (time.perf_counter() - t0) / 1e9
    run(8)
    return max(run(iters), run(iters))


def run_ours(args, rank, world, local_rank):
    import visionx_slam_b200 as vx
    from visionx_slam_b200.shard import (aggregate_throughput, host_barrier, init_host_group, max_over_ranks,
                                         sequence_seed, shard_sequences)
    from visionx_slam_b200.synth import synth_sequence

    # torch.distributed (gloo, CPU tensors) carries only the barrier and the max over ranks: the data path has no
    # collective, so no NCCL communicator is created
    init_host_group(world)
    if world > 1:
        # several ranks share the host: keep this rank's threads (and, by first touch, its pinned
        # staging buffers) on the CPUs NVML reports as local to its GPU
        try:
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        except Exception as e:  # no NVML / no permission: run unbound
            sys.stderr.write(f"bench.py: CPU affinity not set ({e!r})\n")

    B, CH = args.batch, args.chunks
    passes = args.steps * CH
    # every rank owns an independent sequence (BASELINE config 5: sequences sharded across GPUs)
    my_sequence = shard_sequences(world, world, rank)[0]  # one sequence per rank
    base = synth_sequence(W, H, min(B, 32), seed=sequence_seed(2026, my_sequence))
    reps = (B + len(base) - 1) // len(base)
    # forward-backward tiling keeps consecutive frames 1-3 px apart across the repeats
    seq = np.concatenate([base if r % 2 == 0 else base[::-1] for r in range(reps)])[:B]

    h = vx.VxOrb(device_id=local_rank, max_width=W, max_height=H, max_batch=B)  # raises without a B200: no CPU fallback
    cap = h.kp_cap

    def barrier():
        h.sync()
        host_barrier(world)
        h.sync()

    def resident_timed(radius, n_warm, n_timed):
        for _ in range(n_warm):
            h.resident_run(radius)
        barrier()
        l0 = h.launch_count()
        h.timer_begin()
        for _ in range(n_timed):
            h.resident_run(radius)
        ms = h.timer_end()
        return ms, h.launch_count() - l0

    # -------- value: chunk resident in HBM, brute-force matcher ------------------------------------
    h.resident_upload(seq)
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ms_resident, launches_timed = resident_timed(0.0, args.warmup * CH, passes)
    barrier()
    # instrumented pass over the same resident chunk: CUDA events around every kernel family on the
    # library's stream.  It runs the stages strictly one after another (the production schedule above lets
    # the blur run beside FAST + select on a side stream and pipelines passes over the slots), so a stage's
    # time is that stage alone; `value` comes from the un-instrumented loop above, the roofline and the stage
    # table from this one.
    n_instr = max(4, min(passes, 24))
    h.stage_times_reset(True)
    h.timer_begin()
    for _ in range(n_instr):
        h.resident_run(0.0)
    ms_serial = h.timer_end()
    stages = h.stage_times()
    h.stage_times_reset(False)
    h.resident_run(0.0)
    h.resident_run(0.0)  # second pass: frame 0 is matched against the carried last frame
    res = h.resident_fetch()
    n_kp = float(np.mean([len(r[0]) for r in res]))
    n_match = float(np.mean([len(r[2]) for r in res]))
    parity = None
    if rank == 0:
        parity = [check_parity(vx, res, seq, 0.0, [0, 1, B // 2, B - 1])]

    # -------- e2e: pinned host frames in, keypoints + descriptors + matches out ------------------
    slots = h.track_slots()
    inflight_max = max(1, min(slots, 3))  # chunks in flight: with 3 the H2D of chunk i+2 hides behind the kernels of chunk i
    pin_in = [vx.PinnedArray((B, H, W), np.uint8) for _ in range(inflight_max)]
    for p in pin_in:
        p.array[:] = seq
    outs = [dict(kp=vx.PinnedArray((B, cap), vx.KEYPOINT_DTYPE), desc=vx.PinnedArray((B, cap, 32), np.uint8),
                 nk=vx.PinnedArray((B,), np.int32), mt=vx.PinnedArray((B, cap), vx.MATCH_DTYPE),
                 nm=vx.PinnedArray((B,), np.int32)) for _ in range(inflight_max)]

    def e2e_chunks(n, radius):
        inflight = []
        for i in range(n):
            k = i % inflight_max
            if len(inflight) == inflight_max:
                h.track_wait(inflight.pop(0))
            o = outs[k]
            inflight.append(h.track_submit(pin_in[k].array, radius, o["kp"].array, o["desc"].array, o["nk"].array,
                                           o["mt"].array, o["nm"].array))
        for s in inflight:
            h.track_wait(s)

    def e2e_timed(radius, n_warm, n_timed):
        h.track_reset()
        e2e_chunks(max(n_warm, inflight_max), radius)
        barrier()
        h.timer_begin()
        t0 = time.perf_counter()
        e2e_chunks(n_timed, radius)
        dev = h.timer_end()
        wall = 1e3 * (time.perf_counter() - t0)
        return dev, wall

    ms_e2e_dev, ms_e2e_wall = e2e_timed(0.0, args.warmup * CH, passes)
    ms_e2e = max(ms_e2e_dev, ms_e2e_wall)
    barrier()
    clocks = sampler.stop()
    h2d_chunk = int(B * W * H)
    head = min(cap, NKP)  # rows per frame the streaming path copies back (tails beyond n_features only on demand)
    d2h_chunk = int(B * (head * (20 + 32 + 12) + 16) + 16)
    e2e_nkp = float(np.mean(outs[0]["nk"].array))
    if rank == 0:  # the streamed result of the last chunk written to outs[k] is the same sequence: check it too
        k = (passes - 1) % inflight_max
        o = outs[k]
        streamed = [(o["kp"].array[i, :o["nk"].array[i]], o["desc"].array[i, :o["nk"].array[i]],
                     o["mt"].array[i, :o["nm"].array[i]]) for i in range(B)]
        parity.append(dict(check_parity(vx, streamed, seq, 0.0, [0, B - 1]), path="vxorb_track_submit (e2e)"))

    # -------- BASELINE config 2's windowed matcher, as extra keys ----------------------------------
    win_passes = max(CH, min(passes, 40))
    ms_res_win, _ = resident_timed(RADIUS, 3, win_passes)
    h.stage_times_reset(True)
    for _ in range(4):
        h.resident_run(RADIUS)
    stages_win = h.stage_times()
    h.stage_times_reset(False)
    h.resident_run(RADIUS)
    h.resident_run(RADIUS)
    res_win = h.resident_fetch()
    if rank == 0:
        parity.append(check_parity(vx, res_win, seq, RADIUS, [0, B - 1]))
    win_dev, win_wall = e2e_timed(RADIUS, 3, win_passes)
    ms_e2e_win = max(win_dev, win_wall)

    popc_peak = h.measure_popc_rate()
    link_gbs = fabric_probe(local_rank, world, h2d_chunk, d2h_chunk, host_barrier)

    # -------- max over ranks ----------------------------------------------------------------------
    ms_resident_max, ms_e2e_max, ms_res_win_max, ms_e2e_win_max, inv_link_max = max_over_ranks(
        [ms_resident, ms_e2e, ms_res_win, ms_e2e_win, 1.0 / link_gbs], world)
    value = aggregate_throughput(B * passes, ms_resident_max, world)
    e2e_value = aggregate_throughput(B * passes, ms_e2e_max, world)
    link_min_gbs = 1.0 / inv_link_max  # the slowest rank's link sets the step (max over ranks)
    fabric_ceiling = world * link_min_gbs * 1e9 / (W * H)

    # -------- roofline of the dominant extraction kernel family, per-stage table ------------------
    peak, peak_src = peaks()
    ab = algorithmic_bytes(int(round(n_kp)))
    table = {}
    for s, (ms, n_launch) in stages.items():
        e = {"ms_per_chunk": ms / n_instr, "launches_per_chunk": n_launch / n_instr,
             "share": ms / ms_serial if ms_serial > 0 else None}  # of the serial (instrumented) pass
        if ab[s] > 0 and ms > 0:
            e["algorithmic_bytes_per_frame"] = ab[s]
            e["achieved_gbs"] = ab[s] * B * n_instr / (ms * 1e-3) / 1e9
            e["frac_of_hbm_peak"] = e["achieved_gbs"] / peak
        if s == "match" and ms > 0:
            pairs = float(B) * n_instr * n_kp * n_kp  # (query, train) pairs evaluated: all pairs, as the reference's matcher
            e["bound"] = "integer pipe: POPC 16 lanes/clk/SM (measured with vxorb_measure_popc_rate)"
            e["popc_peak_per_s_measured"] = popc_peak
            e["popc_algorithmic_per_s"] = pairs * 8 / (ms * 1e-3)   # 8 x popc.b32 per pair
            e["popc_issued_per_s"] = pairs * 4 / (ms * 1e-3)        # Harley-Seal: the kernel issues 4 per pair
            e["frac_of_popc_peak_algorithmic_8_per_pair"] = e["popc_algorithmic_per_s"] / popc_peak
            e["frac_of_popc_peak_issued_4_per_pair"] = e["popc_issued_per_s"] / popc_peak
            ms_w = stages_win["match"][0]
            e["windowed_r32_ms_per_chunk"] = ms_w / 4
        table[s] = e
    hbm_stages = [s for s in ("pyramid", "fast", "blur", "describe") if stages[s][0] > 0]
    dom = max(hbm_stages, key=lambda s: stages[s][0])
    dom_ms, dom_launches = stages[dom]
    ext_ms = sum(stages[s][0] for s in EXTRACT_STAGES)
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": table[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
        "frac": table[dom]["achieved_gbs"] / peak,
        "traffic": TRAFFIC.get(dom) if B == 128 else None,  # the ncu capture was taken at batch 128
        "peak_source": peak_src,
        "bytes_per_launch": ab[dom] * B / max(1.0, dom_launches / n_instr),
        "avg_launch_ms": dom_ms / max(1, dom_launches),
        "extraction_total": {
            "algorithmic_bytes_per_frame": sum(ab[s] for s in hbm_stages),
            "achieved_gbs": sum(ab[s] for s in hbm_stages) * B * n_instr / (ext_ms * 1e-3) / 1e9,
            "ms_per_chunk_serial": ext_ms / n_instr,
        },
    }
    roofline["extraction_total"]["frac"] = roofline["extraction_total"]["achieved_gbs"] / peak

    line = None
    if rank == 0:
        cpu = per_call = c3 = c4 = None
        if world == 1 and not args.no_cpu_baseline:
            thr = host_threads()
            fps, n, dt, kind = cpu_reference(list(base), 12.0, thr, windowed=False)
            cpu = {"value": fps, "unit": "frames/s", "cores": thr, "kind": kind,
                   "sample": f"{n} frames of the same synthetic sequence in {dt:.1f} s; cv2 detectAndCompute + "
                             f"BFMatcher.knnMatch(k=2) + ratio (brute force, orb_matcher.cpp:22-36), {thr} worker "
                             f"threads x 1 OpenCV thread"}
            if kind == "reference":
                fps1, n1, dt1, _ = cpu_reference(list(base), 4.0, 1, windowed=False)
                cpu["single_thread_value"] = fps1
        if world == 1 and not args.no_configs:
            per_call = bench_per_call(vx, peak)
            c3 = bench_c3(vx, peak, popc_peak)
            c4 = bench_c4(vx, peak, popc_peak)
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_resident_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": make_config(B, CH),
            "observed": {"keypoints_per_frame": n_kp, "matches_per_frame": n_match},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d_chunk * CH,
                    "d2h_bytes_per_step": d2h_chunk * CH, "ms_per_step": ms_e2e_max / args.steps,
                    "ms_device_events": ms_e2e_dev / args.steps, "ms_wall": ms_e2e_wall / args.steps,
                    "chunks_in_flight": inflight_max, "keypoints_per_frame": e2e_nkp,
                    # host->device link rate the streaming path sustained (this rank); a PCIe Gen5 x16 link tops
                    # out near 52-55 GB/s with the D2H running beside it (scratch/h2d_test.cu), which caps e2e
                    "h2d_gbs": h2d_chunk * passes / (ms_e2e * 1e-3) / 1e9,
                    "d2h_gbs": d2h_chunk * passes / (ms_e2e * 1e-3) / 1e9,
                    # host fabric probe in the same run: pinned copies of one chunk's bytes in both directions on every
                    # rank at once, no library code; the slowest rank's host->device rate is the yardstick for the e2e rate
                    "fabric_probe": {"h2d_gbs_slowest_rank_with_d2h": link_min_gbs, "frames_per_s_at_probe_rate": fabric_ceiling,
                                     "e2e_vs_probe": e2e_value / fabric_ceiling,
                                     "note": "what plain pinned cudaMemcpyAsync loops (torch, no library code) deliver on every rank "
                                             "at once with the same bytes per chunk in both directions; the streaming path moving "
                                             "~1.0x of it means the host link / memory fabric, not a kernel, bounds e2e"}},
            "gpu_launches": int(launches_timed),
            "parity_checked": True, "parity": parity,
            "windowed": {"note": "BASELINE config 2's matcher (square window r = 32 px, cv knnMatch(mask)) instead of brute force",
                         "value": aggregate_throughput(B * win_passes, ms_res_win_max, world),
                         "e2e": aggregate_throughput(B * win_passes, ms_e2e_win_max, world),
                         "ms_per_chunk": ms_res_win_max / win_passes, "unit": "frames/s"},
            "schedule": {"ms_per_chunk_pipelined": ms_resident / passes, "ms_per_chunk_serial_instrumented": ms_serial / n_instr,
                         "note": "value is timed on the production schedule: passes pipelined over the handle's pipeline slots, the blur on "
                                 "a side stream beside FAST + select; the stage table and the roofline come from a serial "
                                 "pass on one stream with events around every stage.  In the production schedule FAST's "
                                 "persistent grid takes 3 CTAs per SM, leaving room for the other slots' kernels (8 % slower "
                                 "alone, 2 % faster pipelined) and every kernel asks for the same 75 % shared-memory split so that "
                                 "CTAs of different kernels can share an SM (+3 %); alone, as in the serial pass, FAST takes "
                                 "all 5 and every kernel the driver's own split"},
            "roofline": roofline, "stages": table, "cpu_baseline": cpu, "clocks": clocks,
            "per_call": per_call, "c3": c3, "c4": c4,
        }
    h.close()
    host_barrier(world)
    return json.dumps(line) if line is not None else None


def _load_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of every stage's kernels, from the committed
    `ncu --set full` capture of this command at batch 128 (profiles/traffic.json, written by profiles/traffic.py)."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return {}
    with open(p) as f:
        t = json.load(f)
    return {k: v["dram_bytes_per_launch"] for k, v in t.items() if not k.startswith("_")}


TRAFFIC = _load_traffic()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=128, help="frames per chunk (one vxorb_track_submit / resident pass)")
    ap.add_argument("--chunks", type=int, default=8, help="chunks per step")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the per_call / c3 / c4 keys")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: libraries that write to fd 1 are sent to stderr for the duration
    # of the run, the line goes to the saved descriptor
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    try:
        if args.impl == "reference":
            line = run_reference(args, rank, world)
        else:
            line = run_ours(args, rank, world, local_rank)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    if line is not None:
        print(line, flush=True)


if __name__ == "__main__":
    main()
