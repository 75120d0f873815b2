#!/usr/bin/env python
"""bench.py — ORB extract+match frames/s on BASELINE.json's config 2 (streaming 640x480 TUM-shape
sequence, 1000 kp/frame, frame-to-frame windowed matching), one rank per B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path (8-level pyramid, FAST+NMS, retainBest, Harris, IC angle,
7x7 blur, rBRIEF, windowed Hamming 2-NN + ratio test) over one chunk of `--batch` consecutive
frames of a synthetic sequence.  Three numbers come out of one run:

  value  frames/s with the chunk already resident in HBM (vxorb_resident_run), CUDA events on the
         library's own streams, max over ranks.  The K passes are pipelined over the handle's three
         slots exactly as the streaming path pipelines chunks (pass i+1 waits for pass i's last-frame
         descriptors before its first pair is matched); nothing crosses PCIe.  The working set of a chunk (pyramid + blurred
         pyramid + lists, ~2.3 MB/frame) is larger than the 126 MB L2 at the default batch.
  e2e    the same metric through the public streaming call (vxorb_track_submit/_wait) with pinned
         HOST frames in and keypoints, descriptors and matches copied back to the host inside the
         timed region, pipelined over the handle's slots.
  roofline / stages: per-stage CUDA-event times of a separate, serial instrumented pass over the same chunk,
         each against its algorithmic bytes and MEASURED_PEAKS.json.

`--impl reference` times the reference's own CPU path (cv2 = the OpenCV the reference's
ORBExtractor/ORBMatcher wrap, core/feature/orb_extractor.cpp:13, orb_matcher.cpp:22-36) on the
host cores; if cv2 is missing it falls back to the C oracle port.  No file under /root/reference
is read at run time.
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
WORKLOAD = "config2: streaming 640x480 TUM-shape synthetic sequence, 1000 kp/frame, 8 levels x1.2, " \
           "frame-to-frame windowed (r=32 px) Hamming 2-NN + ratio 0.8"

# ---- algorithmic bytes per frame (SURVEY.md 8d staged-minimum model), computed from the geometry ----


def level_sizes(w=W, h=H, n_levels=8, sf=1.2):
    sf = float(np.float32(sf))
    out = []
    for l in range(n_levels):
        inv = np.float32(1.0) / np.float32(sf ** l)
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
    return out


def algorithmic_bytes(n_kp=NKP):
    px = [a * b for a, b in level_sizes()]
    s = sum(px)
    return {
        "pyramid": (s - px[-1]) + (s - px[0]),   # read every source level + write levels 1..7
        "fast": s,                                # one read of every level pixel
        "select": 0,                              # list work, no image traffic in the model
        "blur": 2 * s,                            # read + write every level
        "describe": n_kp * 512 + n_kp * (32 + 24),  # 512 gathers + descriptor + keypoint record
        "match": 0,                               # integer-pipe bound: popc ops, not bytes
    }


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
            self._stop.wait(0.05)

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
    how the CPU path uses many cores: cv::ORB itself is serial).  Returns (frames/s, frames, s)."""
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
        "config": {"workload": WORKLOAD, "reference_path": "cv2 (OpenCV) ORB_create(1000,1.2,8).detectAndCompute + "
                   "BFMatcher(NORM_HAMMING).knnMatch(k=2) + ratio 0.8, brute force as orb_matcher.cpp:22-36"},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": thr, "kind": kind,
                         "sample": f"{total_n} frames of the synthetic 640x480 sequence in {total_t:.1f} s, "
                                   f"{thr} worker threads x 1 OpenCV thread"},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    return json.dumps(line)


# ---- our arm -------------------------------------------------------------------------------------

def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    import visionx_slam_b200 as vx
    from visionx_slam_b200.synth import synth_sequence

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the ORB path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        # several ranks share the host: keep this rank's threads (and, by first touch, its pinned
        # staging buffers) on the CPUs NVML reports as local to its GPU
        try:
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local_rank))
        except Exception as e:  # no NVML / no permission: run unbound
            sys.stderr.write(f"bench.py: CPU affinity not set ({e!r})\n")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B = args.batch
    # every rank owns an independent sequence (BASELINE config 5: sequences sharded across GPUs)
    from visionx_slam_b200.shard import aggregate_throughput, max_over_ranks, sequence_seed, shard_sequences
    my_sequence = shard_sequences(world, world, rank)[0]  # one sequence per rank
    base = synth_sequence(W, H, min(B, 32), seed=sequence_seed(2026, my_sequence))
    reps = (B + len(base) - 1) // len(base)
    # forward-backward tiling keeps consecutive frames 1-3 px apart across the repeats
    seq = np.concatenate([base if r % 2 == 0 else base[::-1] for r in range(reps)])[:B]

    h = vx.VxOrb(device_id=local_rank, max_width=W, max_height=H, max_batch=B)
    cap = h.kp_cap
    lib_launches0 = h.launch_count()

    # -------- value: chunk resident in HBM --------------------------------------------------------
    h.resident_upload(seq)
    for _ in range(args.warmup):
        h.resident_run(RADIUS)
    h.sync()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    l0 = h.launch_count()
    h.timer_begin()
    for _ in range(args.steps):
        h.resident_run(RADIUS)
    ms_resident = h.timer_end()
    launches_timed = h.launch_count() - l0
    barrier()
    # instrumented pass over the same resident chunk: CUDA events around every kernel family on the
    # library's stream.  It runs the stages strictly one after another (the production schedule above lets
    # the blur run beside FAST + select on a side stream), so a stage's time is that stage alone; `value`
    # comes from the un-instrumented loop above, the roofline and the stage table from this one.
    h.stage_times_reset(True)
    h.timer_begin()
    for _ in range(args.steps):
        h.resident_run(RADIUS)
    ms_serial = h.timer_end()
    stages = h.stage_times()
    h.stage_times_reset(False)
    # the reference's own matcher is brute force (orb_matcher.cpp:22-25): the same instrumented pass with
    # radius 0 gives the match stage of the all-pairs kernel, reported against the measured popcount peak
    bf_steps = max(2, min(5, args.steps))
    h.resident_run(0.0)
    h.sync()
    h.stage_times_reset(True)
    for _ in range(bf_steps):
        h.resident_run(0.0)
    h.sync()
    stages_bf = h.stage_times()
    h.stage_times_reset(False)
    h.resident_run(RADIUS)  # leave the windowed result resident for the fetch below
    barrier()
    res = h.resident_fetch()
    n_kp = float(np.mean([len(r[0]) for r in res]))
    n_match = float(np.mean([len(r[2]) for r in res[1:]])) if B > 1 else 0.0

    # -------- e2e: pinned host frames in, keypoints + descriptors + matches out ------------------
    slots = h.track_slots()
    chunks = max(1, min(slots, 3))  # chunks in flight: with 3 the H2D of chunk i+2 hides behind the kernels of chunk i
    pin_in = [vx.PinnedArray((B, H, W), np.uint8) for _ in range(chunks)]
    for p in pin_in:
        p.array[:] = seq
    outs = [dict(kp=vx.PinnedArray((B, cap), vx.KEYPOINT_DTYPE), desc=vx.PinnedArray((B, cap, 32), np.uint8),
                 nk=vx.PinnedArray((B,), np.int32), mt=vx.PinnedArray((B, cap), vx.MATCH_DTYPE),
                 nm=vx.PinnedArray((B,), np.int32)) for _ in range(chunks)]

    def e2e_steps(n):
        inflight = []
        for i in range(n):
            k = i % chunks
            if len(inflight) == chunks:
                h.track_wait(inflight.pop(0))
            o = outs[k]
            inflight.append(h.track_submit(pin_in[k].array, RADIUS, o["kp"].array, o["desc"].array, o["nk"].array,
                                           o["mt"].array, o["nm"].array))
        for s in inflight:
            h.track_wait(s)

    h.track_reset()
    e2e_steps(max(args.warmup, chunks))
    barrier()
    h.timer_begin()
    t0 = time.perf_counter()
    e2e_steps(args.steps)
    ms_e2e_dev = h.timer_end()
    ms_e2e_wall = 1e3 * (time.perf_counter() - t0)
    ms_e2e = max(ms_e2e_dev, ms_e2e_wall)
    barrier()
    clocks = sampler.stop()
    h2d = int(B * W * H)
    d2h = int(B * (cap * (20 + 32 + 12) + 8) + 16)
    e2e_nkp = float(np.mean(outs[0]["nk"].array))

    popc_peak = h.measure_popc_rate()

    # -------- max over ranks ----------------------------------------------------------------------
    ms_resident_max, ms_e2e_max = max_over_ranks([ms_resident, ms_e2e], world, device="cuda")
    value = aggregate_throughput(B * args.steps, ms_resident_max, world)
    e2e_value = aggregate_throughput(B * args.steps, ms_e2e_max, world)

    # -------- roofline of the dominant extraction kernel family, per-stage table ------------------
    peak, peak_src = peaks()
    ab = algorithmic_bytes(int(round(n_kp)))
    table = {}
    for s, (ms, n_launch) in stages.items():
        per_step_ms = ms / args.steps
        e = {"ms_per_step": per_step_ms, "launches_per_step": n_launch / args.steps,
             "share": ms / ms_serial if ms_serial > 0 else None}  # of the serial (instrumented) step
        if ab[s] > 0 and ms > 0:
            e["algorithmic_bytes_per_frame"] = ab[s]
            e["achieved_gbs"] = ab[s] * B * args.steps / (ms * 1e-3) / 1e9
            e["frac_of_hbm_peak"] = e["achieved_gbs"] / peak
        if s == "match" and ms > 0:
            pairs = (B - 1 + 1) * args.steps  # frame 0 matches the carried last frame from step 2 on
            popc = n_kp * n_kp * 8 * pairs
            e["allpairs_equiv_popc_per_s"] = popc / (ms * 1e-3)
            e["popc_peak_per_s_measured"] = popc_peak
            e["frac_of_popc_peak_allpairs_equiv"] = e["allpairs_equiv_popc_per_s"] / popc_peak
            ms_bf = stages_bf["match"][0]
            if ms_bf > 0:
                popc_bf = n_kp * n_kp * 8 * B * bf_steps  # 8 x 32-bit popcounts per (query, train) pair, all pairs
                e["bruteforce"] = {"ms_per_step": ms_bf / bf_steps, "popc_per_s": popc_bf / (ms_bf * 1e-3),
                                   "frac_of_popc_peak": popc_bf / (ms_bf * 1e-3) / popc_peak,
                                   "note": "all-pairs 2-NN of the same 128 pairs (radius 0), the matcher the reference "
                                           "itself runs; bound: integer pipe, POPC 16 lanes/clk/SM"}
        table[s] = e
    hbm_stages = [s for s in ("pyramid", "fast", "blur", "describe") if stages[s][0] > 0]
    dom = max(hbm_stages, key=lambda s: stages[s][0])
    dom_ms, dom_launches = stages[dom]
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": table[dom]["achieved_gbs"], "peak": peak, "unit": "GB/s",
        "frac": table[dom]["achieved_gbs"] / peak,
        "traffic": TRAFFIC.get(dom) if B == 128 else None,  # the ncu capture was taken at batch 128
        "peak_source": peak_src,
        "bytes_per_launch": ab[dom] * B / max(1.0, dom_launches / args.steps),
        "avg_launch_ms": dom_ms / max(1, dom_launches),
        "extraction_total": {
            "algorithmic_bytes_per_frame": sum(ab[s] for s in hbm_stages),
            "achieved_gbs": sum(ab[s] for s in hbm_stages) * B * args.steps /
            (sum(stages[s][0] for s in ("pyramid", "fast", "select", "blur", "describe")) * 1e-3) / 1e9,
        },
    }
    roofline["extraction_total"]["frac"] = roofline["extraction_total"]["achieved_gbs"] / peak

    line = None
    if rank == 0:
        cpu = None
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
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_resident_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "batch_frames_per_step_per_gpu": B, "keypoints_per_frame": n_kp,
                       "matches_per_frame": n_match, "sharding": "one independent sequence per GPU, no collective",
                       "l2_policy": f"inputs larger than L2: {B} frames x ~2.3 MB working set per step"},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e_max / args.steps, "ms_device_events": ms_e2e_dev / args.steps,
                    "ms_wall": ms_e2e_wall / args.steps, "chunks_in_flight": chunks, "keypoints_per_frame": e2e_nkp,
                    # host->device link rate the streaming path sustained (this rank); a PCIe Gen5 x16 link tops
                    # out near 52-55 GB/s with the D2H running beside it (scratch/h2d_test.cu), which caps e2e
                    "h2d_gbs": h2d / (ms_e2e / args.steps * 1e-3) / 1e9},
            "gpu_launches": int(launches_timed),
            "schedule": {"ms_per_step_overlapped": ms_resident / args.steps, "ms_per_step_serial_instrumented": ms_serial / args.steps,
                         "note": "value is timed on the production schedule: passes pipelined over the handle's 3 slots, the blur on "
                                 "a side stream beside FAST + select; the stage table and the roofline come from a serial "
                                 "pass on one stream with events around every stage"},
            "roofline": roofline, "stages": table, "cpu_baseline": cpu, "clocks": clocks,
        }
    h.close()
    if world > 1:
        dist.destroy_process_group()
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
    ap.add_argument("--batch", type=int, default=128, help="frames per step per GPU")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: libraries that write to fd 1 (NCCL prints its version banner
    # there) are sent to stderr for the duration of the run, the line goes to the saved descriptor
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
