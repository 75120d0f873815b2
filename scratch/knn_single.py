"""device time of single-pair brute-force 2-NN at config-4 sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import visionx_slam_b200 as vx
rng = np.random.default_rng(0)
with vx.VxOrb(max_train=16384) as h:
    for nq, nt in ((4000, 16384), (4000, 4000), (2500, 16384), (8000, 8000)):
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8); t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
        sid_q, sid_t = h.set_upload(q), h.set_upload(t)
        mt = np.zeros(nq, vx.MATCH_DTYPE)
        for _ in range(5): h.match_ids_into(sid_q, sid_t, mt)
        h.stage_times_reset(True)
        for _ in range(30): h.match_ids_into(sid_q, sid_t, mt)
        ms = h.stage_times()["match"][0] / 30
        h.stage_times_reset(False)
        print(os.environ.get("VXORB_LIB", "default")[-16:], nq, nt, round(ms, 4), flush=True)
        h.set_free(sid_q); h.set_free(sid_t)
