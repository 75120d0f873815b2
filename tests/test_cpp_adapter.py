"""The C++ host side (visionx_slam_b200/cpp): the reference-facing classes CudaORBExtractor /
CudaORBMatcher, compiled against tests/stubs (no OpenCV/Eigen/glog C++ headers in this image).

CPU: the adapters compile with the reference's warning flags (-Wall -Wextra -Wpedantic,
CMakeLists.txt:11-15) and the driver links against libvxorb.so.
GPU: tests/cpp/adapter_main.cpp drives Extract on every frame and Match(last, curr) the way
tracking.cpp:42,208 does; results are compared with the oracle bit for bit."""
import os
import struct
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "visionx_slam_b200", "cpp")
EXE = os.path.join(CPP, "build", "adapter_main")


def _build():
    subprocess.check_call(["make", "-C", CPP, "-s"])


def test_adapters_compile_and_link():
    _build()
    assert os.path.exists(EXE)
    out = subprocess.run(["ldd", EXE], capture_output=True, text=True).stdout
    assert "libvxorb.so" in out and "not found" not in out.split("libvxorb.so")[1].split("\n")[0]


def test_adapters_derive_from_reference_interfaces():
    h = open(os.path.join(CPP, "cuda_orb_extractor.h")).read()
    assert "class CudaORBExtractor : public FeatureExtractor" in h and "void Extract(Frame& frame) override;" in h
    m = open(os.path.join(CPP, "cuda_orb_matcher.h")).read()
    assert "class CudaORBMatcher : public FeatureMatcher" in m
    assert "int Match(const Frame::Ptr& last, const Frame::Ptr& curr, std::vector<cv::DMatch>& matches) override;" in m


def _run_adapter(tmp_path, frames, initial_max=None):
    n, hh, w = frames.shape[:3]
    ch = frames.shape[3] if frames.ndim == 4 else 1
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        f.write(struct.pack("<4i", n, w, hh, ch))
        f.write(np.ascontiguousarray(frames).tobytes())
    cmd = [EXE, str(fin), str(fout)] + ([str(initial_max)] if initial_max else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    buf = open(fout, "rb").read()
    pos = 0
    feats = []
    for _ in range(n):
        (k,) = struct.unpack_from("<i", buf, pos)
        pos += 4
        rec = np.frombuffer(buf, np.dtype([("x", "<f8"), ("y", "<f8"), ("r", "<f4")]), k, pos)
        pos += 20 * k
        desc = np.frombuffer(buf, np.uint8, 32 * k, pos).reshape(k, 32)
        pos += 32 * k
        feats.append((rec, desc))
    matches = []
    for _ in range(2):
        per = []
        for _ in range(n - 1):
            (m,) = struct.unpack_from("<i", buf, pos)
            pos += 4
            per.append(np.frombuffer(buf, np.dtype([("q", "<i4"), ("t", "<i4"), ("d", "<f4")]), m, pos))
            pos += 12 * m
        matches.append(per)
    assert pos == len(buf)
    return feats, matches[0], matches[1], r.stderr


@pytest.mark.gpu
def test_cpp_adapter_matches_oracle(tmp_path):
    import oracle
    from conftest import matches_as_rows
    from visionx_slam_b200.synth import synth_sequence
    _build()
    frames = synth_sequence(640, 480, 3, seed=21)
    frames = np.concatenate([frames, np.zeros((1, 480, 640), np.uint8)])  # blank frame: 0 features, 0 matches
    feats, m_plain, m_filt, log = _run_adapter(tmp_path, frames)
    ref = [oracle.orb_extract(f) for f in frames]
    for (rec, desc), (okp, odesc) in zip(feats, ref):
        assert len(rec) == len(okp)
        assert np.array_equal(rec["x"], okp["x"].astype(np.float64)) and np.array_equal(rec["y"], okp["y"].astype(np.float64))
        assert np.array_equal(rec["r"], okp["response"])
        assert np.array_equal(desc, odesc)
    for i in range(1, len(frames)):
        om = oracle.match(ref[i - 1][1], ref[i][1], 0.8)
        got = m_plain[i - 1]
        assert np.array_equal(np.stack([got["q"], got["t"], got["d"].astype(np.int32)], 1).reshape(-1, 3), matches_as_rows(om))
        of = oracle.postfilter(om)
        gf = m_filt[i - 1]
        assert np.array_equal(np.stack([gf["q"], gf["t"], gf["d"].astype(np.int32)], 1).reshape(-1, 3), matches_as_rows(of))
    assert len(feats[3][0]) == 0 and len(m_plain[2]) == 0
    # CheckImageQuality / CheckFeatureDistribution (tracking.cpp:93-139) answered by the adapter
    infos = [l.split()[1:] for l in log.splitlines() if l.startswith("FRAMEINFO")]
    assert len(infos) == len(frames)
    for (i, mean, sd, occ, q_ok, d_ok), f, (okp, _) in zip(infos, frames, ref):
        om, osd = oracle.image_stats(f)
        assert float(mean) == om and float(sd) == osd
        oocc, _ = oracle.grid_coverage(np.stack([okp["x"], okp["y"]], 1) if len(okp) else None, 640, 480)
        assert int(occ) == oocc
        assert int(q_ok) == int(30 <= om <= 225 and osd >= 20) and int(d_ok) == int(oocc >= 12.5)
    assert infos[3][4] == "0" and infos[3][5] == "0"  # blank frame fails both checks
    # the depth look-up of CreateLandmarksFromDepth served from the positions Extract left on the device
    deps = [l.split()[1:] for l in log.splitlines() if l.startswith("DEPTH")]
    assert len(deps) == len(frames)
    yy, xx = np.mgrid[0:480, 0:640]
    for (i, ok, valid, total), (okp, _) in zip(deps, ref):
        depth = ((xx * 37 + yy * 101 + int(i) * 977) % 60000).astype(np.uint16)
        if len(okp) == 0:
            assert ok == "0"  # the blank frame has no resident descriptors: the caller keeps its own loop
            continue
        want = oracle.depth_lookup(np.stack([okp["x"], okp["y"]], 1), depth)
        acc = 0.0
        for v in want.tolist():  # plain sequential double adds like the C++ loop (Python's sum() compensates)
            acc += v
        assert ok == "1" and int(valid) == int((want > 0).sum()) and float(total) == acc
    # the shared-handle matcher served the non-empty pairs from the descriptors Extract left on the device
    path = [l for l in log.splitlines() if l.startswith("MATCHPATH")][0]
    assert "resident=4 upload=0" in path, path  # 2 pairs x (Match + MatchFiltered); the blank pair returns early
    assert "Too few matches" not in log  # an empty side returns before the warning (orb_matcher.cpp:18-20)


@pytest.mark.gpu
def test_cpp_adapter_bgr(tmp_path):
    import oracle
    from visionx_slam_b200.synth import synth_scene
    _build()
    rng = np.random.default_rng(5)
    g = synth_scene(320, 240, 9)
    bgr = np.stack([np.clip(g.astype(int) + rng.integers(-20, 20, g.shape), 0, 255).astype(np.uint8) for _ in range(3)], 2)
    other = np.ascontiguousarray(bgr[::-1, ::-1])  # unrelated second frame: few ratio-test survivors
    feats, m_plain, _, log = _run_adapter(tmp_path, np.stack([bgr, other]))
    okp, odesc = oracle.orb_extract(oracle.bgr2gray(bgr))
    assert len(feats[0][0]) == len(okp) and np.array_equal(feats[0][1], odesc)
    okp2, odesc2 = oracle.orb_extract(oracle.bgr2gray(other))
    om = oracle.match(odesc, odesc2, 0.8)
    assert len(m_plain[0]) == len(om)
    assert ("Too few matches" in log) == (len(om) < 50)  # orb_matcher.cpp:38-40 warning kept


@pytest.mark.gpu
def test_cpp_adapter_grows_to_the_frame_size(tmp_path):
    """cv::ORB takes frames of any size: an extractor built for 96 x 96 rebuilds its handle on the first 640 x 480
    frame and then gives exactly what one built large enough gives, resident matching included."""
    from visionx_slam_b200.synth import synth_sequence
    _build()
    frames = synth_sequence(640, 480, 3, seed=33)
    feats, m_plain, m_filt, log = _run_adapter(tmp_path, frames)
    feats2, m_plain2, m_filt2, log2 = _run_adapter(tmp_path, frames, initial_max=96)
    assert "REBUILDS 0" in log and "REBUILDS 1" in log2
    for (a, da), (b, db) in zip(feats, feats2):
        assert len(a) > 900 and np.array_equal(a, b) and np.array_equal(da, db)
    for a, b in zip(m_plain + m_filt, m_plain2 + m_filt2):
        assert len(a) > 0 and np.array_equal(a, b)
    path = [l for l in log2.splitlines() if l.startswith("MATCHPATH")][0]
    assert "resident=4 upload=0" in path, path
