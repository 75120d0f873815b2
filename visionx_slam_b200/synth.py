"""Deterministic synthetic frames (numpy only — no cv2, so it runs on the GPU box).

Shapes follow BASELINE.json's configs (TUM fr1_desk 640x480, KITTI 1241x376, 1920x1080).  The
scene is SURVEY.md Appendix C.1's: 300 random filled rectangles, sigma=1 Gaussian, N(0,3) noise —
enough texture that cv::ORB fills every per-level quota.  All float arithmetic is elementwise
float32 in a fixed order, so the bytes are reproducible across machines.
"""
import hashlib

import numpy as np


def _gauss_sep(img, sigma=1.0, radius=3):
    xs = np.arange(-radius, radius + 1, dtype=np.float64)
    k = np.exp(-(xs * xs) / (2.0 * sigma * sigma))
    k = (k / k.sum()).astype(np.float32)
    h, w = img.shape
    p = np.pad(img, ((0, 0), (radius, radius)), mode="reflect")
    acc = np.zeros_like(img)
    for i in range(2 * radius + 1):
        acc = acc + p[:, i:i + w] * k[i]
    p = np.pad(acc, ((radius, radius), (0, 0)), mode="reflect")
    out = np.zeros_like(img)
    for i in range(2 * radius + 1):
        out = out + p[i:i + h, :] * k[i]
    return out


def synth_scene(w, h, seed, n_rect=None, noise=3.0):
    """One w x h uint8 frame."""
    rng = np.random.default_rng(seed)
    if n_rect is None:
        n_rect = max(300, int(300 * (w * h) / (640 * 480)))
    img = np.zeros((h, w), np.float32)
    for _ in range(n_rect):
        x = int(rng.integers(0, w))
        y = int(rng.integers(0, h))
        a, b = (int(v) for v in rng.integers(5, 60, 2))
        c = float(rng.integers(0, 256))
        img[y:y + b + 1, x:x + a + 1] = c
    img = _gauss_sep(img)
    if noise > 0:
        img = img + rng.normal(0.0, noise, (h, w)).astype(np.float32)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_sequence(w, h, n_frames, seed, max_step=3, noise=2.0):
    """n_frames crops of one larger scene translated 1..max_step px per frame (bouncing), each
    with its own sensor noise: the streaming TUM-shape sequence of BASELINE config 2."""
    rng = np.random.default_rng(seed + 7919)
    margin = 96
    scene = synth_scene(w + 2 * margin, h + 2 * margin, seed, noise=0.0).astype(np.float32)
    frames = np.empty((n_frames, h, w), np.uint8)
    x, y, dx, dy = margin, margin, 1, 1
    for i in range(n_frames):
        crop = scene[y:y + h, x:x + w]
        if noise > 0:
            crop = crop + rng.normal(0.0, noise, (h, w)).astype(np.float32)
        frames[i] = np.clip(np.rint(crop), 0, 255).astype(np.uint8)
        sx, sy = (int(v) for v in rng.integers(1, max_step + 1, 2))
        if not (0 <= x + dx * sx <= 2 * margin):
            dx = -dx
        if not (0 <= y + dy * sy <= 2 * margin):
            dy = -dy
        x += dx * sx
        y += dy * sy
    return frames


def synth_stereo_pair(w=1241, h=376, seed=31):
    """Rectified stereo pair of BASELINE config 3's shape: the right view is the left scene shifted left
    by a row-dependent disparity (8..63 px) plus +-2 grey levels of noise."""
    left = synth_scene(w, h, seed)
    rng = np.random.default_rng(seed)
    right = np.empty_like(left)
    for y in range(h):
        right[y] = np.roll(left[y], -(8 + (y * 56) // h))
    right = np.clip(right.astype(np.int16) + rng.integers(-2, 3, right.shape), 0, 255).astype(np.uint8)
    return left, right


def synth_stereo_sequence(w, h, pairs, seed, max_disparity=48):
    """(left_0, right_0, left_1, right_1, ...): the left views are a synth_sequence, every right view is its left
    view shifted by a row-dependent disparity (8..max_disparity px) plus +-2 grey levels of noise."""
    lefts = synth_sequence(w, h, pairs, seed=seed)
    rng = np.random.default_rng(seed)
    out = np.empty((2 * pairs, h, w), np.uint8)
    for k in range(pairs):
        right = np.empty_like(lefts[k])
        for y in range(h):
            right[y] = np.roll(lefts[k][y], -(8 + (y * (max_disparity - 8)) // h))
        out[2 * k] = lefts[k]
        out[2 * k + 1] = np.clip(right.astype(np.int16) + rng.integers(-2, 3, right.shape), 0, 255).astype(np.uint8)
    return out


def sha256(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
