"""Seeded synthetic inputs for the ORB front-end (no datasets are available offline).

Frames follow SURVEY.md §8(d): band-limited noise blended with an 8x nearest-upsampled block
image (thousands of FAST@20 corners per level), plus a low-contrast gradient region (cells that
take the minThFAST retry, reference src/ORBextractor.cc:1118-1125) and a constant region (cells
that stay empty).  Pure numpy so that tests, bench.py and the GPU box generate identical bytes.
"""
from __future__ import annotations

import hashlib

import numpy as np

# EuRoC / TUM-VI / KITTI stereo constants from the reference's YAMLs
CONFIGS = {
    # name: (rows, cols, nfeatures, mbf, fx)
    "C1": dict(rows=480, cols=752, nfeatures=1000, mbf=47.90639, fx=435.2047),   # Examples/Monocular/EuRoC.yaml
    "C2": dict(rows=480, cols=752, nfeatures=1200, mbf=47.90639, fx=435.2047),   # Examples/Stereo/EuRoC.yaml:9,28
    "C3": dict(rows=512, cols=512, nfeatures=1000, mbf=19.3079, fx=190.978),     # Examples/Stereo/TUM_512.yaml:9,69
    "C4": dict(rows=376, cols=1241, nfeatures=2000, mbf=386.1448, fx=718.856),   # Examples/Stereo/KITTI00-02.yaml:9,28
    "C5": dict(rows=2160, cols=3840, nfeatures=8000, mbf=386.1448, fx=718.856),
}
ORB_DEFAULTS = dict(scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7)


def _binomial5(a: np.ndarray) -> np.ndarray:
    """Separable [1,4,6,4,1]/16 blur with edge replication (float32)."""
    k = np.array([1, 4, 6, 4, 1], np.float32) / 16.0
    p = np.pad(a, ((0, 0), (2, 2)), mode="edge")
    a = sum(k[i] * p[:, i:i + a.shape[1]] for i in range(5))
    p = np.pad(a, ((2, 2), (0, 0)), mode="edge")
    return sum(k[i] * p[i:i + a.shape[0], :] for i in range(5))


def synth_frame(rows: int, cols: int, seed: int, bands: bool = True) -> np.ndarray:
    rng = np.random.default_rng(seed)
    noise = rng.integers(0, 256, (rows, cols), dtype=np.uint8).astype(np.float32)
    tex = _binomial5(noise)
    tex = (tex - 128.0) * 2.2 + 128.0
    br, bc = (rows + 7) // 8, (cols + 7) // 8
    blocks = rng.integers(0, 256, (br, bc), dtype=np.uint8)
    blocks = np.repeat(np.repeat(blocks, 8, axis=0), 8, axis=1)[:rows, :cols].astype(np.float32)
    img = 0.6 * tex + 0.4 * blocks
    if bands:
        # low-contrast gradient region (lower-left quarter): FAST@20 finds nothing, FAST@7 some
        y0, x1 = rows // 2, cols // 2
        gx = np.linspace(64.0, 192.0, x1, dtype=np.float32)[None, :]
        lowc = gx + (tex[y0:, :x1] - 128.0) * 0.16
        img[y0:, :x1] = lowc
        # constant region (upper-right): no corners at any threshold
        img[: rows // 4, (cols * 3) // 5:] = 128.0
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def synth_stereo_pair(rows: int, cols: int, seed: int, max_disp: int = 40):
    """Right image = left shifted by a per-row-block disparity in [1, max_disp] plus +-2 grey levels."""
    left = synth_frame(rows, cols, seed)
    rng = np.random.default_rng(seed + 1_000_003)
    right = np.empty_like(left)
    band = 32
    for y0 in range(0, rows, band):
        d = int(rng.integers(1, max_disp + 1))
        blk = left[y0:y0 + band]
        sh = np.empty_like(blk)
        sh[:, : cols - d] = blk[:, d:]
        sh[:, cols - d:] = blk[:, cols - 1:cols]
        right[y0:y0 + band] = sh
    jitter = rng.integers(-2, 3, (rows, cols), dtype=np.int16)
    right = np.clip(right.astype(np.int16) + jitter, 0, 255).astype(np.uint8)
    return left, right


def synth_rectify_maps(rows: int, cols: int, src_rows: int | None = None, src_cols: int | None = None, seed: int = 0):
    """CV_32FC1 maps shaped like cv::initUndistortRectifyMap's (Examples/Stereo/stereo_euroc.cc:117-118): a
    small rotation + radial (k1, k2) distortion about a principal point, in float32 arithmetic.  Near the
    corners the maps leave the source image (remap's BORDER_CONSTANT taps)."""
    src_rows = src_rows or rows; src_cols = src_cols or cols
    rng = np.random.default_rng(seed + 77)
    fx = np.float32(0.61 * cols); fy = np.float32(0.61 * cols)
    cx = np.float32(cols / 2 + rng.uniform(-6, 6)); cy = np.float32(rows / 2 + rng.uniform(-6, 6))
    k1 = np.float32(-0.28 + rng.uniform(-0.02, 0.02)); k2 = np.float32(0.07)
    th = np.float32(rng.uniform(-0.01, 0.01))
    y, x = np.mgrid[0:rows, 0:cols].astype(np.float32)
    xn = (x - cx) / fx; yn = (y - cy) / fy
    xr = np.cos(th) * xn - np.sin(th) * yn; yr = np.sin(th) * xn + np.cos(th) * yn
    r2 = xr * xr + yr * yr
    k = np.float32(1) + k1 * r2 + k2 * r2 * r2
    sx = np.float32(src_cols / cols); sy = np.float32(src_rows / rows)
    mapx = ((xr * k * fx + cx) * sx + np.float32(rng.uniform(-1.5, 1.5))).astype(np.float32)
    mapy = ((yr * k * fy + cy) * sy + np.float32(rng.uniform(-1.5, 1.5))).astype(np.float32)
    return np.ascontiguousarray(mapx), np.ascontiguousarray(mapy)


def synth_color_frame(rows: int, cols: int, seed: int, channels: int = 3) -> np.ndarray:
    """A colour camera frame: the gray synthetic frame with per-channel gain / offset / noise (alpha = 255)."""
    g = synth_frame(rows, cols, seed).astype(np.float32)
    rng = np.random.default_rng(seed + 4242)
    out = np.empty((rows, cols, channels), np.uint8)
    for c in range(3):
        ch = g * np.float32(rng.uniform(0.8, 1.15)) + np.float32(rng.uniform(-12, 12)) + rng.integers(-3, 4, (rows, cols))
        out[..., c] = np.clip(np.rint(ch), 0, 255).astype(np.uint8)
    if channels == 4:
        out[..., 3] = 255
    return out


def synth_map_points(desc: np.ndarray, kps_xy: np.ndarray, octaves: np.ndarray, n: int, seed: int,
                     rows: int, cols: int, nlevels: int = 8):
    """Synthetic map points for SearchByProjection: descriptors of real keypoints with 0..40 random
    bit flips, projected near (2/3) or far from (1/3) their keypoint."""
    rng = np.random.default_rng(seed)
    src = rng.integers(0, desc.shape[0], n)
    d = desc[src].copy()
    bits = np.unpackbits(d, axis=1)
    nflip = rng.integers(0, 41, n)
    for i in range(n):
        pos = rng.choice(256, nflip[i], replace=False)
        bits[i, pos] ^= 1
    d = np.packbits(bits, axis=1)
    near = rng.random(n) < 0.67
    px = np.where(near, kps_xy[src, 0] + rng.normal(0, 2.0, n), rng.uniform(0, cols, n)).astype(np.float32)
    py = np.where(near, kps_xy[src, 1] + rng.normal(0, 2.0, n), rng.uniform(0, rows, n)).astype(np.float32)
    lvl = np.where(near, np.clip(octaves[src] + rng.integers(0, 2, n), 0, nlevels - 1),
                   rng.integers(0, nlevels, n)).astype(np.int32)
    viewcos = np.where(rng.random(n) < 0.5, 0.9, 0.999).astype(np.float32)
    depth = rng.uniform(0.5, 60.0, n).astype(np.float32)
    inview = (rng.random(n) < 0.95).astype(np.uint8)
    return dict(desc=np.ascontiguousarray(d), projx=px, projy=py, projxr=(px - rng.uniform(0, 30, n)).astype(np.float32),
                level=lvl, viewcos=viewcos, depth=depth, inview=inview)


def synth_vocabulary(k: int = 10, L: int = 4, seed: int = 0, early_leaves: float = 0.02, stop_words: float = 0.03):
    """A vocabulary tree shaped like ORBvoc (k-ary, depth L) in the row order of its text file
    (TemplatedVocabulary.h:1338-1415): child descriptors = parent's with random bit flips (fewer deeper down), idf-like
    weights, a few stopped words (weight 0) and a few leaves above depth L.  Returns dict(k, L, parent, is_leaf, desc, weight)."""
    rng = np.random.default_rng(seed + 31337)
    parent, leaf, desc, weight = [], [], [], []
    frontier = [(0, np.zeros(256, np.uint8), 0)]          # (node id, descriptor bits, level)
    frontier[0] = (0, rng.integers(0, 2, 256).astype(np.uint8), 0)
    nid = 0
    while frontier:
        nxt = []
        for pid, pbits, lvl in frontier:
            for _ in range(k):
                nid += 1
                bits = pbits ^ (rng.random(256) < (0.5 if lvl == 0 else 0.25 / lvl)).astype(np.uint8)
                is_leaf = lvl + 1 == L or (lvl + 1 >= 2 and rng.random() < early_leaves)
                parent.append(pid); leaf.append(is_leaf); desc.append(np.packbits(bits))
                weight.append(0.0 if (is_leaf and rng.random() < stop_words) else float(rng.uniform(0.5, 9.0)))
                if not is_leaf:
                    nxt.append((nid, bits, lvl + 1))
        frontier = nxt
    return dict(k=k, L=L, parent=np.array(parent, np.int32), is_leaf=np.array(leaf, np.uint8),
                desc=np.stack(desc).astype(np.uint8), weight=np.array(weight, np.float64))


def sha256(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
