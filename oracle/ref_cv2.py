"""ORACLE A (test infrastructure): the reference's ORB front-end restated in Python on top of the REAL
OpenCV primitives (cv2 4.13.0 in the build container) — the same calls the reference makes:
cv2.resize / copyMakeBorder / FastFeatureDetector / GaussianBlur / fastAtan2 / norm / BFMatcher.

Used only (a) by oracle/make_golden.py to produce tests/golden/*.npz and (b) by
tests/test_oracle_vs_cv2.py to pin the dependency-free C oracle (oracle/orc_*.c) against OpenCV.
It is slow (Python loops) and never timed.  Follows reference src/ORBextractor.cc and src/Frame.cc;
line numbers are cited per function.
"""
from __future__ import annotations

import ctypes
import ctypes.util
import math
import pathlib

import cv2
import numpy as np

f32 = np.float32
EDGE = 19
HALF_PATCH = 15
PATCH = 31

_libm = ctypes.CDLL(ctypes.util.find_library("m"))
_libm.cosf.restype = ctypes.c_float; _libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float; _libm.sinf.argtypes = [ctypes.c_float]

_PAT = np.array([int(v) for v in "".join(
    l for l in (pathlib.Path(__file__).resolve().parents[1] / "include" / "orb_pattern_31.inc").read_text().splitlines()
    if not l.lstrip().startswith(("/*", "*"))).replace("\n", "").split(",") if v.strip()], np.int32).reshape(512, 2)


def cvround(v) -> int:
    """cvRound: round-half-even."""
    return int(np.rint(v))


class Params:
    """ORBextractor::ORBextractor, src/ORBextractor.cc:468-571."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, iniThFAST=20, minThFAST=7):
        self.nfeatures, self.nlevels, self.ini, self.min = nfeatures, nlevels, iniThFAST, minThFAST
        sfd = float(f32(scaleFactor))  # double member initialised from a float argument
        sf = [f32(1.0)]
        for i in range(1, nlevels):
            sf.append(f32(float(sf[-1]) * sfd))                      # :488 float*double -> float
        self.sf = np.array(sf, f32)
        self.sig2 = self.sf * self.sf
        self.isf = (f32(1.0) / self.sf).astype(f32)
        self.isig2 = (f32(1.0) / self.sig2).astype(f32)
        factor = f32(1.0 / sfd)                                      # :509
        nd = f32(f32(nfeatures) * (f32(1) - factor)) / f32(f32(1) - f32(math.pow(float(factor), float(nlevels))))  # :511
        nd = f32(nd)
        self.nper = []
        s = 0
        for _ in range(nlevels - 1):
            self.nper.append(cvround(nd)); s += self.nper[-1]
            nd = f32(nd * factor)
        self.nper.append(max(nfeatures - s, 0))
        vmax = int(math.floor(HALF_PATCH * math.sqrt(2.0) / 2 + 1))
        vmin = int(math.ceil(HALF_PATCH * math.sqrt(2.0) / 2))
        um = [0] * (HALF_PATCH + 1)
        for v in range(vmax + 1):
            um[v] = cvround(math.sqrt(HALF_PATCH * HALF_PATCH - v * v))
        v0 = 0
        for v in range(HALF_PATCH, vmin - 1, -1):
            while um[v0] == um[v0 + 1]:
                v0 += 1
            um[v] = v0
            v0 += 1
        self.umax = um


def compute_pyramid(img: np.ndarray, P: Params):
    """ComputePyramid, src/ORBextractor.cc:1664-1715.  Returns padded planes (the `temp` Mats)."""
    planes = []
    for level in range(P.nlevels):
        scale = P.isf[level]
        w, h = cvround(f32(img.shape[1]) * scale), cvround(f32(img.shape[0]) * scale)
        if level != 0:
            prev = planes[level - 1][EDGE:-EDGE, EDGE:-EDGE]
            cur = cv2.resize(prev, (w, h), interpolation=cv2.INTER_LINEAR)
            temp = cv2.copyMakeBorder(cur, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101 | cv2.BORDER_ISOLATED)
        else:
            temp = cv2.copyMakeBorder(img, EDGE, EDGE, EDGE, EDGE, cv2.BORDER_REFLECT_101)
        planes.append(temp)
    return planes


def detect_level(level_img: np.ndarray, P: Params):
    """Cell loop of ComputeKeyPointsOctTree, src/ORBextractor.cc:1038-1143 with real cv::FAST."""
    minB = EDGE - 3
    maxBX, maxBY = level_img.shape[1] - EDGE + 3, level_img.shape[0] - EDGE + 3
    W = f32(35)
    width, height = f32(maxBX - minB), f32(maxBY - minB)
    nCols, nRows = int(width / W), int(height / W)
    wCell, hCell = int(math.ceil(f32(width / f32(nCols)))), int(math.ceil(f32(height / f32(nRows))))
    f_ini = cv2.FastFeatureDetector_create(P.ini, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    f_min = cv2.FastFeatureDetector_create(P.min, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    out = []
    stats = dict(cells=0, retry=0, empty=0)
    for i in range(nRows):
        iniY = minB + i * hCell
        maxY = iniY + hCell + 6
        if iniY >= maxBY - 3:
            continue
        maxY = min(maxY, maxBY)
        for j in range(nCols):
            iniX = minB + j * wCell
            maxX = iniX + wCell + 6
            if iniX >= maxBX - 6:
                continue
            maxX = min(maxX, maxBX)
            sub = level_img[iniY:maxY, iniX:maxX]
            stats["cells"] += 1
            kps = f_ini.detect(sub, None)
            if len(kps) == 0:
                stats["retry"] += 1
                kps = f_min.detect(sub, None)
                if len(kps) == 0:
                    stats["empty"] += 1
            for kp in kps:
                out.append((int(kp.pt[0]) + j * wCell, int(kp.pt[1]) + i * hCell, int(kp.response)))
    return out, stats


class _Node:
    __slots__ = ("x0", "y0", "x1", "y1", "keys", "no_more", "seq")


def distribute_octree(K, minX, maxX, minY, maxY, N, reverse_ties=False):
    """DistributeOctTree, src/ORBextractor.cc:688-1034, literal port on a Python list used as std::list
    (index 0 = front).  Tie-break (size, creation seq) — see orc_octree.c."""
    seq = [0]

    def mk(x0, y0, x1, y1):
        n = _Node(); n.x0, n.y0, n.x1, n.y1 = x0, y0, x1, y1
        n.keys = []; n.no_more = False; n.seq = -1
        return n

    def divide(p):
        hx = int(math.ceil(f32(p.x1 - p.x0) / f32(2))); hy = int(math.ceil(f32(p.y1 - p.y0) / f32(2)))
        mx, my = p.x0 + hx, p.y0 + hy
        c = [mk(p.x0, p.y0, mx, my), mk(mx, p.y0, p.x1, my), mk(p.x0, my, mx, p.y1), mk(mx, my, p.x1, p.y1)]
        for k in p.keys:
            x, y = K[k][0], K[k][1]
            if x < mx:
                c[0 if y < my else 2].keys.append(k)
            else:
                c[1 if y < my else 3].keys.append(k)
        for n in c:
            if len(n.keys) == 1:
                n.no_more = True
        return c

    nIni = int(np.round(f32(maxX - minX) / f32(maxY - minY)))  # C round(): half away from zero
    r = float(f32(maxX - minX) / f32(maxY - minY))
    nIni = int(math.floor(r + 0.5))
    hX = f32(maxX - minX) / f32(nIni)
    L = []
    ini = []
    for i in range(nIni):
        n = mk(int(f32(hX * f32(i))), 0, int(f32(hX * f32(i + 1))), maxY - minY)
        n.seq = seq[0]; seq[0] += 1
        L.append(n); ini.append(n)
    for k, kp in enumerate(K):
        ini[int(f32(kp[0]) / hX)].keys.append(k)
    L2 = []
    for n in L:
        if len(n.keys) == 1:
            n.no_more = True; L2.append(n)
        elif len(n.keys) > 1:
            L2.append(n)
    L = L2

    def push_children(L, c, vec):
        nexp = 0
        for n in c:
            if len(n.keys) > 0:
                n.seq = seq[0]; seq[0] += 1
                L.insert(0, n)
                if len(n.keys) > 1:
                    nexp += 1
                    vec.append(n)
        return nexp

    finish = False
    while not finish:
        prevSize = len(L)
        nToExpand = 0
        vec = []
        todo = [n for n in L if not n.no_more]     # children go to the front: never revisited
        for n in todo:
            c = divide(n)
            nToExpand += push_children(L, c, vec)
            L.remove(n)
        if len(L) >= N or len(L) == prevSize:
            finish = True
        elif len(L) + nToExpand * 3 > N:
            while not finish:
                prevSize = len(L)
                prev = sorted(vec, key=lambda n: (len(n.keys), -n.seq if reverse_ties else n.seq))
                vec = []
                for n in reversed(prev):
                    c = divide(n)
                    push_children(L, c, vec)
                    L.remove(n)
                    if len(L) >= N:
                        break
                if len(L) >= N or len(L) == prevSize:
                    finish = True
    res = []
    for n in L:
        best = n.keys[0]
        for k in n.keys[1:]:
            if K[k][2] > K[best][2]:
                best = k
        res.append(K[best])
    return res


def ic_angle(plane: np.ndarray, x: int, y: int, umax) -> float:
    """IC_Angle, src/ORBextractor.cc:91-138; (x, y) in padded-plane coordinates."""
    m01 = 0; m10 = 0
    row = plane[y].astype(np.int64)
    us = np.arange(-HALF_PATCH, HALF_PATCH + 1)
    m10 += int((us * row[x - HALF_PATCH:x + HALF_PATCH + 1]).sum())
    for v in range(1, HALF_PATCH + 1):
        d = umax[v]
        us = np.arange(-d, d + 1)
        p = plane[y + v, x - d:x + d + 1].astype(np.int64)
        m = plane[y - v, x - d:x + d + 1].astype(np.int64)
        m01 += v * int((p - m).sum())
        m10 += int((us * (p + m)).sum())
    return cv2.fastAtan2(float(m01), float(m10))


def orb_descriptor(blurred: np.ndarray, x: int, y: int, angle_deg) -> np.ndarray:
    """computeOrbDescriptor, src/ORBextractor.cc:150-203 (float32, no FMA, libm cosf/sinf)."""
    factorPI = f32(math.pi / 180.0)
    ang = f32(f32(angle_deg) * factorPI)
    a = f32(_libm.cosf(ctypes.c_float(float(ang)))); b = f32(_libm.sinf(ctypes.c_float(float(ang))))
    px = _PAT[:, 0].astype(f32); py = _PAT[:, 1].astype(f32)
    ry = np.rint((px * b).astype(f32) + (py * a).astype(f32)).astype(np.int32)
    rx = np.rint((px * a).astype(f32) - (py * b).astype(f32)).astype(np.int32)
    vals = blurred[y + ry, x + rx].astype(np.int32)
    bits = (vals[0::2] < vals[1::2]).astype(np.uint8)          # 256 tests
    return np.packbits(bits.reshape(32, 8)[:, ::-1], axis=1).reshape(32)


KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])


def extract(img: np.ndarray, P: Params, lapping=(0, 0), reverse_ties=False):
    """ORBextractor::operator(), src/ORBextractor.cc:1534-1659.  Returns dict of all stages."""
    planes = compute_pyramid(img, P)
    minB = EDGE - 3
    cands, sels, angles, blurs, stats = [], [], [], [], []
    for level in range(P.nlevels):
        limg = planes[level][EDGE:-EDGE, EDGE:-EDGE]
        c, st = detect_level(limg, P)
        cands.append(c); stats.append(st)
        s = distribute_octree(c, minB, limg.shape[1] - EDGE + 3, minB, limg.shape[0] - EDGE + 3, P.nper[level], reverse_ties)
        sels.append(s)
    for level in range(P.nlevels):
        angles.append([ic_angle(planes[level], s[0] + minB + EDGE, s[1] + minB + EDGE, P.umax) for s in sels[level]])
    n = sum(len(s) for s in sels)
    kps = np.zeros(n, KP_DTYPE); desc = np.zeros((n, 32), np.uint8)
    mono, stereo = 0, n - 1
    for level in range(P.nlevels):
        if len(sels[level]) == 0:
            blurs.append(None); continue
        limg = planes[level][EDGE:-EDGE, EDGE:-EDGE]
        work = cv2.GaussianBlur(limg.copy(), (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
        blurs.append(work)
        scale = P.sf[level]
        sps = int(f32(PATCH) * scale)
        for i, s in enumerate(sels[level]):
            xl, yl = s[0] + minB, s[1] + minB
            d = orb_descriptor(work, xl, yl, angles[level][i])
            x, y = f32(xl), f32(yl)
            if level != 0:
                x = f32(x * scale); y = f32(y * scale)
            if lapping[0] <= x <= lapping[1]:
                dst = stereo; stereo -= 1
            else:
                dst = mono; mono += 1
            kps[dst] = (x, y, sps, angles[level][i], s[2], level, -1)
            desc[dst] = d
    return dict(mono=mono, kps=kps, desc=desc, planes=planes, cands=cands, sels=sels, angles=angles, blurs=blurs, stats=stats)


def stereo_matches(resL, resR, P: Params, mb: float, mbf: float):
    """Frame::ComputeStereoMatches, src/Frame.cc:1027-1276, with cv2.norm for the SAD."""
    kL, kR, dL, dR = resL["kps"], resR["kps"], resL["desc"], resR["desc"]
    N = len(kL)
    uRight = np.full(N, -1, f32); depth = np.full(N, -1, f32)
    nRows = resL["planes"][0].shape[0] - 2 * EDGE
    rows = [[] for _ in range(nRows)]
    for iR in range(len(kR)):
        r = f32(2.0) * P.sf[kR["octave"][iR]]
        maxr = int(math.ceil(f32(kR["y"][iR] + r))); minr = int(math.floor(f32(kR["y"][iR] - r)))
        for yi in range(minr, maxr + 1):
            rows[yi].append(iR)
    mb, mbf = f32(mb), f32(mbf)
    maxD = f32(mbf / mb)
    vd = []
    pop = np.array([bin(i).count("1") for i in range(256)], np.int32)
    for iL in range(N):
        lvl = int(kL["octave"][iL]); vL = kL["y"][iL]; uL = kL["x"][iL]
        cand = rows[int(vL)]
        if not cand:
            continue
        minU = f32(uL - maxD); maxU = uL
        if maxU < 0:
            continue
        best, bestR = 100, 0
        for iR in cand:
            if kR["octave"][iR] < lvl - 1 or kR["octave"][iR] > lvl + 1:
                continue
            uR = kR["x"][iR]
            if minU <= uR <= maxU:
                d = int(pop[dL[iL] ^ dR[iR]].sum())
                if d < best:
                    best, bestR = d, iR
        if best < 75:
            uR0 = kR["x"][bestR]
            s = P.isf[lvl]
            rnd = lambda v: f32(math.floor(float(v) + 0.5))  # roundf for positive values
            suL, svL, suR0 = rnd(f32(kL["x"][iL] * s)), rnd(f32(kL["y"][iL] * s)), rnd(f32(uR0 * s))
            w = 5; Lr = 5
            imL = resL["planes"][lvl][EDGE:-EDGE, EDGE:-EDGE]; imR = resR["planes"][lvl][EDGE:-EDGE, EDGE:-EDGE]
            IL = imL[int(svL) - w:int(svL) + w + 1, int(suL) - w:int(suL) + w + 1]
            if suR0 + Lr - w < 0 or suR0 + Lr + w + 1 >= imR.shape[1]:
                continue
            bestS, bestinc = 2 ** 31 - 1, 0
            dists = [f32(0)] * 11
            for inc in range(-Lr, Lr + 1):
                IR = imR[int(svL) - w:int(svL) + w + 1, int(suR0) + inc - w:int(suR0) + inc + w + 1]
                dist = f32(cv2.norm(IL, IR, cv2.NORM_L1))
                if dist < bestS:
                    bestS, bestinc = int(dist), inc
                dists[Lr + inc] = dist
            if bestinc in (-Lr, Lr):
                continue
            d1, d2, d3 = dists[Lr + bestinc - 1], dists[Lr + bestinc], dists[Lr + bestinc + 1]
            delta = f32(f32(d1 - d3) / f32(f32(2.0) * f32(f32(d1 + d3) - f32(f32(2.0) * d2))))
            if delta < -1 or delta > 1:
                continue
            bestuR = f32(P.sf[lvl] * f32(f32(suR0 + f32(bestinc)) + delta))
            disp = f32(uL - bestuR)
            if 0 <= disp < maxD:
                if disp <= 0:
                    disp = f32(0.01); bestuR = f32(float(uL) - 0.01)
                depth[iL] = f32(mbf / disp); uRight[iL] = bestuR
                vd.append((bestS, iL))
    if vd:
        vd.sort()
        median = f32(vd[len(vd) // 2][0])
        th = f32(f32(f32(1.5) * f32(1.4)) * median)
        for s_, i_ in reversed(vd):
            if f32(s_) < th:
                break
            uRight[i_] = -1; depth[i_] = -1
    return uRight, depth
