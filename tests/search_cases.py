"""Input builders shared by oracle/make_ref_golden.py (which runs them through the compiled reference, oracle/_ref/libref_match.so)
and tests/test_ref_golden.py (which runs them through the C oracle on the CPU and through the CUDA path on the GPU and compares with
the vectors the reference produced).  Every builder is a pure function of its seed; `mod` is any module with KP_DTYPE (orc_py or
the package)."""
import numpy as np

from test_gpu_search import SF, SIGMA2, _feature_vector, _flip, _frame, _moved, _projected_rows

ROWS, COLS = 480, 752
BOUNDS = (0.0, float(COLS), 0.0, float(ROWS))


def projection_frame(mod, seed=12):
    """ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono), stereo, forward (src/ORBmatcher.cc:2469-2711)"""
    rng = np.random.default_rng(seed)
    n = 1000
    lk, ld, _, _ = _frame(mod, rng, n, ROWS, COLS)
    ck, cd, _ = _moved(mod, rng, lk, ld, ROWS, COLS)
    uR = np.where(rng.random(n) < 0.7, ck["x"] - rng.uniform(1, 40, n), -1).astype(np.float32)
    invz = np.where(rng.random(n) < 0.97, 2.0 ** -rng.integers(1, 6, n), -0.1).astype(np.float32)   # exact under 1 / (1 / invz)
    last = dict(desc=_flip(rng, ld.copy(), 0, 10),
                u=(lk["x"] + 4.0 + rng.normal(0, 3, n)).astype(np.float32), v=(lk["y"] - 3.0 + rng.normal(0, 3, n)).astype(np.float32),
                invz=invz, octave=lk["octave"].astype(np.int32), angle=lk["angle"],
                valid=(rng.random(n) < 0.85).astype(np.uint8), obs=(rng.random(n) < 0.9).astype(np.uint8))
    held = (rng.random(n) < 0.03).astype(np.uint8)
    return dict(ck=ck, cd=cd, uR=uR, last=last, held=held, th=7.0, mbf=40.0, fwd=True, bwd=False, orient=True)


def bow(mod, seed=31):
    """ORBmatcher::SearchByBoW(KeyFrame*, Frame&, ...) (src/ORBmatcher.cc:314-547)"""
    rng = np.random.default_rng(seed)
    n = 1200
    k1, d1, pool, src = _frame(mod, rng, n, ROWS, COLS)
    k2, d2, perm = _moved(mod, rng, k1, d1, ROWS, COLS, flips=(0, 50))
    node1 = (src % 89).astype(np.int32); node2 = node1[perm]
    node2 = np.where(rng.random(n) < 0.05, rng.integers(0, 89, n), node2).astype(np.int32)
    kf1 = dict(kps=k1, desc=d1, valid=(rng.random(n) < 0.7).astype(np.uint8))
    kf2 = dict(kps=k2, desc=d2, valid=None)
    kf1["fv_node"], kf1["fv_start"], kf1["fv_idx"] = _feature_vector(rng, node1, 0.02)
    kf2["fv_node"], kf2["fv_start"], kf2["fv_idx"] = _feature_vector(rng, node2, 0.02)
    return dict(kf1=kf1, kf2=kf2, nnratio=0.7, orient=True)


def fuse_sim3(mod, seed=73):
    """ORBmatcher::Fuse(KeyFrame*, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:2051-2185)"""
    rng = np.random.default_rng(seed)
    n = 1800
    kk, kd, _, _ = _frame(mod, rng, n, ROWS, COLS, quant=0.5)
    ck, cd, _ = _moved(mod, rng, kk, kd, ROWS, COLS)
    mp = _projected_rows(mod, rng, kk, kd, n, ROWS, COLS)
    return dict(ck=ck, cd=cd, mp=mp, th=4.0, isig2=(1.0 / SIGMA2).astype(np.float32))


def sim3(mod, seed=81):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:2201-2434)"""
    rng = np.random.default_rng(seed)
    n = 1400
    k1, d1, _, _ = _frame(mod, rng, n, ROWS, COLS)
    k2, d2, perm = _moved(mod, rng, k1, d1, ROWS, COLS, flips=(0, 25))
    inv = np.empty(n, np.int64); inv[perm] = np.arange(n)
    mk = lambda d, ka, idx: dict(desc=_flip(rng, d.copy(), 0, 10), u=(ka["x"][idx] + rng.normal(0, 1.5, n)).astype(np.float32),
                                 v=(ka["y"][idx] + rng.normal(0, 1.5, n)).astype(np.float32),
                                 level=np.clip(ka["octave"][idx] + rng.choice([0, 0, 1], n), 0, 7).astype(np.int32),
                                 valid=(rng.random(n) < 0.8).astype(np.uint8))
    rows1 = mk(d1, k2, inv); rows2 = mk(d2, k1, perm)
    return dict(k1=k1, d1=d1, k2=k2, d2=d2, rows1=rows1, rows2=rows2, th=7.5)


def run_oracle(name, orc, o, c):
    """through orc_py on the Oracle handle `o` (liborc.so, or the proxy that routes to the compiled reference)"""
    if name == "projection_frame":
        g = orc.Grid(o, c["ck"], *BOUNDS)
        nm, a = orc.search_by_projection_frame(g, c["cd"], c["uR"], SF, c["last"], c["th"], c["mbf"], c["fwd"], c["bwd"], c["orient"], c["held"])
        return nm, a
    if name == "bow":
        return orc.search_by_bow(o, 0, c["kf1"], c["kf2"], c["nnratio"], c["orient"])
    if name == "fuse_sim3":
        nf, bi, _ = orc.fuse(orc.Grid(o, c["ck"], *BOUNDS), c["cd"], None, SF, c["isig2"], c["mp"], c["th"])
        return nf, bi
    if name == "sim3":
        return orc.search_by_sim3(orc.Grid(o, c["k1"], *BOUNDS), c["d1"], orc.Grid(o, c["k2"], *BOUNDS), c["d2"], SF, c["rows1"], c["rows2"], c["th"])
    raise KeyError(name)


def run_gpu(name, pkg, c):
    if name == "projection_frame":
        return pkg.SearchByProjectionFrame(c["ck"], c["cd"], c["uR"], BOUNDS, SF, c["last"], c["th"], c["mbf"], c["fwd"], c["bwd"], c["orient"], c["held"])
    if name == "bow":
        return pkg.SearchByBoW(0, c["kf1"], c["kf2"], c["nnratio"], c["orient"])
    if name == "fuse_sim3":
        nf, bi, _ = pkg.Fuse(c["ck"], c["cd"], None, BOUNDS, SF, c["isig2"], c["mp"], c["th"])
        return nf, bi
    if name == "sim3":
        return pkg.SearchBySim3(c["k1"], c["d1"], c["k2"], c["d2"], BOUNDS, SF, c["rows1"], c["rows2"], c["th"])
    raise KeyError(name)


CASES = {"projection_frame": projection_frame, "bow": bow, "fuse_sim3": fuse_sim3, "sim3": sim3}


def digest(c):
    """hash of every array of a case (nested dicts included): the goldens are tied to these exact inputs"""
    import hashlib
    h = hashlib.sha256()
    def walk(x):
        if isinstance(x, dict):
            for k in sorted(x):
                h.update(k.encode()); walk(x[k])
        elif isinstance(x, np.ndarray):
            h.update(np.ascontiguousarray(x).tobytes())
        else:
            h.update(repr(x).encode())
    walk(c)
    return h.hexdigest()
