"""Stage-by-stage GPU-vs-oracle diff (debug helper, run on the GPU box)."""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parents[1]))
import numpy as np
import _pkg
pkg = _pkg.load(); orc = _pkg.load_oracle(); orc.build(); o = orc.Oracle()
rows, cols, nf, lap, seed = 480, 752, 1000, (0, 1000), 0
if len(sys.argv) > 1: rows, cols, nf, seed = map(int, sys.argv[1:5])
img = pkg.synth.synth_frame(rows, cols, seed)
E = orc.Extractor(nf, oracle=o); mo, ko, do = E(img, lap)
X = pkg.ORBextractor(nf); mg, kg, dg = X(img, lap)
print("mono", mg, mo, "n", len(kg), len(ko))
def arr(c): return np.stack([c["x"], c["y"], c["score"]], 1).astype(np.int32) if len(c) else np.empty((0,3),np.int32)
for l in range(8):
    pg, po = X.pyramid_level(l, padded=True), E.level_plane(l)
    d = pg != po
    print(f"L{l} plane shape {pg.shape} {po.shape} mism {d.sum()}", end=" ")
    if d.any():
        ys, xs = np.nonzero(d); print("first", ys[0], xs[0], pg[ys[0], xs[0]], po[ys[0], xs[0]], "rows", ys.min(), ys.max(), "cols", xs.min(), xs.max(), end=" ")
    cg, co = X.level_keys(l, 0), arr(E.level_candidates(l))
    print(f"| cand {len(cg)} {len(co)} eq {np.array_equal(cg, co)}", end=" ")
    if not np.array_equal(cg, co):
        sg = set(map(tuple, cg)); so = set(map(tuple, co))
        print("only_gpu", len(sg - so), "only_orc", len(so - sg), list(sg - so)[:3], list(so - sg)[:3], end=" ")
        m = min(len(cg), len(co)); nz = np.nonzero((cg[:m] != co[:m]).any(1))[0]
        if len(nz): print("first diff idx", nz[0], cg[nz[0]], co[nz[0]], end=" ")
    (sg_, ag), (so_, ao) = X.level_keys(l, 1), E.level_selected(l)
    so_ = arr(so_)
    print(f"| sel {len(sg_)} {len(so_)} eq {np.array_equal(sg_, so_)}", end=" ")
    if not np.array_equal(sg_, so_):
        a = set(map(tuple, sg_)); b = set(map(tuple, so_)); print("only_gpu", len(a - b), "only_orc", len(b - a), end=" ")
    if len(ag) == len(ao): print("| ang mism", int((ag != ao).sum()), end=" ")
    b = E.level_blurred(l)
    if b is not None:
        bg = X.blurred_level(l); print("| blur mism", int((bg != b).sum()), end="")
    print()
if len(kg) == len(ko):
    for f in ko.dtype.names: print(f, int((kg[f] != ko[f]).sum()), end="; ")
    print("desc rows differing", int((dg != do).any(1).sum()))
# cell (0,0) of level 0 in detail
cg, co = X.level_keys(0, 0), arr(E.level_candidates(0))
print("GPU first 12:", cg[:12].tolist())
print("ORC first 12:", co[:12].tolist())
p = E.level_plane(0)[19:, 19:]
sub = p[16:16 + 44, 16:16 + 42]
print("oracle FAST(20) on cell00:", [tuple(int(v) for v in c) for c in o.fast(sub, 20)][:12])
print("oracle FAST(20) no-NMS on cell00 count:", len(o.fast(sub, 20, False)))
