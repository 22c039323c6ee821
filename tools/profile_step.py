#!/usr/bin/env python
"""One resident extraction step repeated a few times, for ncu: `ncu --launch-skip <13 * (reps - 1)> -c 13 ... python tools/profile_step.py [frames] [reps] [C1..C5]`
captures exactly the last, warm step (7 pyramid launches, FAST, quadtree, assemble, blur, describe = 13 launches per step of C1)."""
import importlib, sys
from pathlib import Path
import numpy as np
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
pkg = importlib.import_module("orb-slam3_b200")
nfr = int(sys.argv[1]) if len(sys.argv) > 1 else 32
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
c = pkg.synth.CONFIGS[sys.argv[3] if len(sys.argv) > 3 else "C1"]
frames = np.stack([pkg.synth.synth_frame(c["rows"], c["cols"], s) for s in range(nfr)])
X = pkg.ORBextractor(c["nfeatures"])
X.upload(frames)
for _ in range(reps):
    X.run_resident((0, 1000)); X.sync()
print("launches", X.launches() if hasattr(X, "launches") else "n/a")
X.close()
