"""ORBX_ZEROCOPY experiment: results of orbx_extract_batch on pinned frames must not depend on how the frames reach the device."""
import importlib, os, sys
from pathlib import Path
import numpy as np, torch
sys.path.insert(0, str(Path(__file__).resolve().parents[2]))
pkg = importlib.import_module("orb-slam3_b200")
frames = np.stack([pkg.synth.synth_frame(480, 752, s) for s in range(8)] * 40)            # 320 frames: several chunks
pin = torch.from_numpy(frames).pin_memory().numpy()
X = pkg.ORBextractor(1000)
out = X.extract_batch(pin, (0, 1000))
ref = X.extract_batch(frames, (0, 1000))
ok = all(np.array_equal(a, b) for a, b in zip(out, ref))
print("zerocopy" if os.environ.get("ORBX_ZEROCOPY") else "copy", "identical to pageable path:", ok)
sys.exit(0 if ok else 1)
