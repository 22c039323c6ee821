"""CPU suite: bench.py's reference arm / JSON contract and the N>1 sharding path on gloo (world_size 2)."""
import json
import os
import pathlib
import subprocess
import sys

import numpy as np
import pytest

ROOT = pathlib.Path(__file__).resolve().parents[1]
REQUIRED = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline", "gpu_launches"}


def _run(cmd, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, env=e, timeout=600)


def test_reference_arm_json_contract():
    r = _run([sys.executable, "bench.py", "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-frames-per-core", "1"])
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    j = json.loads(lines[0])
    assert REQUIRED <= set(j) and j["impl"] == "reference" and j["unit"] == "frames/s" and j["value"] > 0
    assert j["cpu_baseline"]["kind"] in ("reference", "port") and j["cpu_baseline"]["cores"] >= 1
    if (ROOT / "oracle" / "_ref" / "libref_native.so").exists():
        assert j["cpu_baseline"]["kind"] == "reference" and j["cpu_baseline"]["port_vs_cv2"] > 0
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["e2e"]["value"] == j["value"]
    assert "752x480" in j["config"]["workload"] and j["vs_baseline"] is None


def test_reference_arm_under_torchrun_only_rank0_prints():
    r = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
              "--master-port", "29533", "bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
              "--ref-frames-per-core", "1"])
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1 and json.loads(lines[0])["n_gpus"] == 2


def test_ours_arm_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    r = _run([sys.executable, "bench.py", "--steps", "1", "--warmup", "0"])
    assert r.returncode != 0 and "no CPU path" in (r.stderr + r.stdout)


def test_shard_range_covers_units_exactly(pkg):
    from orbslam3_b200.shard import shard_range
    for n in (0, 1, 7, 8, 9, 1000):
        for world in (1, 2, 3, 8):
            got = [i for r in range(world) for i in shard_range(n, r, world)]
            assert got == list(range(n))
            sizes = [len(shard_range(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(5, 2, 2)


_WORKER = r"""
import os, sys, json
sys.path.insert(0, os.environ["ORBX_ROOT"])
import numpy as np, torch.distributed as dist
import _pkg
pkg = _pkg.load_synth_only()
import importlib.util
spec = importlib.util.spec_from_file_location("shard", os.path.join(os.environ["ORBX_ROOT"], "orb-slam3_b200", "shard.py"))
shard = importlib.util.module_from_spec(spec); spec.loader.exec_module(shard)
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
orc = _pkg.load_oracle(); o = orc.Oracle()
units = shard.shard_range(5, rank, world)           # 5 frames over 2 ranks: 3 + 2
E = orc.Extractor(300, oracle=o)
n = 0
for u in units:                                     # each rank extracts only its own frames (no exchange)
    n += len(E(pkg.synth.synth_frame(256, 320, u))[1])
shard.barrier()
elapsed = 10.0 * (rank + 1)                         # pretend device times: max over ranks = 20 ms
thr = shard.whole_job_throughput(len(units), elapsed)
tot = shard.dist_reduce(n, "sum")
if rank == 0:
    print(json.dumps({"thr": thr, "total_kp": tot, "world": world}))
dist.destroy_process_group()
"""


def test_gloo_world2_sharded_run(tmp_path, orc):
    w = tmp_path / "worker.py"
    w.write_text(_WORKER)
    r = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
              "--master-port", "29534", str(w)], env={"ORBX_ROOT": str(ROOT)})
    assert r.returncode == 0, r.stderr[-3000:]
    j = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert j["world"] == 2 and abs(j["thr"] - 5 / 0.020) < 1e-6        # 5 units / max(10, 20) ms
    o = orc.Oracle()
    import _pkg
    synth = _pkg.load_synth_only().synth
    E = orc.Extractor(300, oracle=o)
    assert j["total_kp"] == sum(len(E(synth.synth_frame(256, 320, u))[1]) for u in range(5))
