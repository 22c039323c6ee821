"""Import helper: the package directory is `orb-slam3_b200/` (not a valid identifier)."""
import importlib.util
import pathlib
import sys

ROOT = pathlib.Path(__file__).resolve().parent


def load():
    if "orbslam3_b200" in sys.modules:
        return sys.modules["orbslam3_b200"]
    d = ROOT / "orb-slam3_b200"
    spec = importlib.util.spec_from_file_location("orbslam3_b200", d / "__init__.py", submodule_search_locations=[str(d)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["orbslam3_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


def load_oracle():
    """ORACLE binding (tests / smoke / bench cpu_baseline only)."""
    p = str(ROOT / "oracle")
    if p not in sys.path:
        sys.path.insert(0, p)
    import orc_py
    return orc_py


def load_synth_only():
    """The synthetic-input generator without touching liborbx.so (reference arm / CPU-only paths)."""
    import types
    if "orbslam3_b200" in sys.modules:
        return sys.modules["orbslam3_b200"]
    d = ROOT / "orb-slam3_b200"
    spec = importlib.util.spec_from_file_location("orbslam3_b200_synth", d / "synth.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    ns = types.SimpleNamespace(synth=m)
    return ns
