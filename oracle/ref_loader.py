"""Loader for oracle/_ref (the reference's own C++ searchsorted, compiled by build_ref.py).
TEST INFRASTRUCTURE ONLY — see oracle/nerf_oracle.py header."""
import glob
import importlib.util
import os

_HERE = os.path.dirname(os.path.abspath(__file__))


def load_reference_searchsorted():
    """Return the compiled reference module (exports searchsorted_cpu_wrapper) or None."""
    cands = sorted(glob.glob(os.path.join(_HERE, "_ref", "tss_ref_cpu*.so")))
    if not cands:
        return None
    import torch  # noqa: F401  (libtorch symbols must be loaded first)
    spec = importlib.util.spec_from_file_location("tss_ref_cpu", cands[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
