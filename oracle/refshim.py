"""TEST INFRASTRUCTURE ONLY -- import the unmodified reference (`/root/reference/sbayes`) in this
container so that (a) the numpy restatement in `oracle/sbayes_oracle.py` can be validated against it and
(b) golden vectors can be generated (`oracle/make_golden.py`).  `/root/reference` does not exist on the GPU
box: nothing under `tests/ -m gpu`, `bench.py` or `__graft_entry__.smoke()` imports this module.

Recipe = SURVEY.md Appendix B: four stub modules (matplotlib, fastcluster, pyproj, pycldf) that the
reference imports at module import time but never uses on the MCMC path, and three monkeypatches for
APIs removed from Python 3.11 / pandas 3 (`open(..., 'rU')`, `random.sample(set, k)`, `DataFrame.applymap`).
"""
import builtins
import os
import random
import sys
import types

REFERENCE_ROOT = os.environ.get("SBAYES_REFERENCE_ROOT", "/root/reference")
_installed = False


def available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "sbayes"))


def _stub(name, **attrs):
    if name in sys.modules:
        return sys.modules[name]
    m = types.ModuleType(name)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


def install():
    """Register stubs + monkeypatches and put the reference on sys.path (idempotent)."""
    global _installed
    if _installed:
        return
    if not available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    try:
        import matplotlib  # noqa: F401
    except ImportError:
        dummy = lambda *a, **k: None
        plt = _stub("matplotlib.pyplot", get_cmap=dummy, Normalize=dummy, gca=dummy)
        col = _stub("matplotlib.collections", LineCollection=object)
        _stub("matplotlib", pyplot=plt, collections=col)
    try:
        import fastcluster  # noqa: F401
    except ImportError:
        _stub("fastcluster", linkage=lambda *a, **k: None)
    try:
        import pyproj  # noqa: F401
    except ImportError:
        tr = _stub("pyproj.transformer", Transformer=object)
        crs = _stub("pyproj.crs", CRS=object)
        _stub("pyproj", CRS=object, Transformer=object, transformer=tr, crs=crs)
    try:
        import pycldf  # noqa: F401
    except ImportError:
        _stub("pycldf", StructureDataset=object, Dataset=object)

    # (1) open(..., 'rU')  (preprocessing.py:39)
    _open = builtins.open

    def open_no_u(file, mode="r", *a, **k):
        if isinstance(mode, str) and "U" in mode:
            mode = mode.replace("U", "") or "r"
        return _open(file, mode, *a, **k)

    builtins.open = open_no_u

    # (2) random.sample(set, k)  (zone_sampling.py:1015)
    _sample = random.sample

    def sample_any(population, k, **kw):
        if isinstance(population, (set, frozenset)):
            population = sorted(population)
        return _sample(population, k, **kw)

    random.sample = sample_any

    # (3) DataFrame.applymap  (util.py:356,369)
    import pandas as pd
    if not hasattr(pd.DataFrame, "applymap"):
        pd.DataFrame.applymap = pd.DataFrame.map

    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    _installed = True


def modules():
    """Return the reference's hot-path modules as a namespace."""
    install()
    import sbayes.model as model
    import sbayes.util as util
    import sbayes.preprocessing as preprocessing
    import sbayes.sampling.zone_sampling as zone_sampling
    import sbayes.sampling.mcmc_generative as mcmc_generative
    import sbayes.mcmc_setup as mcmc_setup
    import sbayes.postprocessing as postprocessing
    return types.SimpleNamespace(model=model, util=util, preprocessing=preprocessing,
                                 zone_sampling=zone_sampling, mcmc_generative=mcmc_generative,
                                 mcmc_setup=mcmc_setup, postprocessing=postprocessing)
