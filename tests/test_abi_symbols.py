"""CPU: the C-ABI library builds (nvcc cross-compile), loads, and exports every symbol include/sbayes_b200.h
declares.  No compute call is made here (no GPU)."""
import os
import re

from conftest import ROOT


def test_header_symbols_exported(built_lib):
    from contact_zones_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'sbayes_b200.h')).read()
    declared = set(re.findall(r'\b(sb_[a-z_0-9]+)\s*\(', header))
    declared -= {'sb_ctx'}
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for s in declared:
        assert hasattr(built_lib, s), s
    assert built_lib.sb_version() == 100


def test_no_gpu_fails_loudly(built_lib):
    """Without a CUDA device the product path raises (no CPU fallback)."""
    import numpy as np
    import pytest
    from contact_zones_b200.engine import Engine
    from contact_zones_b200._lib import SbError
    if built_lib.sb_device_count() > 0:
        pytest.skip('a GPU is present')
    with pytest.raises(SbError):
        Engine(np.zeros((4, 2), np.uint8), np.ones((2, 2), bool), n_zones=1)


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, 'contact_zones_b200')
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith('.py'):
                src = open(os.path.join(dirpath, f)).read()
                assert 'sbayes_oracle' not in src and 'refshim' not in src, f
