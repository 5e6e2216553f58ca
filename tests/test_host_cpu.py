"""CPU tests of the host side: the C-ABI library loads and exports every symbol declared in
include/sgnn_b200.h; state_dict keys match the reference's checkpoint format; no product
module imports the oracle."""
import ctypes
import os
import re

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from sgnn_b200 import _lib as L
    lib = ctypes.CDLL(L.LIB_PATH)
    hdr = open(os.path.join(ROOT, "include", "sgnn_b200.h")).read()
    names = sorted(set(re.findall(r"\b(sgnn_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 12
    for n in names:
        assert hasattr(lib, n), n
    assert lib.sgnn_sm_arch() == 100


def test_state_dict_keys_match_reference_checkpoints():
    from sgnn_b200.single_scale import LearnedSimulator
    from tests.util import Fixture
    f = Fixture("single_h64")
    sd = f.state_dict()
    stats = f.stats()
    sim = LearnedSimulator(2, 21, 3, 64, 5, 1, 64, 15.0, stats, 1, 9)
    mine = sim.state_dict()
    assert list(mine.keys()) == list(sd.keys())
    for k in sd:
        assert mine[k].shape == sd[k].shape, k
    sim.load_state_dict(sd)
    assert sum(p.numel() for p in sim.parameters()) == 160524


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "sgnn_b200")):
        for fn in files:
            if fn.endswith(".py"):
                src = open(os.path.join(dirpath, fn)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), fn
