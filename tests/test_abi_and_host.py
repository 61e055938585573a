"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/viba_b200.h declares (no compute calls without a GPU), compute entry points fail loudly
without a device, and the host-only symbolic phase is consistent with the oracle's structure."""
import ctypes as C
import os
import re
import numpy as np
import pytest
from tests.oracle_lib import Oracle
from visual_inertial_bundle_adjustment_b200 import capi, synth, graph as G

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "viba_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    lib = capi.load()
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/viba_b200.h but not exported"
    for n in capi.EXPORTS:
        assert n in names, f"{n} bound by capi.py but not declared in the header"


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(capi.VbError) as e:
        capi.Context(0)
    assert e.value.code == 2 and "no CPU fallback" in str(e.value)


def test_settings_default_match_reference_library_defaults():
    lib = capi.load()
    s = G.vb_settings()
    lib.vb_settings_default(C.byref(s))
    # lib/small_thing/Optimizer.h:40-78
    assert (s.damping, s.damping_adjust_on_fail, s.damping_adjust_on_good, s.damping_adjust_on_average) == (1e-5, 2.5, 0.7, 1.5)
    assert (s.damping_min, s.damping_max, s.min_relative_cost_reduction) == (1e-9, 1e8, 0.3)
    assert (s.max_step_factor_attempts, s.try_sub_step, s.max_num_iterations) == (2, 1, 50)


@pytest.mark.parametrize("name", ["tiny", "small"])
def test_symbolic_phase_structure(oracle, name):
    s = synth.make_session(**synth.CONFIGS[name])
    o = Oracle(oracle, s.graph)
    order, data_size, lumps, perm = capi.symbolic_only(s.graph)
    assert order == o.order
    assert sorted(perm) == list(range(o.nblocks))
    assert np.array_equal(perm[:s.n_points], np.arange(s.n_points))  # elimRanges = [0, P) first
    st = capi.symbolic_stats(s.graph)
    assert st["levels"] >= 1 and st["lumps"] >= st["levels"] and data_size >= o.data_size - 9 * s.n_points
    # nested dissection never loses a block and RCM (leaf = 0) gives the same order/dataSize class
    st0 = capi.symbolic_stats(s.graph, nd_leaf_blocks=0)
    assert st0["order"] == st["order"]


def test_malformed_graph_is_rejected():
    s = synth.make_session(**synth.CONFIGS["tiny"])
    g = s.graph
    bad = g.var_index.copy()
    bad[s.ids["pose"][0]] = bad[s.ids["pose"][1]]  # duplicate registration index
    keep = g.var_index
    g.var_index = bad
    try:
        with pytest.raises(capi.VbError) as e:
            capi.symbolic_only(g)
        assert e.value.code == 1
    finally:
        g.var_index = keep


def test_registration_rule_follows_reference_order():
    """registerAllVariables: per store, per argument position, per factor in insertion order
    (lib/small_thing/Optimizer.cpp:25-39, Factor.h:737-758)."""
    s = synth.make_session(**synth.CONFIGS["tiny"])
    g = s.graph
    nxt = s.n_points
    idx = {int(v): i for i, v in enumerate(s.ids["point"])}
    for st in g.stores:
        for a in range(st.vars.shape[1]):
            for v in st.vars[:, a]:
                v = int(v)
                if v in idx or g.var_index[v] == -2:
                    continue
                idx[v] = nxt; nxt += 1
    for v, i in idx.items():
        assert g.var_index[v] == i
