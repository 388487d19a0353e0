"""CPU-side checks of the boundary: the library loads, exports every symbol the header
declares, and the host logic that needs no GPU (validation order, short-circuits, error
strings, sysfs parsing of the weight map) behaves like the reference."""
import os
import re

import pytest

import topo
from __graft_entry__ import ROOT, package

tp = package()


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "tp_allocator.h")).read()
    declared = set(re.findall(r"\b(tp_[a-z_]+)\s*\(", hdr))
    assert {"tp_create", "tp_allocate", "tp_allocate_batch", "tp_search_shard", "tp_decode", "tp_destroy",
            "tp_strerror"} <= declared
    L = tp.lib()
    for sym in declared:
        assert hasattr(L, sym), sym


def test_error_strings_match_reference():
    # besteffort_policy.go:36-43
    assert tp.strerror(-1) == "allocation size can not be negative"
    assert tp.strerror(-2) == "available devices count less than allocation size"
    assert tp.strerror(-3) == "must_include devices size is more than allocation size"
    assert tp.strerror(-4) == "must_include length should be less than or equal to avilable device size"
    assert tp.strerror(-5) == "Init method must be called before Allocate"
    assert tp.strerror(-6) == "No candidate subset found with matching criteria"


def test_uninitialised_policy_validation_order():
    pol = tp.BestEffortPolicy()
    ids = [f"test{i}" for i in range(1, 9)]
    for args, msg in [((ids, [], 0), "allocation size can not be negative"),
                      ((ids, [], -3), "allocation size can not be negative"),
                      ((ids[:2], [], 3), "available devices count less than allocation size"),
                      ((ids, ids[:3], 2), "must_include devices size is more than allocation size"),
                      ((ids, [], 2), "Init method must be called before Allocate")]:
        with pytest.raises(tp.AllocatorError, match=msg):
            pol.Allocate(*args)


def test_init_without_devices_fails_like_reference():
    pol = tp.BestEffortPolicy()
    with pytest.raises(tp.AllocatorError, match="Besteffort Policy init failed to initialize p2pWeights"):
        pol.Init([], "/nonexistent")


@pytest.mark.skipif(not os.path.isdir("/root/reference/testdata"), reason="reference trees only exist in the build container")
@pytest.mark.parametrize("name,tree", [("mi308", "topology-parsing-mi308/topology/nodes"),
                                       ("mi210", "topo-mi210-xgmi-pcie/nodes"),
                                       ("cpx", "topo-mi300-cpx/topology/nodes")])
def test_host_sysfs_reader_matches_golden(name, tree):
    """The C++ reader (host/besteffort_policy.cpp, device.go:107-252) builds the same map."""
    devs, pairs = topo.fixture_topology(name)
    pol = tp.BestEffortPolicy()
    try:
        pol.Init([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], "/root/reference/testdata/" + tree)
    except tp.AllocatorError as e:  # no GPU here: tp_create fails after the parse
        assert "GPU policy init failed" in str(e)
    assert pol.pairs() == sorted(pairs)
    assert pol.weight_rows() == topo.golden("topologies.json")[name]["weight_rows"]


@pytest.mark.parametrize("name", ["mi308", "mi210", "cpx"])
def test_host_sysfs_reader_on_materialised_tree(name, tmp_path):
    """Same check from the distilled trees of tests/golden/sysfs_trees.json.gz (what the GPU box parses)."""
    devs, pairs = topo.fixture_topology(name)
    pol = tp.BestEffortPolicy()
    try:
        pol.Init([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs],
                 topo.materialize_sysfs_tree(name, str(tmp_path)))
    except tp.AllocatorError as e:  # no GPU here: tp_create fails after the parse
        assert "GPU policy init failed" in str(e)
    assert pol.pairs() == sorted(pairs)
