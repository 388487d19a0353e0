"""The advisory mode's checker (oracle/exhaustive.c: ref_exhaustive_global_optimum) on CPU.

Pins the unconstrained optimum against a brute force written here (itertools over the id
strings, weights straight from the pair list), and records the two facts SURVEY.md section
8(f) rank 4 states: on fully linked topologies the reference's pick IS a global optimum; with
links missing from p2pWeights it often is not.
"""
import itertools

import pytest

import topo
from oracle.binding import RefPolicy

FIXTURES = ["mi308", "mi210", "cpx"]


def make(name):
    devs, pairs = topo.fixture_topology(name) if name in FIXTURES else getattr(topo, name + "_topology")()
    pol = RefPolicy()
    pol.InitPairs(devs, pairs)
    node = {d.Id: d.NodeId for d in devs}
    w = {(min(a, b), max(a, b)): v for a, b, v in pairs}
    return pol, node, w


def score(ids, node, w):
    """DeviceSet.TotalWeight (device.go:254-273): every pair once, absent pair = 0."""
    ns = [node[i] for i in ids]
    return sum(w.get((min(a, b), max(a, b)), 0) for a, b in itertools.combinations(ns, 2))


def brute(av, rq, size, node, w):
    free = [i for i in av if i not in rq]
    return min(score(list(c) + rq, node, w) for c in itertools.combinations(free, size - len(rq)))


@pytest.mark.parametrize("name", ["mi308", "mi210", "missing_links", "hives"])
def test_against_brute_force(name):
    pol, node, w = make(name)
    ids = list(node)
    gen = topo.SplitMix64(0x60 + len(ids))
    done = 0
    while done < 60:
        av, rq, k = topo.random_request(gen, ids, kmax=5)
        if len(av) > 18:
            av = av[:18]
            if any(r not in av for r in rq) or len(av) <= k:
                continue
        out = pol.GlobalOptimum(av, rq, k)
        assert len(out) == k and len(set(out)) == k and set(rq) <= set(out) <= set(av)
        if not pol.last["total"]:  # short-circuit: nothing searched
            continue
        assert out[k - len(rq):] == rq  # required last, request order
        assert pol.last["scored"] == pol.last["total"]
        assert pol.last["score"] == score(out, node, w) == brute(av, rq, k, node, w)
        done += 1


@pytest.mark.parametrize("name", FIXTURES + ["c4", "hives"])
def test_reference_pick_is_optimal_on_fully_linked_topologies(name):
    pol, node, _ = make(name)
    ids = list(node)
    gen = topo.SplitMix64(0xAD + len(ids))
    for _ in range(40):
        av, rq, k = topo.random_request(gen, ids, kmax=5 if len(ids) > 40 else 7)
        pol.Allocate(av, rq, k)
        if not pol.last["searched"]:
            continue
        ref_score = pol.last["score"]
        pol.GlobalOptimum(av, rq, k)
        assert pol.last["score"] == ref_score


def test_reference_pick_can_miss_the_optimum_when_links_are_missing():
    pol, node, _ = make("missing_links")
    ids = list(node)
    gen = topo.SplitMix64(0x155)
    searched = beaten = 0
    for _ in range(200):
        av, rq, k = topo.random_request(gen, ids, kmax=6)
        pol.Allocate(av, rq, k)
        if not pol.last["searched"]:
            continue
        ref_score = pol.last["score"]
        pol.GlobalOptimum(av, rq, k)
        assert pol.last["score"] <= ref_score
        searched += 1
        beaten += pol.last["score"] < ref_score
    assert searched > 100 and beaten > searched // 10, (searched, beaten)
