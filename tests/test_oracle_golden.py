"""Pins the CPU oracle (oracle/ref_policy.c) to the reference's own known answers.

Every case of the reference's allocator tests is replayed from
tests/golden/reference_cases.json (besteffort_policy_test.go:53-160,
device_test.go:90-169), on weights frozen from the reference's fixture trees.
Also proves, on CPU, the claim the GPU path rests on: scoring every subset
under the Family constraint with the RefKey tie-break (oracle/exhaustive.c)
returns exactly what the reference's BFS returns.
"""
import pytest

import topo
from oracle.binding import OracleError, RefPolicy

FIXTURES = ["mi308", "mi210", "cpx"]


def make_policy(devs, pairs):
    pol = RefPolicy()
    pol.InitPairs(devs, pairs)
    return pol


@pytest.fixture(scope="module")
def policies():
    return {n: make_policy(*topo.fixture_topology(n)) for n in FIXTURES}


def test_weight_map_shape(policies):
    t = topo.golden("topologies.json")
    # device_test.go:90-108: 31 "from" rows on mi308; :110-123: 4 groups
    assert policies["mi308"].weight_rows() == 31 == t["mi308"]["weight_rows"]
    assert policies["mi308"].group_count() == 4
    assert len(policies["mi210"].pairs()) == 28 and {w for _, _, w in policies["mi210"].pairs()} == {40, 80}
    assert len(policies["cpx"].pairs()) == 1953 and {w for _, _, w in policies["cpx"].pairs()} == {30, 40, 50}


@pytest.mark.parametrize("name", FIXTURES)
def test_reference_known_answers(policies, name):
    pol = policies[name]
    for case in topo.golden("reference_cases.json")[name]:
        out = pol.Allocate(case["available"], case.get("required") or [], case["size"])
        assert len(out) == case["size"]
        if case.get("expected") is not None:  # besteffort_policy_test.go:196-209 compares sorted
            assert sorted(out) == sorted(case["expected"]), case["line"]
        assert out == case["oracle"]["ids"]
        assert pol.last["score"] == case["oracle"]["score"]
        assert pol.last["n_candidates"] == case["oracle"]["n_candidates"]


def test_candidate_counts(policies):
    # device_test.go:125-169: size 3 -> 4 candidates, size 12 -> 12 candidates
    pol = policies["mi308"]
    ids = [d.Id for d in pol.devices]
    pol.Allocate(ids, [], 3)
    assert pol.last["n_candidates"] == 4
    pol.Allocate(ids, [], 12)
    assert pol.last["n_candidates"] == 12


def test_survey_vectors(policies):
    for c in topo.golden("survey_cases.json"):
        pol = policies[c["topology"]]
        by_node = {d.NodeId: d.Id for d in pol.devices}
        node_of = {d.Id: d.NodeId for d in pol.devices}
        out = pol.Allocate([by_node[n] for n in c["available_nodes"]], [by_node[n] for n in c["required_nodes"]],
                           c["size"])
        assert sorted(node_of[i] for i in out) == sorted(c["expected_nodes"])
        assert pol.last["score"] == c["score"] and pol.last["n_candidates"] == c["n_candidates"]
        assert out == c["oracle_ids"]


@pytest.mark.parametrize("name", FIXTURES)
def test_frozen_random_cases(policies, name):
    pol = policies[name]
    for c in topo.golden("random_cases.json")[name]:
        assert pol.Allocate(c["available"], c["required"], c["size"]) == c["ids"]
        assert pol.last["score"] == c["score"]


def test_validation_order_and_short_circuits(policies):
    pol = policies["mi210"]
    ids = [d.Id for d in pol.devices]
    checks = [  # besteffort_policy.go:90-124, in order
        ((ids, [], 0), "allocation size can not be negative"),
        ((ids[:2], [], 3), "available devices count less than allocation size"),
        ((ids, ids[:3], 2), "must_include devices size is more than allocation size"),
        ((ids[:4], ["zz"], 3), "No candidate subset found with matching criteria"),
    ]
    for args, msg in checks:
        with pytest.raises(OracleError, match=msg):
            pol.Allocate(*args)
    # the fourth check (len(required) > len(available), :102-104) is unreachable: the first three
    # already imply len(required) <= size <= len(available)
    # len(avail) == size returns avail as given (no sort, no search)
    assert pol.Allocate(["test3", "test1"], [], 2) == ["test3", "test1"]
    assert not pol.last["searched"]
    assert pol.Allocate(ids, ["test7", "test2"], 2) == ["test7", "test2"]
    with pytest.raises(OracleError, match="Init method must be called"):
        RefPolicy().Allocate(ids, [], 2)


SYNTH = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology,
         "cpx64": topo.cpx64_topology}


@pytest.mark.parametrize("name", FIXTURES + list(SYNTH))
def test_exhaustive_family_refkey_equals_bfs(policies, name):
    """exhaustive + Family + RefKey == the reference BFS."""
    pol = policies[name] if name in policies else make_policy(*SYNTH[name]())
    ids = [d.Id for d in pol.devices]
    gen = topo.SplitMix64(0xA110C ^ len(name))
    n = 0
    while n < 120:
        av, req, k = topo.random_request(gen, ids, kmax=6 if len(ids) > 40 else 8)
        out = pol.Allocate(av, req, k)
        if not pol.last["searched"]:
            continue
        score = pol.last["score"]
        assert pol.AllocateExhaustive(av, req, k, threads=2) == out, (av, req, k)
        assert pol.last["score"] == score
        n += 1
