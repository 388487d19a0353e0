"""GPU parity of the advisory mode (tp_global_optimum) against oracle/exhaustive.c's
ref_exhaustive_global_optimum: same ordered ids, same score, every subset scored, on the fixture and
synthetic topologies, through both exhaustive kernels; plus the full-size BASELINE searches, where
the unconstrained optimum must equal the reference's pick (fully linked topologies)."""
import pytest

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

pytestmark = pytest.mark.gpu
tp = package()

FIXTURES = ["mi308", "mi210", "cpx"]
SYNTH = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology}


def make(name, kernel=None):
    devs, pairs = topo.fixture_topology(name) if name in FIXTURES else SYNTH[name]()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_AUTO if kernel is None else kernel)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return ref, pol


@pytest.mark.parametrize("name", FIXTURES + list(SYNTH))
def test_random_requests(name):
    ref, pol = make(name)
    _, lex = make(name, tp.KERNEL_LEX)
    ids = [d.Id for d in ref.devices]
    gen = topo.SplitMix64(0x6107 ^ len(ids))
    kernels, beaten = set(), 0
    for _ in range(120):
        av, rq, k = topo.random_request(gen, ids, kmax=6 if len(ids) > 40 else 8)
        want = ref.GlobalOptimum(av, rq, k)
        got = pol.GlobalOptimum(av, rq, k)
        assert got == want, (av, rq, k)
        if not ref.last["total"]:  # short-circuit
            assert pol.last["searched"] == 0
            continue
        assert pol.last["score"] == ref.last["score"]
        assert pol.last["subsets_scored"] == pol.last["subsets_total"] == ref.last["total"]
        assert pol.last["groups_touched"] == k - len(rq)  # every device its own group
        kernels.add(pol.last["kernel"])
        assert lex.GlobalOptimum(av, rq, k) == want and lex.last["key"] == pol.last["key"]
        adv = pol.last["score"]
        pol.Allocate(av, rq, k)  # the reference's pick never beats it; ties on fully linked topologies
        assert adv <= pol.last["score"]
        if name != "missing":
            assert adv == pol.last["score"]
        beaten += adv < pol.last["score"]
    assert kernels & {tp.KERNEL_SPLIT, tp.KERNEL_TRI} or len(ids) <= 8
    if name == "missing":
        assert beaten > 5
    pol.close()
    lex.close()


def test_validation_and_limits():
    ref, pol = make("cpx")
    ids = [d.Id for d in ref.devices]
    assert pol.GlobalOptimum(ids[:3], [], 3) == ids[:3] and pol.last["searched"] == 0
    for args, msg in [((ids, [], 0), "allocation size can not be negative"),
                      ((ids[:2], [], 3), "available devices count less than allocation size"),
                      ((ids, ids[:3], 2), "must_include devices size is more than allocation size"),
                      ((ids[:6], ["amdgpu_xcp_60"], 3), "No candidate subset found with matching criteria"),
                      ((ids, [], 20), "search space too large")]:  # k' = 20 > TP_MAX_SUBSET
        with pytest.raises(tp.AllocatorError, match=msg):
            pol.GlobalOptimum(*args)
    pol.close()


def test_full_size_searches():
    """C3 (8 of 63 CPX partitions) and C4 (8 of 80) without the Family restriction: 7-bit digits, so
    the winner's sequence spills into key.lo; the optimum equals the reference's pick."""
    for name, want_total in (("cpx", 3872894697), ("c4", 28987537150)):
        ref, pol = make(name)
        ids = [d.Id for d in ref.devices]
        want = ref.GlobalOptimum(ids, [], 8)
        got = pol.GlobalOptimum(ids, [], 8)
        assert got == want and pol.last["score"] == ref.last["score"]
        assert pol.last["subsets_scored"] == want_total == ref.last["total"]
        assert pol.last["kernel"] == tp.KERNEL_TRI and pol.last["key"][1] != 0
        adv = pol.last["score"]
        pol.Allocate(ids, [], 8)
        assert adv == pol.last["score"]
        pol.close()
