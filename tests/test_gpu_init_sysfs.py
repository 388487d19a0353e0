"""The reference's own test flow on the GPU box, sysfs tree to ids: BestEffortPolicy.Init(devs, topoDir)
parses a KFD topology tree (host C++ reader, device.go:107-252), hands the weights to the GPU
(tp_create), and Allocate answers the reference's known-answer cases
(besteffort_policy_test.go:162-216 drives exactly this: Init from a fixture folder, then Allocate).
The trees are materialised from tests/golden/sysfs_trees.json.gz (the fields the policy reads from
the reference's fixtures; /root/reference does not exist on the GPU box)."""
import pytest

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

pytestmark = pytest.mark.gpu
tp = package()


@pytest.mark.parametrize("name", ["mi308", "mi210", "cpx"])
def test_init_from_sysfs_then_allocate(name, tmp_path):
    tree = topo.materialize_sysfs_tree(name, str(tmp_path))
    devs, pairs = topo.fixture_topology(name)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.Init([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], tree)  # sysfs -> weights -> GPU
    assert pol.pairs() == sorted(pairs)
    assert pol.weight_rows() == topo.golden("topologies.json")[name]["weight_rows"]
    ref = RefPolicy()
    ref.Init(devs, tree)  # the oracle from the same tree
    for case in topo.golden("reference_cases.json")[name]:
        out = pol.Allocate(case["available"], case.get("required") or [], case["size"])
        assert len(out) == case["size"]
        if case.get("expected") is not None:  # the reference compares as sets (besteffort_policy_test.go:196-209)
            assert sorted(out) == sorted(case["expected"]), case["line"]
        assert out == case["oracle"]["ids"] and pol.last["score"] == case["oracle"]["score"]
        assert out == ref.Allocate(case["available"], case.get("required") or [], case["size"])
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x1217 + len(ids))
    for _ in range(40):
        av, rq, k = topo.random_request(gen, ids, kmax=6)
        assert pol.Allocate(av, rq, k) == ref.Allocate(av, rq, k)
    pol.close()


def test_init_from_empty_tree_disables_the_policy(tmp_path):
    """Init error path (besteffort_policy.go:72-74): no weights -> error, Allocate then says invalidInit."""
    pol = tp.BestEffortPolicy(cuda_device=0)
    devs, _ = topo.fixture_topology("mi210")
    with pytest.raises(tp.AllocatorError, match="Besteffort Policy init failed to initialize p2pWeights"):
        pol.Init([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], str(tmp_path))
    with pytest.raises(tp.AllocatorError, match="Init method must be called before Allocate"):
        pol.Allocate([d.Id for d in devs], [], 2)
    pol.close()
