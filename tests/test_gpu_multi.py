"""One process, one Allocate, several GPUs (tp_options.extra_devices): the reference policy is a single object
in a single process (manager.go:101, amdgpu.go:300-319), so the sharded search has to work behind
the same call.  Shards run on every visible GPU; on a one-GPU box the extra shards are put on GPU 0
as well, which exercises the same host path (worker threads, per-GPU policies, min over the shards'
packed winners)."""
import pytest

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

pytestmark = pytest.mark.gpu
tp = package()


def gpus():
    import torch
    return torch.cuda.device_count()


def multi_policy(devs, pairs, n_shards, min_subsets=1):
    extra = [s % gpus() for s in range(1, n_shards)]
    pol = tp.BestEffortPolicy(cuda_device=0, extra_devices=extra, multi_gpu_min_subsets=min_subsets)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return pol


@pytest.mark.parametrize("n_shards", [2, 3, 8])
def test_sharded_allocate_matches_reference(n_shards):
    devs, pairs = topo.fixture_topology("cpx")
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = multi_policy(devs, pairs, n_shards)
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x3617 + n_shards)
    sharded = 0
    for _ in range(40):
        av, rq, k = topo.random_request(gen, ids, kmax=8, p_choices=(0.3, 0.5))
        want = ref.Allocate(av, rq, k)
        got = pol.Allocate(av, rq, k)
        assert got == want, (av, rq, k)
        if ref.last["searched"]:
            assert pol.last["score"] == ref.last["score"]
            if pol.last["kernel"] != tp.KERNEL_FAMILY:
                assert pol.last["subsets_scored"] == pol.last["subsets_total"]
                sharded += pol.last["gpu_launches"] > 1
    assert sharded > 10
    pol.close()


def test_full_size_and_two_word_keys_over_all_gpus():
    """C3 at full size over max(2, visible GPUs) shards, and a winner whose group sequence needs both key words."""
    n = max(2, gpus())
    devs, pairs = topo.fixture_topology("cpx")
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = multi_policy(devs, pairs, n, min_subsets=0)  # default threshold (2^31): C3 is above it
    ids = [d.Id for d in devs]
    assert pol.Allocate(ids, [], 8) == ref.Allocate(ids, [], 8)
    assert pol.last["subsets_scored"] == 3872894697 and pol.last["gpu_launches"] == n and pol.last["score"] == 840
    assert pol.Allocate(ids[:40], [], 6) == ref.Allocate(ids[:40], [], 6)  # below the threshold: one GPU, one launch
    assert pol.last["gpu_launches"] == 1
    pol.close()
    devs, pairs = topo.wide_topology()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = multi_policy(devs, pairs, n)
    ids = [d.Id for d in devs]
    singles = [ids[g * 8 + (g % 3)] for g in range(16)]
    for av, rq, k in ((singles, [], 10), (singles[1:], [singles[4]], 10)):
        assert pol.Allocate(av, rq, k) == ref.AllocateExhaustive(av, rq, k)
        assert pol.last["key"][1] != 0 and pol.last["gpu_launches"] == n
    pol.close()


def test_bad_extra_device_fails_init():
    devs, pairs = topo.fixture_topology("mi210")
    pol = tp.BestEffortPolicy(cuda_device=0, extra_devices=[63])
    with pytest.raises(tp.AllocatorError, match="GPU policy init failed"):
        pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    pol.close()
