"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle, bit-exact.

Same request -> same ordered device ids and same score as the reference's policy
(oracle/ref_policy.c), on the reference's own cases, the frozen golden vectors and
seeded random requests over fixture and synthetic topologies."""
import pytest

import topo
from __graft_entry__ import package
from oracle.binding import RefPolicy

pytestmark = pytest.mark.gpu
tp = package()

FIXTURES = ["mi308", "mi210", "cpx"]
SYNTH = {"c4": topo.c4_topology, "hives": topo.hives_topology, "missing": topo.missing_links_topology,
         "cpx64": topo.cpx64_topology}


def make_pair(name):
    devs, pairs = topo.fixture_topology(name) if name in FIXTURES else SYNTH[name]()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    return ref, pol


@pytest.fixture(scope="module")
def pols():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = make_pair(name)
        return cache[name]
    yield get
    for _, pol in cache.values():
        pol.close()


def check(ref, pol, av, rq, size):
    want = ref.Allocate(av, rq, size)
    got = pol.Allocate(av, rq, size)
    assert got == want, (av, rq, size, got, want)
    if ref.last["searched"]:
        assert pol.last["searched"] == 1 and pol.last["gpu_launches"] >= 1
        assert pol.last["score"] == ref.last["score"]
        if pol.last["kernel"] != tp.KERNEL_FAMILY:
            assert pol.last["subsets_scored"] == pol.last["subsets_total"]
    else:
        assert pol.last["searched"] == 0


@pytest.mark.parametrize("name", FIXTURES)
def test_reference_cases(pols, name):
    """besteffort_policy_test.go:53-160 replayed on the GPU policy."""
    ref, pol = pols(name)
    for case in topo.golden("reference_cases.json")[name]:
        out = pol.Allocate(case["available"], case.get("required") or [], case["size"])
        assert len(out) == case["size"]
        if case.get("expected") is not None:
            assert sorted(out) == sorted(case["expected"]), case["line"]
        assert out == case["oracle"]["ids"]
        assert pol.last["score"] == case["oracle"]["score"]


def test_survey_vectors(pols):
    for c in topo.golden("survey_cases.json"):
        ref, pol = pols(c["topology"])
        by_node = {d.NodeId: d.Id for d in ref.devices}
        out = pol.Allocate([by_node[n] for n in c["available_nodes"]], [by_node[n] for n in c["required_nodes"]],
                           c["size"])
        assert out == c["oracle_ids"] and pol.last["score"] == c["score"]


@pytest.mark.parametrize("name", FIXTURES)
def test_frozen_random_cases(pols, name):
    _, pol = pols(name)
    for c in topo.golden("random_cases.json")[name]:
        assert pol.Allocate(c["available"], c["required"], c["size"]) == c["ids"]
        assert pol.last["score"] == c["score"]


@pytest.mark.parametrize("name", FIXTURES + list(SYNTH))
def test_random_requests_vs_oracle(pols, name):
    ref, pol = pols(name)
    ids = [d.Id for d in ref.devices]
    gen = topo.SplitMix64(0xB200 ^ (len(name) << 8) ^ len(ids))
    for _ in range(300):
        av, rq, k = topo.random_request(gen, ids, kmax=7 if len(ids) > 40 else 8)
        check(ref, pol, av, rq, k)


def test_edge_cases(pols):
    ref, pol = pols("cpx")
    ids = [d.Id for d in ref.devices]
    check(ref, pol, ids[:9], [], 8)            # one more available than requested
    check(ref, pol, ids, ids[:7], 8)           # k' = 1
    check(ref, pol, ids[:3], [], 3)            # short-circuit: available == size, order as given
    check(ref, pol, list(reversed(ids[:5])), [ids[1], ids[0]], 2)  # short-circuit: required == size
    check(ref, pol, ids[:20], [ids[19], ids[3]], 6)
    for args, msg in [((ids, [], 0), "allocation size can not be negative"),
                      ((ids[:2], [], 3), "available devices count less than allocation size"),
                      ((ids, ids[:3], 2), "must_include devices size is more than allocation size"),
                      ((ids[:6], ["amdgpu_xcp_60"], 3), "No candidate subset found with matching criteria")]:
        with pytest.raises(tp.AllocatorError, match=msg):
            pol.Allocate(*args)
    with pytest.raises(tp.AllocatorError, match="unknown or duplicate device index"):
        pol.Allocate(ids[:6] + ["nope"], ["nope"], 3)  # the reference nil-derefs here


def test_full_size_cpx_8_of_63(pols):
    """BASELINE config C3 at full size: 3 872 894 697 subsets, expected test1 + xcp_1..7 (score 840)."""
    ref, pol = pols("cpx")
    ids = [d.Id for d in ref.devices]
    out = pol.Allocate(ids, [], 8)
    assert out == ["test1"] + [f"amdgpu_xcp_{i}" for i in range(1, 8)]
    assert pol.last["score"] == 840 and pol.last["subsets_scored"] == 3872894697


def test_shards_cover_the_space(pols):
    """Multi-GPU contract on one GPU: min over shard keys == the unsharded key, counts add up."""
    ref, pol = pols("cpx")
    ids = [d.Id for d in ref.devices]
    gen = topo.SplitMix64(77)
    for shards in (2, 3, 8):
        av, rq, k = topo.random_request(gen, ids, kmax=6)
        want = ref.Allocate(av, rq, k)
        if not ref.last["searched"]:
            continue
        keys = [pol.SearchShard(av, rq, k, s, shards) for s in range(shards)]
        assert pol.Decode(av, rq, k, min(keys)) == want


def test_full_size_headline_searches(pols):
    """BASELINE configs at full size against the reference policy's ids (ordered): C4 = 8 of 80
    (28 987 537 150 subsets), 8 of 64 (the 4.4e9 figure of C3) and 7 of 80, every subset counted."""
    for name, n_take, size, total in (("c4", 80, 8, 28987537150), ("cpx64", 64, 8, 4426165368), ("c4", 80, 7, 3176716400)):
        ref, pol = pols(name)
        ids = [d.Id for d in ref.devices][:n_take]
        want = ref.Allocate(ids, [], size)
        got = pol.Allocate(ids, [], size)
        assert got == want, (name, got, want)
        assert pol.last["score"] == ref.last["score"] and pol.last["subsets_scored"] == total
        assert pol.last["kernel"] == tp.KERNEL_TRI
    # ... and with required devices and a ragged pool on the 80-device topology
    ref, pol = pols("c4")
    ids = [d.Id for d in ref.devices]
    pool = [i for n, i in enumerate(ids) if n % 11 != 3]
    for rq, size in (([pool[40]], 8), ([pool[3], pool[70]], 9)):
        assert pol.Allocate(pool, rq, size) == ref.Allocate(pool, rq, size)
        assert pol.last["score"] == ref.last["score"] and pol.last["subsets_scored"] == pol.last["subsets_total"]


@pytest.mark.parametrize("name", ["cpx", "c4", "missing", "hives", "mi308"])
def test_tri_kernel_walk_depths_and_widths(pols, name):
    """Kernel TRI: k' = 6..12 (0..3 walked positions above the triple), both lane-vector widths, pools
    wide enough for leaves above the table cap (the pruned re-walk), required devices, and shards;
    the forced SPLIT kernel must return the same key."""
    ref, pol = pols(name)
    devs = ref.devices
    pairs = topo.fixture_topology(name)[1] if name in FIXTURES else SYNTH[name]()[1]
    spl = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_SPLIT)
    spl.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x7A1 ^ len(ids))
    seen = 0
    for kp, cap in ((6, 80), (7, 60), (8, 44), (9, 30), (10, 26), (11, 24), (12, 22)):
        for trial in range(3):
            n = min(len(ids), cap - gen.below(6))
            pool = sorted(ids, key=lambda _: gen.next())[:n] if trial else ids[:n]
            nreq = gen.below(3) if trial == 2 else 0
            rq, size = pool[:nreq], kp + nreq
            if len(pool) <= size + 1:
                continue
            want = ref.Allocate(pool, rq, size)
            got = pol.Allocate(pool, rq, size)
            assert got == want, (name, kp, len(pool), rq)
            assert pol.last["score"] == ref.last["score"] and pol.last["subsets_scored"] == pol.last["subsets_total"]
            if pol.last["subsets_total"] > 32768:
                assert pol.last["kernel"] == tp.KERNEL_TRI
                seen += 1
            assert spl.Allocate(pool, rq, size) == want and spl.last["key"] == pol.last["key"]
            keys = [pol.SearchShard(pool, rq, size, s, 3) for s in range(3)]
            assert pol.Decode(pool, rq, size, min(keys)) == want
    assert seen >= 6
    spl.close()


@pytest.mark.parametrize("name,sizes", [("mi308", range(4, 13)), ("cpx", (4, 5, 6, 7)), ("c4", (4, 5, 6)),
                                        ("missing", range(4, 11)), ("hives", range(4, 12))])
def test_split_kernel_all_depths(pols, name, sizes):
    """Kernels SPLIT (k' >= 4, every walk depth D = 0..4) and TRI (k' >= 6, every walk depth 0..3): LEX,
    SPLIT (forced) and the default choice (TRI from k' = 6 on) must agree with the oracle and with
    each other, key for key."""
    ref, pol = pols(name)
    devs = ref.devices
    pairs = topo.fixture_topology(name)[1] if name in FIXTURES else SYNTH[name]()[1]
    lex = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_LEX)
    lex.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    spl = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_SPLIT)
    spl.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x5B117 + len(ids))
    seen = set()
    for size in sizes:
        for trial in range(4):
            cap = 26 if size >= 9 else 40 if size >= 7 else len(ids)  # keep C(n,k') small enough for seconds
            pool = ids[:] if trial == 0 and cap == len(ids) else [i for i in ids if gen.chance(0.6)][:cap]
            if len(pool) <= size + 1:
                pool = ids[:size + 3]
            nreq = gen.below(3) if size > 5 else 0
            req = pool[:nreq]
            want = ref.Allocate(pool, req, size)
            got = pol.Allocate(pool, req, size)
            assert got == want, (name, size, pool, req)
            assert pol.last["score"] == ref.last["score"]
            assert pol.last["subsets_scored"] == pol.last["subsets_total"]
            seen.add(pol.last["kernel"])
            assert lex.Allocate(pool, req, size) == want and lex.last["kernel"] == tp.KERNEL_LEX
            assert lex.last["key"] == pol.last["key"]
            assert spl.Allocate(pool, req, size) == want and spl.last["key"] == pol.last["key"]
            assert spl.last["kernel"] in (tp.KERNEL_SPLIT, tp.KERNEL_LEX)
            seen.add(spl.last["kernel"])
            if size - nreq >= 6 and pol.last["subsets_total"] > 32768:
                assert pol.last["kernel"] == tp.KERNEL_TRI, (name, size, len(pool))
    assert tp.KERNEL_SPLIT in seen
    assert tp.KERNEL_TRI in seen or max(sizes) < 6
    lex.close()
    spl.close()


@pytest.mark.parametrize("name", ["mi308", "cpx", "c4", "missing", "hives"])
def test_family_kernel(pols, name):
    """Kernel FAMILY: (a) forced on enumerable requests it must equal the exhaustive kernels' key;
    (b) on requests far beyond enumeration (the reference's own 30-of-63 case and larger) it must
    equal the reference's BFS."""
    ref, pol = pols(name)
    devs = ref.devices
    fam = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_FAMILY)
    fam.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs],
                        topo.fixture_topology(name)[1] if name in FIXTURES else SYNTH[name]()[1])
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0xFA111 + len(ids))
    for _ in range(60):  # (a)
        av, rq, k = topo.random_request(gen, ids, kmax=6)
        want = ref.Allocate(av, rq, k)
        if not ref.last["searched"]:
            continue
        assert fam.Allocate(av, rq, k) == want and fam.last["kernel"] == tp.KERNEL_FAMILY
        assert pol.Allocate(av, rq, k) == want and pol.last["key"] == fam.last["key"]
    big = 0
    for _ in range(40):  # (b)
        av = [i for i in ids if gen.chance(0.85)]
        k = 17 + gen.below(max(min(len(av) - 18, 24), 1))
        if len(av) <= k:
            continue
        rq = av[:gen.below(3)]
        want = ref.Allocate(av, rq, k)
        got = pol.Allocate(av, rq, k)
        assert got == want, (name, len(av), k)
        assert pol.last["score"] == ref.last["score"] and pol.last["subsets_scored"] > 0
        if k - len(rq) > tp.TP_MAX_SUBSET:
            assert pol.last["kernel"] == tp.KERNEL_FAMILY
            big += 1
    assert big > 10
    fam.close()


def test_widest_topology_and_two_word_keys():
    """128 devices (TP_MAX_DEVICES): the M = 128 instantiations of SPLIT, LEX at full width, and
    winners whose group sequence spills into key.lo (16 groups = 4-bit digits, k' >= 9)."""
    devs, pairs = topo.wide_topology()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x128)
    kernels, wide_keys = set(), 0
    cases = [(ids, [], 4), (ids, [ids[5]], 5), (ids[3:], [], 5), (ids, [], 2), (ids, [], 3)]
    # one free device in each of the 16 groups, k' >= 9: the winner touches >= 9 groups, so its 4-bit
    # digits spill into key.lo.  (The reference's BFS would walk 16!/6! orderings here; the oracle's
    # exhaustive scorer, shown equal to the BFS in test_oracle_golden.py, answers instead.)
    singles = [ids[g * 8 + (g % 3)] for g in range(16)]
    wide_cases = [(singles, [], 10), (singles, [singles[4]], 11), (singles[1:], [], 9)]
    for _ in range(4):   # wide pools, k' = 4..5: n - p > 96 -> M = 128
        pool = [i for i in ids if gen.chance(0.9)]
        cases.append((pool, [], 4 + gen.below(2)))
    # deeper walks at the wide lane vectors: (n free, k') -> (M, D)
    cases.append((ids[:100], [ids[7]], 8))   # 99 free, k' = 7: M = 96,  D = 1
    cases.append((ids, [], 7))               # 128 free, k' = 7: M = 128, D = 1
    cases.append((ids[:96], [], 8))          # 96 free,  k' = 8: M = 96,  D = 2
    cases.append((ids[:84], [], 8))          # 84 free,  k' = 8: M = 80,  D = 2
    for av, rq, k in cases + wide_cases:
        exhaustive = (av, rq, k) in wide_cases
        want = ref.AllocateExhaustive(av, rq, k) if exhaustive else ref.Allocate(av, rq, k)
        got = pol.Allocate(av, rq, k)
        assert got == want, (len(av), rq, k)
        assert pol.last["score"] == ref.last["score"]
        kernels.add(pol.last["kernel"])
        wide_keys += pol.last["key"][1] != 0
        # shards of the same request must min-reduce (two words) to the same winner
        keys = [pol.SearchShard(av, rq, k, s, 3) for s in range(3)]
        assert pol.Decode(av, rq, k, min(keys)) == want
    assert tp.KERNEL_SPLIT in kernels and tp.KERNEL_TRI in kernels and wide_keys > 0
    # the same deep walks through SPLIT's wide instantiations (the default picks TRI up to 80 + p devices)
    spl = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_SPLIT)
    spl.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    for av, rq, k in cases[-4:] + wide_cases:
        want = pol.Allocate(av, rq, k)
        assert spl.Allocate(av, rq, k) == want and spl.last["kernel"] == tp.KERNEL_SPLIT
        assert spl.last["key"] == pol.last["key"]
    spl.close()
    pol.close()


@pytest.mark.parametrize("kernel", [tp.KERNEL_SPLIT, tp.KERNEL_TRI])
@pytest.mark.parametrize("name", ["cpx", "missing", "hives"])
def test_work_plan_with_cut_values(name, kernel):
    """The work plan's second level: single values of the first walked position cut by the second one
    (normally only for the heaviest values of large or sharded searches) is forced for every value by
    the tuning knobs, at walk depths 2, 3 and 4; answers, scored counts and the min over shards must
    not change."""
    devs, pairs = topo.fixture_topology(name) if name in FIXTURES else SYNTH[name]()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    pol = tp.BestEffortPolicy(cuda_device=0, kernel=tp.KERNEL_SPLIT if kernel == tp.KERNEL_SPLIT else tp.KERNEL_AUTO)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    ids = [d.Id for d in ref.devices]
    gen = topo.SplitMix64(0xC07 + len(ids))
    pol.set_tuning(item_cap=200, item_cap1=150)
    for size, cap in ((8, 30), (9, 26), (10, 24)):  # k' = size: D = 2, 3, 4
        for _ in range(3):
            pool = [i for i in ids if gen.chance(0.7)][:cap]
            if len(pool) <= size + 2:
                pool = ids[:size + 4]
            want = ref.Allocate(pool, [], size)
            got = pol.Allocate(pool, [], size)
            assert got == want, (name, size, pool)
            assert pol.last["score"] == ref.last["score"]
            assert pol.last["kernel"] == (kernel if pol.last["subsets_total"] > 32768 or kernel == tp.KERNEL_SPLIT
                                          else tp.KERNEL_LEX)
            assert pol.last["subsets_scored"] == pol.last["subsets_total"]
            keys = [pol.SearchShard(pool, [], size, s, 5) for s in range(5)]
            assert pol.Decode(pool, [], size, min(keys)) == want
    pol.close()
