"""Shared test helpers: golden topologies, synthetic topologies, seeded requests.

Nothing here reads /root/reference: fixtures come from tests/golden/*.json
(written by tests/golden/make_golden.py in the build container).
"""
from __future__ import annotations

import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle.binding import Device, get_test_devices  # noqa: E402

GOLDEN = os.path.join(HERE, "golden")
_MASK = (1 << 64) - 1


class SplitMix64:
    """splitmix64, the RNG SURVEY.md section 8(d) names for the seeded workloads."""

    def __init__(self, seed: int):
        self.s = seed & _MASK

    def next(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & _MASK
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
        return z ^ (z >> 31)

    def below(self, n: int) -> int:
        return self.next() % n

    def chance(self, p: float) -> bool:
        return (self.next() >> 11) * (1.0 / (1 << 53)) < p


def random_request(gen: SplitMix64, ids, kmax=8, p_choices=(0.25, 0.5, 0.75), k_choices=None,
                   p_required=0.5, max_required=2):
    """(available, required, size): random availability mask, size in [1,kmax], 0-2 required."""
    while True:
        p = p_choices[gen.below(len(p_choices))]
        av = [i for i in ids if gen.chance(p)]
        k = k_choices[gen.below(len(k_choices))] if k_choices else 1 + gen.below(kmax)
        if len(av) <= k:
            continue
        # shuffle the request order: kubelet gives no ordering guarantee
        for i in range(len(av) - 1, 0, -1):
            j = gen.below(i + 1)
            av[i], av[j] = av[j], av[i]
        req = []
        if gen.chance(p_required):
            nreq = min(1 + gen.below(max_required), k - 1)
            pool = list(av)
            for _ in range(nreq):
                req.append(pool.pop(gen.below(len(pool))))
        return av, req, k


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def fixture_topology(name):
    """(devices, pairs) of a reference fixture, from tests/golden/topologies.json."""
    t = golden("topologies.json")[name]
    return get_test_devices(*t["test_devices"]), [tuple(p) for p in t["pairs"]]


def pair_weight(a: Device, b: Device, link_type: int) -> int:
    """calculatePairWeight (device.go:135-157)."""
    return (10 if a.DevId == b.DevId else 20) + (10 if link_type == 11 else 40 if link_type == 2 else 50) \
        + (10 if a.NumaNode == b.NumaNode else 20)


def synthetic_topology(groups, per_group, numa_count=2, start_node=2, link=lambda a, b: 11, pad=2,
                       drop=None):
    """groups x per_group partitions; ParentIds zero-padded so Go string order == numeric.

    link(a, b) -> KFD link type (11 XGMI, 2 PCIe, else other) or None for a missing link.
    drop(a, b) -> True to leave the pair out of p2pWeights (absent => weight 0).
    """
    devs = []
    node = start_node
    per_numa = max(groups // numa_count, 1)
    for g in range(groups):
        for j in range(per_group):
            ident = f"test{g + 1:0{pad}d}" if j == 0 else f"amdgpu_xcp_{g * per_group + j}"
            devs.append(Device(ident, node, g // per_numa, str(g)))
            node += 1
    pairs = []
    for i, a in enumerate(devs):
        for b in devs[i + 1:]:
            if drop and drop(a, b):
                continue
            ty = link(a, b)
            if ty is None:
                continue
            pairs.append((a.NodeId, b.NodeId, pair_weight(a, b, ty)))
    return devs, pairs


def c4_topology():
    """BASELINE config C4: 10 groups x 8, NodeIds 2..81, NUMA = g // 5, all XGMI."""
    return synthetic_topology(10, 8, numa_count=2)


def cpx64_topology():
    """CPX fixture plus a 64th synthetic partition (NodeId 65, group 8): the 8-of-64 search."""
    devs, pairs = fixture_topology("cpx")
    extra = Device("amdgpu_xcp_63", 65, devs[-1].NumaNode, devs[-1].DevId)
    pairs = list(pairs) + [(d.NodeId, 65, pair_weight(d, extra, 11)) for d in devs]
    return devs + [extra], pairs


def hives_topology():
    """8 groups x 4; XGMI inside hives of two GPUs, PCIe across; NUMA interleaved."""
    devs, _ = synthetic_topology(8, 4)
    for d in devs:
        d.NumaNode = int(d.DevId) % 2
    pairs = []
    for i, a in enumerate(devs):
        for b in devs[i + 1:]:
            ty = 11 if int(a.DevId) // 2 == int(b.DevId) // 2 else 2
            pairs.append((a.NodeId, b.NodeId, pair_weight(a, b, ty)))
    return devs, pairs


def missing_links_topology(seed=7, frac=0.15):
    """8 groups x 4 with ~15% of the pairs absent from p2pWeights (weight 0)."""
    gen = SplitMix64(seed)
    return synthetic_topology(8, 4, drop=lambda a, b: gen.chance(frac))


def wide_topology():
    """16 groups x 8 = 128 devices (the ABI's maximum), 4 NUMA nodes, XGMI inside pairs of groups, PCIe across."""
    devs, _ = synthetic_topology(16, 8, numa_count=4)
    pairs = []
    for i, a in enumerate(devs):
        for b in devs[i + 1:]:
            ty = 11 if int(a.DevId) // 2 == int(b.DevId) // 2 else 2
            pairs.append((a.NodeId, b.NodeId, pair_weight(a, b, ty)))
    return devs, pairs


def materialize_sysfs_tree(name, root):
    """Write fixture `name` of tests/golden/sysfs_trees.json.gz (see make_sysfs_trees.py) under `root` as the
    KFD topology layout the policy parses (device.go:107-252); returns the nodes directory."""
    import gzip
    with gzip.open(os.path.join(GOLDEN, "sysfs_trees.json.gz"), "rt") as f:
        nodes = json.load(f)[name]
    top = os.path.join(root, name, "nodes")
    for node, entry in nodes.items():
        d = os.path.join(top, node)
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "properties"), "w") as f:
            f.write("cpu_cores_count 0\n" + (f"drm_render_minor {entry['minor']}\n" if entry.get("minor") is not None else ""))
        for sub in ("io_links", "p2p_links"):
            for ln, (a, b, ty) in (entry.get(sub) or {}).items():
                ld = os.path.join(d, sub, ln)
                os.makedirs(ld, exist_ok=True)
                with open(os.path.join(ld, "properties"), "w") as f:
                    f.write("".join(f"{k} {v}\n" for k, v in (("type", ty), ("node_from", a), ("node_to", b)) if v is not None))
    return top
