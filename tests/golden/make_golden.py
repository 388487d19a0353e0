#!/usr/bin/env python
"""Generate tests/golden/*.json from the reference's fixture trees.

Run HERE (build container, /root/reference mounted); the GPU box has no
/root/reference, so what the tests need from the trees is frozen into small
JSON files:

  topologies.json       per fixture: the getTestDevices() parameters of the
                        reference test (besteffort_policy_test.go:26-51) and the
                        p2pWeights map the policy builds from the tree
                        (device.go:220-252) as (from, to, weight) triples.
  reference_cases.json  the reference's own known-answer cases
                        (besteffort_policy_test.go:53-160, device_test.go:90-169)
                        + what the oracle returns for them (order, score, #cand).
  survey_cases.json     the extra vectors of SURVEY.md section 8(c).
  random_cases.json     seeded random requests answered by the oracle's BFS
                        restatement (frozen regression data).

The sysfs parse is done twice -- by oracle/ref_policy.c and by an independent
Python reader using the reference's regular expressions verbatim -- and the two
must agree before anything is written.
"""
import json
import os
import re
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.binding import RefPolicy, get_test_devices  # noqa: E402
import topo  # noqa: E402

REF = "/root/reference/testdata/"
FIXTURES = {
    # name: (devCount, partitionCountPerDev, numanodeCount, startNodeId, endNodeId, tree)
    "mi308": (4, 8, 2, 2, 33, "topology-parsing-mi308/topology/nodes"),   # besteffort_policy_test.go:27-34
    "mi210": (8, 1, 2, 2, 9, "topo-mi210-xgmi-pcie/nodes"),               # :35-42
    "cpx": (8, 8, 2, 2, 64, "topo-mi300-cpx/topology/nodes"),             # :43-50
}


def go_parse_int(s):
    """strconv.ParseInt(s, 0, 32) on a digit string."""
    v = int(s, 8) if len(s) > 1 and s[0] == "0" else int(s)
    if v > 2**31 - 1:
        raise ValueError(s)
    return v


def last_match(path, regexes):
    res = [0] * len(regexes)
    with open(path) as f:
        for line in f.read().split("\n"):
            for i, r in enumerate(regexes):
                m = r.search(line)
                if m:
                    res[i] = go_parse_int(m.group(1))
    return res


def python_pairs(devs, tree):
    """Independent restatement of device.go:159-252 for cross-checking the C parser."""
    by_node = {d.NodeId: d for d in devs}
    weights = {}
    minor = [re.compile(r"drm_render_minor\s(\d+)")]
    link = [re.compile(r"node_from\s(\d+)"), re.compile(r"node_to\s(\d+)"), re.compile(r"type\s(\d+)")]
    for node in sorted(n for n in os.listdir(tree) if n[0].isdigit()):
        prop = os.path.join(tree, node, "properties")
        try:
            if last_match(prop, minor)[0] <= 0:
                continue
        except (OSError, ValueError):
            continue
        for sub in ("io_links", "p2p_links"):
            d = os.path.join(tree, node, sub)
            if not os.path.isdir(d):
                continue
            for ln in sorted(n for n in os.listdir(d) if n[0].isdigit()):
                try:
                    f, t, ty = last_match(os.path.join(d, ln, "properties"), link)
                except (OSError, ValueError):
                    continue
                a, b = (f, t) if f < t else (t, f)
                if a not in by_node or b not in by_node:
                    continue
                da, db = by_node[a], by_node[b]
                w = (10 if da.DevId == db.DevId else 20) + (10 if ty == 11 else 40 if ty == 2 else 50) \
                    + (10 if da.NumaNode == db.NumaNode else 20)
                weights[(a, b)] = w
    return sorted((a, b, w) for (a, b), w in weights.items())


def main():
    topologies = {}
    policies = {}
    for name, (dc, pc, nc, s, e, tree) in FIXTURES.items():
        devs = get_test_devices(dc, pc, nc, s, e)
        pol = RefPolicy()
        pol.Init(devs, REF + tree)
        pairs = pol.pairs()
        assert pairs == python_pairs(devs, REF + tree), name
        topologies[name] = {"test_devices": [dc, pc, nc, s, e], "source": "testdata/" + tree,
                            "weight_rows": pol.weight_rows(), "groups": pol.group_count(),
                            "pairs": [list(p) for p in pairs]}
        policies[name] = pol
        print(name, len(devs), "devices", len(pairs), "pairs")
    with open(os.path.join(HERE, "topologies.json"), "w") as f:
        json.dump(topologies, f, separators=(",", ":"))

    def answer(name, case):
        pol = policies[name]
        ids = [d.Id for d in pol.devices]
        av = case.get("available") or ids
        if case.get("filtered"):
            av = [a for a in av if a not in case["filtered"]]
        out = pol.Allocate(av, case.get("required") or [], case["size"])
        exp = case.get("expected")
        if exp is not None:
            assert sorted(out) == sorted(exp), (name, case, out)
        assert len(out) == case["size"]
        return dict(case, available=av, oracle={"ids": out, **pol.last})

    xcp = lambda a, b: [f"amdgpu_xcp_{i}" for i in range(a, b + 1)]  # noqa: E731
    t6 = ["test3", "test4", "test5", "test6", "test7", "test8"]
    ref_cases = {
        "mi308": [  # besteffort_policy_test.go:54-70 (size only)
            {"line": "55-59", "size": 1}, {"line": "60-64", "size": 3}, {"line": "65-69", "size": 12}],
        "mi210": [  # :71-97
            {"line": "72-77", "size": 1, "expected": ["test1"]},
            {"line": "78-83", "size": 3, "expected": ["test1", "test2", "test3"]},
            {"line": "84-89", "size": 5, "expected": ["test1", "test2", "test3", "test4", "test5"]},
            {"line": "90-96", "size": 3, "available": t6, "expected": ["test5", "test6", "test7"]}],
        "cpx": [  # :98-160
            {"line": "99-104", "size": 1, "expected": ["test8"]},
            {"line": "105-110", "size": 3, "expected": ["test8"] + xcp(57, 58)},
            {"line": "111-116", "size": 5, "expected": ["test8"] + xcp(57, 60)},
            {"line": "117-123", "size": 3, "available": t6, "expected": ["test5", "test6", "test7"]},
            {"line": "124-130", "size": 3, "available": t6, "required": ["test5"],
             "expected": ["test5", "test6", "test7"]},
            {"line": "131-134", "size": 30},
            {"line": "135-139", "size": 8, "expected": ["test1"] + xcp(1, 7)},
            {"line": "140-145", "size": 7, "expected": ["test8"] + xcp(57, 62)},
            {"line": "146-152", "size": 4, "filtered": ["test8"] + xcp(57, 58), "expected": xcp(59, 62)},
            {"line": "153-159", "size": 10, "filtered": ["test1", "test2", "test3", "test4", "test8", "amdgpu_xcp_57"],
             "expected": ["test5"] + xcp(33, 39) + xcp(58, 59)}],
    }
    out = {name: [answer(name, c) for c in cases] for name, cases in ref_cases.items()}
    # device_test.go:125-169 pins the candidate counts 4 and 12 on mi308
    assert out["mi308"][1]["oracle"]["n_candidates"] == 4 and out["mi308"][2]["oracle"]["n_candidates"] == 12
    assert topologies["mi308"]["weight_rows"] == 31 and topologies["mi308"]["groups"] == 4  # :90-123
    with open(os.path.join(HERE, "reference_cases.json"), "w") as f:
        json.dump(out, f, indent=1)

    # SURVEY.md section 8(c), second table: NodeId-level vectors
    def rng(spec):
        res = []
        for part in spec.split(","):
            a, _, b = part.partition("-")
            res += list(range(int(a), int(b or a) + 1))
        return res

    survey = [
        ("mi308", "2,4-6,8-10,12,14,21-25,29,31-33", "", 4, "29,31-33", 180, 6),
        ("mi308", "2,4,6,10-12,15,18,20-22,30,33", "", 5, "2,10-12,15", 340, 12),
        ("mi308", "3,6,8,10,15,17,21-22,29,31,33", "17", 4, "3,6,8,17", 210, 8),
        ("mi308", "2-5,15-17,19-20,25-27,29", "3", 3, "2-4", 90, 4),
        ("cpx", "2-3,5,8,10-12,16-17,20,22-24,29,33,35,41,45,51,53,56,61", "", 6, "10-12,16-17,29", 500, 316),
        ("cpx", "8,11,13-14,17,22,28,34,38,41,44,51,54,57,61,64", "", 6, "34,38,41,51,54,57", 540, 1184),
        ("cpx", "7,12-13,15,17,20,22,28,37,40,42,44,48,50,54-56,58,61-63", "61", 5, "50,54-56,61", 340, 94),
        ("cpx", "2,4-6,9-10,19,21,26,28,33,35,37,41-42,46,53,55,57,63", "63", 6, "2,4-6,9,63", 550, 85),
        ("mi210", "3,5-9", "", 4, "6-9", 240, 360),
        ("mi210", "3,7-9", "9", 2, "7,9", 40, 3),
        ("mi210", "2-9", "", 2, "2,3", 40, 56),
        ("cpx", "2,10,18,26,34,42,50,58", "", 4, "2,10,18,26", 240, 1680),
    ]
    sv = []
    for name, av, rq, k, exp, score, ncand in survey:
        pol = policies[name]
        by_node = {d.NodeId: d.Id for d in pol.devices}
        node_of = {d.Id: d.NodeId for d in pol.devices}
        res = pol.Allocate([by_node[n] for n in rng(av)], [by_node[n] for n in rng(rq)] if rq else [], k)
        assert sorted(node_of[i] for i in res) == sorted(rng(exp)), (name, av, res)
        assert pol.last["score"] == score and pol.last["n_candidates"] == ncand, (name, av, pol.last)
        sv.append({"topology": name, "available_nodes": rng(av), "required_nodes": rng(rq) if rq else [],
                   "size": k, "expected_nodes": rng(exp), "score": score, "n_candidates": ncand,
                   "oracle_ids": res})
    with open(os.path.join(HERE, "survey_cases.json"), "w") as f:
        json.dump(sv, f, indent=1)

    # frozen random regression cases (answers from the BFS restatement)
    rnd = {}
    for name, pol in policies.items():
        ids = [d.Id for d in pol.devices]
        gen = topo.SplitMix64(0x5EED0001 ^ len(ids))
        cases = []
        while len(cases) < 150:
            av, rq, k = topo.random_request(gen, ids, kmax=8)
            res = pol.Allocate(av, rq, k)
            if not pol.last["searched"]:
                continue
            score = pol.last["score"]
            ex = pol.AllocateExhaustive(av, rq, k)
            assert ex == res and pol.last["score"] == score, (name, av, rq, k, res, ex)
            cases.append({"available": av, "required": rq, "size": k, "ids": res, "score": score})
        rnd[name] = cases
    with open(os.path.join(HERE, "random_cases.json"), "w") as f:
        json.dump(rnd, f, separators=(",", ":"))
    print("golden files written")


if __name__ == "__main__":
    main()
