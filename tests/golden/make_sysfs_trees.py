#!/usr/bin/env python
"""Distil the reference's three KFD fixture trees into tests/golden/sysfs_trees.json.gz.

Run HERE (build container, /root/reference mounted).  The GPU box has no /root/reference, and the
`-m gpu` test of BestEffortPolicy.Init(devs, topoDir) needs a sysfs tree to parse; what is frozen
is not a copy of the trees but the only fields the policy reads from them (device.go:107-252):
per node its drm_render_minor, per io_links / p2p_links entry its node_from, node_to and type.
tests/topo.py:materialize_sysfs_tree() writes them back as a directory tree of `properties`
files in the layout the reference globs (nodes/<n>/properties, nodes/<n>/{io,p2p}_links/<l>/properties).
"""
import gzip
import json
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/testdata/"
TREES = {"mi308": "topology-parsing-mi308/topology/nodes", "mi210": "topo-mi210-xgmi-pcie/nodes",
         "cpx": "topo-mi300-cpx/topology/nodes"}
FIELDS = {"minor": re.compile(r"drm_render_minor\s(\d+)"), "from": re.compile(r"node_from\s(\d+)"),
          "to": re.compile(r"node_to\s(\d+)"), "type": re.compile(r"type\s(\d+)")}


def last(path, keys):
    out = {}
    with open(path) as f:
        for line in f.read().split("\n"):
            for k in keys:
                m = FIELDS[k].search(line)
                if m:
                    out[k] = m.group(1)  # kept as the digit string the file holds
    return out


def main():
    trees = {}
    for name, rel in TREES.items():
        root, nodes = REF + rel, {}
        for node in sorted(os.listdir(root)):
            if not node[0].isdigit():
                continue
            entry = {"minor": last(os.path.join(root, node, "properties"), ["minor"]).get("minor")}
            for sub in ("io_links", "p2p_links"):
                d = os.path.join(root, node, sub)
                if os.path.isdir(d):
                    entry[sub] = {ln: [last(os.path.join(d, ln, "properties"), ["from", "to", "type"]).get(k)
                                       for k in ("from", "to", "type")]
                                  for ln in sorted(os.listdir(d)) if ln[0].isdigit()}
            nodes[node] = entry
        trees[name] = nodes
        print(name, len(nodes), "nodes", sum(len(e.get(s, {})) for e in nodes.values() for s in ("io_links", "p2p_links")), "links")
    with gzip.open(os.path.join(HERE, "sysfs_trees.json.gz"), "wt") as f:
        json.dump(trees, f, separators=(",", ":"))


if __name__ == "__main__":
    main()
