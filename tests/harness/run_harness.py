"""Write the input file of tests/harness/abi_harness and run it (used by bench.py's latency leg and the GPU test)."""
import json
import os
import subprocess
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
BINARY = os.path.join(HERE, "abi_harness")


def build(target="all"):
    subprocess.run(["make", "-C", HERE, "-s", target], check=True)
    return BINARY


def write_input(path, name, devs, pairs, avail, req, size):
    with open(path, "w") as f:
        f.write(f"config {name}\ndevices {len(devs)}\n")
        for d in devs:
            f.write(f"{d.Id} {d.NodeId} {d.NumaNode} {d.DevId}\n")
        f.write(f"pairs {len(pairs)}\n")
        for a, b, w in pairs:
            f.write(f"{a} {b} {w}\n")
        f.write(f"request {len(avail)} {len(req)} {size}\n{' '.join(avail)}\n{' '.join(req)}\n")


def run(name, devs, pairs, avail, req, size, sequential=10000, threads=8, per_thread=2000, binary=None, env=None):
    """-> the harness's JSON line as a dict (p50/p99/mean of the native tp_allocate calls, the string-level policy,
    concurrent callers and the CPU restatement of the reference's BFS)."""
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "input.txt")
        write_input(path, name, devs, pairs, avail, req, size)
        r = subprocess.run([binary or BINARY, path, str(sequential), str(threads), str(per_thread)], capture_output=True,
                           text=True, env=env)
    if r.returncode != 0 or not r.stdout.strip():
        raise RuntimeError(f"abi_harness failed ({r.returncode}): {r.stdout[-500:]} {r.stderr[-2000:]}")
    return json.loads(r.stdout.strip().splitlines()[-1])
