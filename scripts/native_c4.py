"""Native end-to-end latency of the headline request (8 of 80) and of 8 of 64: tp_allocate from C++, clock_gettime."""
import os, sys, json
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests")); sys.path.insert(0, os.path.join(R, "tests", "harness"))
import topo, run_harness
run_harness.build()
for name, mk in (("c4_8of80", topo.c4_topology), ("cpx64_8of64", topo.cpx64_topology)):
    devs, pairs = mk()
    ids = [d.Id for d in devs]
    out = run_harness.run(name, devs, pairs, ids, [], 8, sequential=300, threads=2, per_thread=20)
    print(name, {k: out[k] for k in ("c_abi_sequential", "policy_strings_sequential", "c_abi_concurrent", "mismatches")}, flush=True)
