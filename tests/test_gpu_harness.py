"""The native harness (tests/harness/abi_harness.cpp) replays the cgo call sequence of INTEGRATION.md against the
C ABI -- create, sequential calls, concurrent callers, destroy -- and checks every answer against the oracle."""
import os
import sys

import pytest

import topo

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "harness"))
import run_harness  # noqa: E402


@pytest.mark.parametrize("name,fixture,pick,req,size", [
    ("c1_mi210_2of8", "mi210", None, 0, 2),
    ("c2_mesh_4of8", "cpx", [f"test{i}" for i in range(1, 9)], 0, 4),
    ("cpx_5of24_required", "cpx", "first24", 1, 5)])
def test_native_sequence(name, fixture, pick, req, size):
    run_harness.build()
    devs, pairs = topo.fixture_topology(fixture)
    ids = [d.Id for d in devs]
    avail = ids[:24] if pick == "first24" else (pick or ids)
    res = run_harness.run(name, devs, pairs, avail, avail[:req], size, sequential=600, threads=8, per_thread=150)
    assert res["mismatches"] == 0 and res["c_abi_sequential"]["calls"] == 600
    assert res["c_abi_concurrent"]["calls"] == 8 * 150 and res["c_abi_concurrent"]["threads"] == 8
