#!/usr/bin/env python
"""bench.py -- subsets scored/sec of the GPU preferred-allocation search (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (all N): BASELINE config C4 -- synthetic 80-partition topology (10 GPUs x 8, 2 NUMA
nodes, all XGMI), GetPreferredAllocation for 8 of 80 = C(80,8) = 28 987 537 150 subsets per
step, sharded by rank range over the N GPUs (strong scaling), one min-all-reduce of the packed
winner per step.  A step is one complete search; the result of every step is checked.

  value     subsets/s with the request descriptor already in HBM (tp_prepare), the kernel (and,
            for N>1, the NCCL min-all-reduce) timed with CUDA events on torch's current stream.
  e2e       the same metric through the reference-facing call, BestEffortPolicy.Allocate with
            host id strings: per step the H2D copy of the request and the D2H read of the winner
            are inside the timed region.
  roofline  integer-ALU roofline of the search kernel: achieved = subsets/s x
            integer ops the kernel executes per subset; peak = IADD3 microbenchmark measured in
            this run (MEASURED_PEAKS.json has no integer figure).  HBM traffic is ~1 KB/launch.
  cpu_baseline  oracle/exhaustive.c (the same enumeration, all host threads) on a bounded
            rank-range sample of the same workload.

--impl reference times the reference side on the host cores: no Go toolchain exists in this
image, so it is the C restatement (oracle/, kind "port"): the exhaustive scorer for the
metric's unit, and the reference's own BFS policy latency for the same request beside it.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "subsets_scored_per_sec"
UNIT = "subsets/s"
WORKLOAD = "C4: synthetic 10x8-partition topology (80 devices), GetPreferredAllocation 8 of 80, C(80,8)=28987537150 subsets/step"


def bench_config(world):
    """The `config` block, identical in both arms (the reference arm runs on the GPU arm's config)."""
    return {"workload": WORKLOAD,
            "sharding": f"subset space split x{world} (item stride), one min-exchange of the packed winner per step"
            if world > 1 else "single GPU",
            "l2": "flushed (256 MiB write) between timed steps"}


def host_threads():
    """Host threads the CPU arm may use: the affinity mask, not OMP_NUM_THREADS (torch.distributed.run
    exports OMP_NUM_THREADS=1 to every rank, which would make the reference arm a 1-thread run)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def workload():
    import topo
    devs, pairs = topo.c4_topology()
    ids = [d.Id for d in devs]
    expected = ["test01"] + [f"amdgpu_xcp_{i}" for i in range(1, 8)]
    return devs, pairs, ids, [], 8, expected, 840


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].startswith("Active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def ncu_dram_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the search kernel, from the
    committed `ncu --set full` summary of this round (profiles/); None when there is none."""
    name = {1: "r1_lex_ncu_full.txt", 2: "r1_split_final_c4_ncu_full.txt", 4: "r2_pair_c4_ncu_full.txt"}.get(kernel)
    path = os.path.join(ROOT, "profiles", name) if name else None
    if not path or not os.path.exists(path):
        return None
    total, unit_scale = 0.0, {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for line in open(path):
        parts = line.split()
        if len(parts) >= 4 and parts[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum") and parts[1] == "=":
            total += float(parts[2]) * unit_scale.get(parts[3], 1.0)
    return total


def percentile(xs, p):
    xs = sorted(xs)
    return xs[min(len(xs) - 1, int(p / 100.0 * len(xs)))]


def cpu_exhaustive(ref, ids, req, size, seconds, threads=0, rate=None):
    """Time oracle/exhaustive.c on ranks [0, X) of the workload, X sized for ~`seconds` (the whole
    space when it fits in that time)."""
    probe = 1 << 28
    threads = threads or host_threads()
    if rate is None:
        t = time.perf_counter()
        ref.AllocateExhaustive(ids, req, size, threads=threads, rank_lo=0, rank_hi=probe)
        dt = time.perf_counter() - t
        rate = ref.last["scored"] / dt
    total = 28987537150
    sample = int(min(max(rate * seconds, probe), total))
    t = time.perf_counter()
    ref.AllocateExhaustive(ids, req, size, threads=threads, rank_lo=0, rank_hi=sample)
    dt = time.perf_counter() - t
    return ref.last["scored"] / dt, ref.last["threads"], ref.last["scored"], dt


def run_reference(args, rank):
    """--impl reference: the CPU side (oracle port; the Go reference cannot run here)."""
    if rank != 0:
        return
    from oracle.binding import RefPolicy
    devs, pairs, ids, req, size, expected, score = workload()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    # the reference's own algorithm (BFS over partition groups) on this request
    lat = []
    for _ in range(200):
        t = time.perf_counter()
        out = ref.Allocate(ids, req, size)
        lat.append((time.perf_counter() - t) * 1e6)
    assert out == expected and ref.last["score"] == score
    n_cand = ref.last["n_candidates"]
    # A step scores the whole C(80,8) when the host does that in <= 8 s (16 cores: ~3.8 s), else the
    # first 8 s worth of ranks; warm-up steps are short (they only spin the OpenMP team up).
    per_step_s = 8.0
    rate = None
    for _ in range(max(args.warmup, 1)):
        rate, _, _, _ = cpu_exhaustive(ref, ids, req, size, 0.5)
    scored, secs, threads = 0, 0.0, 1
    for _ in range(args.steps):
        r, threads, n, dt = cpu_exhaustive(ref, ids, req, size, per_step_s, rate=rate)
        scored += n
        secs += dt
    value = scored / secs
    per_step = scored // max(args.steps, 1)
    sample = (f"{'all' if per_step == 28987537150 else 'ranks [0, ' + str(per_step) + ') of the'} "
              f"28987537150 subsets of C(80,8) per step, oracle/exhaustive.c, {threads} OpenMP threads")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / max(args.steps, 1) * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": bench_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "reference_policy": {"what": "BFS restatement of BestEffortPolicy.Allocate (oracle/ref_policy.c), 1 thread",
                                 "candidates_scored": n_cand, "p50_us": percentile(lat, 50),
                                 "p99_us": percentile(lat, 99)},
            "host_cores": os.cpu_count(), "omp_num_threads_env": os.environ.get("OMP_NUM_THREADS")}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the C1/C2 latency and C3 legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank)

    import torch
    import torch.distributed as dist
    from __graft_entry__ import package
    tp = package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the search has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        cpu_group = dist.new_group(backend="gloo")  # handle exchange and waits that must not occupy a GPU
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run"

    devs, pairs, ids, req, size, expected, score = workload()
    pol = tp.BestEffortPolicy(cuda_device=local_rank)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    # ---- value: device-resident request, CUDA events on torch's current stream ----------
    prep = tp.PreparedSearch(pol, ids, req, size, shard_index=rank, shard_count=world)
    out = torch.zeros(3, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream()
    launches = 0

    # The path's one exchange step -- the min of the shards' packed (score, RefKey) winners -- is fused into
    # the search kernel (its last block stores the winner into every rank's buffer over NVLink and waits for
    # the others'); the NCCL min-all-reduce it replaces is timed further down for comparison.
    exchange = None
    if world > 1:
        def gather_handles(raw):
            got = [None] * world
            dist.all_gather_object(got, raw, group=cpu_group)
            return got
        exchange = tp.Exchange(pol, rank, world, gather_handles)

    def step():
        if exchange is not None:
            return prep.launch_exchange(exchange, stream.cuda_stream, out.data_ptr())
        return prep.launch(stream.cuda_stream, out.data_ptr())

    def step_nccl():
        n = prep.launch(stream.cuda_stream, out.data_ptr())
        dist.all_reduce(out[0:1], op=dist.ReduceOp.MIN)
        return n

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(args.warmup):
        step()
    barrier()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    for e0, e1 in evs:
        flush.fill_(1)  # L2 flush between timed iterations (untimed)
        e0.record(stream)
        launches += step()
        e1.record(stream)
    barrier()
    dev_ms = max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in evs))
    # A search takes a few ms, far less than nvidia-smi's sampling period: keep the same kernel
    # running (untimed) for ~1.5 s so that the clock / throttle record is taken under this load.
    sustain = int(min(max(1500.0 / max(dev_ms / args.steps, 1e-3), 1), 5000))
    t_s = time.perf_counter()
    for _ in range(sustain):
        step()
    barrier()
    sustain_ms = (time.perf_counter() - t_s) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["sampled_over"] = f"warm-up + {args.steps} timed steps + {sustain} untimed steps of the same search ({sustain_ms:.0f} ms)"
    key_hi = int(out[0].item())
    total_subsets = int(out[2].item())  # (with the fused exchange: already the sum over the ranks)
    got = pol.Decode(ids, req, size, (key_hi, 0))
    assert got == expected and (key_hi >> 40) == score, (got, key_hi)
    assert total_subsets == 28987537150, total_subsets
    value = total_subsets * args.steps / (dev_ms * 1e-3)

    # ---- e2e: the reference-facing call, host buffers in, host ids out -------------------
    h2d, d2h = tp.lib().tp_descriptor_bytes(0), 24  # the prepared request up, key + count down

    e2e_out = torch.zeros(3, dtype=torch.int64, device=dev)
    e2e_host = torch.zeros(3, dtype=torch.int64).pin_memory()

    def e2e_step():
        if world == 1:
            return pol.Allocate(ids, req, size)
        # every rank: validate + upload the request (H2D), search its shard with the fused exchange, read the
        # global winner back (D2H), decode it to ids
        q = tp.PreparedSearch(pol, ids, req, size, shard_index=rank, shard_count=world)
        q.launch_exchange(exchange, stream.cuda_stream, e2e_out.data_ptr())
        e2e_host.copy_(e2e_out, non_blocking=True)
        stream.synchronize()
        q.close()
        return pol.Decode(ids, req, size, (int(e2e_host[0]), int(e2e_host[1])))

    for _ in range(min(args.warmup, 2)):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = e2e_step()
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    assert res == expected
    e2e_value = 28987537150 * args.steps / e2e_s

    # ---- A/B of the exchange (untimed for the headline): kernel + NCCL min-all-reduce -----------
    exchange_ab = None
    if world > 1:
        for _ in range(3):
            step_nccl()
        barrier()
        ab = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
        for e0, e1 in ab:
            flush.fill_(1)
            e0.record(stream)
            step_nccl()
            e1.record(stream)
        barrier()
        nccl_ms = max_over_ranks(sum(e0.elapsed_time(e1) for e0, e1 in ab)) / args.steps
        exchange_ab = {"fused_in_kernel_ms_per_step": dev_ms / args.steps, "kernel_plus_nccl_allreduce_ms_per_step": nccl_ms}

    # ---- one process, one call, all GPUs (what a Go plugin can do: it cannot torchrun) -----------
    single_process = None
    if world > 1:
        if rank == 0:
            mp = tp.BestEffortPolicy(cuda_device=0, extra_devices=list(range(1, world)))
            mp.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
            for _ in range(3):
                assert mp.Allocate(ids, req, size) == expected
            t1 = time.perf_counter()
            for _ in range(args.steps):
                mp.Allocate(ids, req, size)
            dt = (time.perf_counter() - t1) / args.steps
            single_process = {"what": "BestEffortPolicy.Allocate of ONE process on all GPUs (tp_options.extra_devices): "
                                      "host ids in, host ids out, shards launched by per-GPU threads, host-side min",
                              "n_gpus": world, "ms_per_call": dt * 1e3, "subsets_per_s": 28987537150 / dt,
                              "kernel_ms_slowest_gpu": mp.last["kernel_ms"], "gpu_launches": mp.last["gpu_launches"]}
            mp.close()
        dist.barrier(group=cpu_group)  # the other ranks wait on the CPU: their GPUs are rank 0's for this leg

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel (this rank's shard, kernel only) ------------
        k_ms, _, info = tp.search_resident(pol, ids, req, size, iters=max(args.steps, 3), shard_index=rank,
                                           shard_count=world)
        k_ms /= max(args.steps, 3)
        shard_subsets = prep.subsets
        packed = info["kernel"] in (tp.KERNEL_SPLIT, tp.KERNEL_TRI)
        if packed:
            h2d = tp.lib().tp_descriptor_bytes(1)  # the work plan (TpPlan) travels with the request descriptor
        peak_iadd = tp.measure_int_peak(local_rank, 0)      # IADD3, 32-bit lane-ops/s (x1e9)
        peak_addmin = tp.measure_int_peak(local_rank, 3)    # VIADDMNMX.U16x2 lane-instr/s (x1e9)
        rate = shard_subsets / (k_ms * 1e-3)
        if packed:
            # TRI and SPLIT execute one 16-bit add and one 16-bit min per subset (VIADDMNMX.U16x2: two
            # subsets per lane-instruction); the peak is that instruction's measured rate x 4 ops.
            ops_per_subset, unit_note = 2, "16-bit integer add/min ops"
            kname = "tp_exhaustive_tri<80,1>" if info["kernel"] == tp.KERNEL_TRI else "tp_exhaustive_split<80,2>"
            peak_tops = peak_addmin * 4 / 1e3
            peak_src = "VIADDMNMX.U16x2 microbenchmark in this run (tp_measure_int_peak variant 3) x 4 ops/lane-instr"
        else:
            # LEX at k'=8: 7 weight gathers + unary term = 8 integer adds + 1 compare per subset
            ops_per_subset, kname, unit_note = 9, "tp_exhaustive_lex<8>", "32-bit integer ops"
            peak_tops = peak_iadd / 1e3
            peak_src = "IADD3 microbenchmark in this run (tp_measure_int_peak variant 0)"
        achieved = rate * ops_per_subset / 1e12
        roofline = {"bound": "int-alu", "achieved": achieved, "peak": peak_tops, "unit": "Tops/s",
                    "frac": achieved / peak_tops, "traffic": ncu_dram_traffic(info["kernel"]), "ops": unit_note,
                    "peak_source": peak_src + "; MEASURED_PEAKS.json has no integer figure",
                    "peak_iadd3_tops": peak_iadd / 1e3, "peak_viaddmnmx_u16x2_glane_instr": peak_addmin,
                    "ops_per_subset": ops_per_subset, "kernel": kname, "kernel_ms": k_ms,
                    "kernel_subsets_per_s": rate,
                    "north_star_15op_frac": rate * 15 / 1e12 / (peak_iadd / 1e3),
                    "hbm_bytes_per_launch_algorithmic": h2d + d2h}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None,
                "dtype": "u16" if packed else "int32", "data": "synthetic",
                "config": bench_config(world),
                "clocks": clocks,
                "e2e": ({"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                         "ms_per_step": e2e_s / args.steps * 1e3, "api": "BestEffortPolicy.Allocate(ids, req, size)"}
                        if world == 1 else
                        # N GPUs behind the reference-facing call: ONE process, one Allocate (a Go plugin cannot
                        # torchrun), every GPU gets the request (H2D) and returns its winner (D2H) inside the call
                        {"value": single_process["subsets_per_s"], "unit": UNIT, "h2d_bytes_per_step": h2d * world,
                         "d2h_bytes_per_step": d2h * world, "ms_per_step": single_process["ms_per_call"],
                         "api": "BestEffortPolicy.Allocate(ids, req, size) of one process over all GPUs "
                                "(tp_options.extra_devices), timed on rank 0 while the other ranks wait on the CPU"}),
                "gpu_launches": launches, "roofline": roofline}
        if exchange_ab:
            line["exchange_ab"] = exchange_ab
        if single_process:
            line["single_process_multi_gpu"] = single_process
            line["e2e_rank_processes"] = {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_s / args.steps * 1e3,
                                          "api": "per rank: tp_prepare (H2D) + tp_launch_exchange + D2H of the global winner + tp_decode"}
        if world == 1 and not args.no_extra:
            line["latency"] = latency_leg(tp, dev)
            line["c3"] = c3_leg(tp, peak_addmin)
            line["c5_pod_storm"] = pod_storm_leg(tp)
        if world == 1 and not args.no_cpu:
            from oracle.binding import RefPolicy
            ref = RefPolicy()
            ref.InitPairs(devs, pairs)
            lat = []
            for _ in range(200):  # the reference's own algorithm on this request, beside the exhaustive numbers
                t = time.perf_counter()
                ref.Allocate(ids, req, size)
                lat.append((time.perf_counter() - t) * 1e6)
            line["reference_policy"] = {"what": "BFS restatement of BestEffortPolicy.Allocate (oracle/ref_policy.c), 1 thread, "
                                                "same request; the exhaustive search returns the same ids",
                                        "candidates_scored": ref.last["n_candidates"], "p50_us": percentile(lat, 50),
                                        "p99_us": percentile(lat, 99)}
            r, threads, n, dt = cpu_exhaustive(ref, ids, req, size, 12.0)
            line["cpu_baseline"] = {"value": r, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"ranks [0, {n}) of C(80,8) in {dt:.1f}s, oracle/exhaustive.c, "
                                              f"{os.cpu_count()} host cpus visible"}
    prep.close()
    if exchange is not None:
        exchange.close()
    pol.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if line:
        print(json.dumps(line))


def latency_leg(tp, dev):
    """p50/p99 GetPreferredAllocation latency on the latency-bound configs C1 and C2."""
    import topo
    from oracle.binding import RefPolicy
    out = {}
    for name, fixture, pick, size in (("c1_mi210_2of8", "mi210", None, 2),
                                      ("c2_mesh_4of8", "cpx", [f"test{i}" for i in range(1, 9)], 4)):
        devs, pairs = topo.fixture_topology(fixture)
        ids = pick or [d.Id for d in devs]
        pol = tp.BestEffortPolicy(cuda_device=dev.index)
        pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
        ref = RefPolicy()
        ref.InitPairs(devs, pairs)
        want = ref.Allocate(ids, [], size)
        av = pol.indices(ids)
        for _ in range(200):
            tp.allocate_indices(pol, av, [], size)
        abi, host, cpu = [], [], []
        for _ in range(10000):  # >= 10 000 sequential calls at the policy boundary
            t = time.perf_counter()
            got_idx, info = tp.allocate_indices(pol, av, [], size)
            abi.append((time.perf_counter() - t) * 1e6)
        for _ in range(3000):
            t = time.perf_counter()
            got = pol.Allocate(ids, [], size)
            host.append((time.perf_counter() - t) * 1e6)
        assert got == want
        for _ in range(3000):
            t = time.perf_counter()
            ref.Allocate(ids, [], size)
            cpu.append((time.perf_counter() - t) * 1e6)
        # 8 concurrent callers on one policy (one goroutine per RPC in the reference's gRPC server)
        import threading
        conc = [[] for _ in range(8)]

        def caller(slot):
            for _ in range(500):
                t0 = time.perf_counter()
                tp.allocate_indices(pol, av, [], size)
                conc[slot].append((time.perf_counter() - t0) * 1e6)

        threads = [threading.Thread(target=caller, args=(s,)) for s in range(8)]
        t0 = time.perf_counter()
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        conc_s = time.perf_counter() - t0
        conc_all = [x for c in conc for x in c]
        # kernel-only time from CUDA events: the latency path records no events, so the same search is
        # timed device-resident through the regular multi-block launch (an upper bound for its kernel)
        tp.search_resident(pol, ids, [], size, iters=3)  # first launch of an instantiation loads its code
        k_ms, _, _ = tp.search_resident(pol, ids, [], size, iters=200)
        native = None
        try:  # the same request from a native caller: no interpreter, no GIL (SURVEY 8(d) latency protocol)
            sys.path.insert(0, os.path.join(ROOT, "tests", "harness"))
            import run_harness
            run_harness.build()
            native = run_harness.run(name, devs, pairs, ids, [], size)
            assert native["mismatches"] == 0
        except Exception as e:  # the harness is tooling: its absence must not take the bench line down
            native = {"unavailable": repr(e)[:200]}
        out[name] = {"native": native, "subsets": info["subsets_total"], "c_abi_p50_us": percentile(abi, 50),
                     "c_abi_mean_us": sum(abi) / len(abi), "kernel_events_us": k_ms / 200 * 1e3,
                     "concurrent8_p50_us": percentile(conc_all, 50), "concurrent8_p99_us": percentile(conc_all, 99),
                     "concurrent8_calls_per_s": len(conc_all) / conc_s,
                     "c_abi_p99_us": percentile(abi, 99), "policy_p50_us": percentile(host, 50),
                     "policy_p99_us": percentile(host, 99),
                     "cpu_reference_bfs_p50_us": percentile(cpu, 50),
                     "cpu_reference_candidates": ref.last["n_candidates"]}
        pol.close()
    out["note"] = ("python columns: wall clock around the ctypes call (interpreter overhead included, concurrent callers share the "
                   "GIL); `native`: tests/harness/abi_harness.cpp, clock_gettime(CLOCK_MONOTONIC) around tp_allocate from C++ "
                   "threads -- the numbers a cgo caller would see")
    return out


def pod_storm_leg(tp, count=10000):
    """BASELINE config C5: 10 000 requests with random availability masks over the CPX topology
    (SURVEY 8d generator, splitmix64 seed 0x5EED0001), answered by ONE tp_allocate_batch call."""
    import topo
    from oracle.binding import RefPolicy
    devs, pairs = topo.fixture_topology("cpx")
    ids = [d.Id for d in devs]
    gen = topo.SplitMix64(0x5EED0001)
    reqs = []
    while len(reqs) < count:
        p = (0.25, 0.5, 0.75)[gen.below(3)]
        av = [i for i in ids if gen.chance(p)]
        k = (1, 2, 4, 8)[gen.below(4)]
        if len(av) < k:
            continue
        reqs.append((av, [av[gen.below(len(av))]] if gen.chance(0.2) else [], k))
    pol = tp.BestEffortPolicy(cuda_device=0)
    pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
    batch = tp.Batch(pol, reqs)
    batch.run()  # warm-up (buffers grow here)
    times = []
    for _ in range(3):
        t = time.perf_counter()
        batch.run()
        times.append(time.perf_counter() - t)
    got, infos = batch.results()
    ref = RefPolicy()
    ref.InitPairs(devs, pairs)
    t = time.perf_counter()
    want = [ref.Allocate(*r) for r in reqs]
    cpu_s = time.perf_counter() - t
    assert got == want
    subsets = sum(i["subsets_scored"] for i in infos)
    best = min(times)
    pol.close()
    return {"requests": count, "batch_ms": best * 1e3, "requests_per_s": count / best,
            "subsets_scored": subsets, "subsets_per_s": subsets / best,
            "gpu_launches": max(i["gpu_launches"] for i in infos),
            "cpu_reference_bfs_requests_per_s": count / cpu_s,
            "note": "wall clock around tp_allocate_batch (host prepare + H2D + kernels + D2H + decode); "
                    "cpu column = oracle BFS restatement, 1 thread, python call overhead included"}


def c3_leg(tp, peak_addmin):
    """BASELINE config C3: CPX fixture 8 of 63 (parity case) and 8 of 64 (the ~4.4e9 figure), each with the
    same roofline arithmetic as the headline kernel (2 sixteen-bit ops per subset against the measured
    VIADDMNMX.U16x2 rate x 4)."""
    import topo
    res = {}
    for name, (devs, pairs), first in (("cpx_8of63", topo.fixture_topology("cpx"), "test1"),
                                       ("cpx_8of64", topo.cpx64_topology(), "test1")):
        ids = [d.Id for d in devs]
        pol = tp.BestEffortPolicy(cuda_device=0)
        pol.InitWithWeights([tp.Device(d.Id, d.NodeId, d.NumaNode, d.DevId) for d in devs], pairs)
        tp.search_resident(pol, ids, [], 8, iters=3)
        ms, key, info = tp.search_resident(pol, ids, [], 8, iters=20)
        got = pol.Decode(ids, [], 8, key)
        assert got[0] == first and info["score"] == 840, (got, info)
        rate = info["subsets_total"] / (ms / 20 * 1e-3)
        achieved, peak = rate * 2 / 1e12, peak_addmin * 4 / 1e3
        res[name] = {"subsets": info["subsets_total"], "kernel_ms": ms / 20, "subsets_per_s": rate, "winner": got,
                     "roofline": {"bound": "int-alu", "achieved": achieved, "peak": peak, "unit": "Tops/s",
                                  "frac": achieved / peak, "ops_per_subset": 2,
                                  "kernel": tp.kernel_name(info["kernel"]),
                                  "peak_source": "VIADDMNMX.U16x2 microbenchmark in this run x 4 ops/lane-instr"}}
        pol.close()
    return res


if __name__ == "__main__":
    main()
