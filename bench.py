#!/usr/bin/env python
"""bench.py -- headline measurements of the two hot paths on B200 (contract: task brief, section 4).

Top level of the JSON line = HP-2, the batched reroute SSSP (BASELINE.json config 5: 224x224 grid, 199,808
links, 50,176 nodes), because that is the path that shards across GPUs.  One step = one batch of `--batch`
origins per GPU, full f64 trees + canonical predecessors written to HBM; N>1 is weak scaling (every rank
gets its own `--batch` origins) and the step then ends with the all-gather of the int32 next-hop tables.
The `tick` object (N=1, rank 0) = HP-1 on BASELINE.json config 4: 159x159 grid (100,488 real + 25,281 source
links), 1M agents departing over 2 h, 14,400 one-second ticks, measured as agent-steps/s.

`--impl reference` times the CPU side instead: oracle/_ref/liblsp.so (the reference's own sp library compiled
from its sources) on every host core for HP-2, and the C restatement of the tick (oracle/tick_oracle.c, one
core: the reference tick has no parallel mode and is Python, which cannot travel to the GPU box) for HP-1.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from spatial_queue_b200 import synth  # noqa: E402

SSSP_GRID = 224
TICK_GRID = 159
TICK_AGENTS = 1_000_000
TICK_TICKS = 14_400
TICK_HORIZON = 7_200
REROUTE_FREQ = 60          # BASELINE.json config 4 / north_star: reroute every 60 s
WEIGHT_FREQ = 100          # run_traffic_simulation.py:83
SHELTER_AFTER = 3 * 3600   # run_traffic_simulation.py:103
# roofline.traffic comes from this round's own `ncu --set full` capture: tools/make_profile_summary.py writes
# dram__bytes_read.sum + dram__bytes_write.sum per launch (per tree / per tick) and the file they came from into
# profiles/traffic.json; with no capture the field is null.
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "traffic.json")


def measured_traffic(kind):
    try:
        return json.load(open(TRAFFIC_FILE)).get(kind)
    except Exception:
        return None


def bind_to_gpu_numa_node(torch, local):
    """N > 1: run this rank on the cores of its GPU's NUMA node, so that the page-locked host buffers of the e2e calls
    (cudaHostAlloc places them where the allocating thread runs) sit behind the GPU's own root complex instead of all
    eight ranks writing through one socket.  Returns a note for the JSON line."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return f"gpu {local}: no NUMA node reported"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"gpu {local}: NUMA node {node} has no allowed cores"
        os.sched_setaffinity(0, cpus)
        return f"gpu {local} -> NUMA node {node} ({len(cpus)} cores)"
    except Exception as ex:
        return f"not bound ({type(ex).__name__}: {ex})"


def sssp_workload_name(n, E, V):
    """The same string in both arms (the driver compares config.workload of `ours` and `reference`)."""
    return f"config 5: batched SSSP reroute sweep, {n}x{n} grid ({E} links, {V} nodes), full trees + canonical predecessors"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[1]) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) < 9:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        busy = [x for x in sm if mx and x > 0.5 * mx[0]] or sm
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
# HP-2 workload
# ---------------------------------------------------------------------------------------------------------------
def sssp_workload(n=SSSP_GRID):
    f = synth.flat_grid(n, seed=1)
    return f.edge_start, f.edge_end, f.edge_fft, n * n


def origins_for(rank, batch, V, step=0):
    perm = np.random.default_rng(1234 + step).permutation(V)
    idx = (np.arange(batch) + rank * batch) % V
    return perm[idx].astype(np.int32)


def _ref_worker(args):
    s, e, w, origins = args
    from oracle import sssp as osssp
    g = osssp.RefGraph(s, e, w, max_vertex_id=int(max(s.max(), e.max())))
    t0 = time.perf_counter()
    g.time_trees(origins)
    return time.perf_counter() - t0


def sssp_reference_trees_per_s(s, e, w, V, n_trees_per_core, cores):
    """The reference's own library on `cores` processes (its only parallel strategy, util/dta_meso.py:513-536)."""
    import multiprocessing as mp
    from oracle import sssp as osssp
    if not osssp.ref_available():
        return None
    chunks = [(s, e, w, origins_for(c, n_trees_per_core, V, step=99)) for c in range(cores)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(cores) as pool:
        pool.map(_ref_worker, chunks)
    wall = time.perf_counter() - t0
    return cores * n_trees_per_core / wall


def calibrate_ref_trees(s, e, w, V, target_s):
    from oracle import sssp as osssp
    g = osssp.RefGraph(s, e, w, max_vertex_id=V - 1)
    t0 = time.perf_counter()
    g.time_trees(origins_for(0, 8, V, step=98))
    per = (time.perf_counter() - t0) / 8
    return max(8, int(target_s / per)), per


# ---------------------------------------------------------------------------------------------------------------
# HP-1 workload
# ---------------------------------------------------------------------------------------------------------------
def tick_bytes(counters, n_links):
    """Algorithmic bytes of a run: 48 B per link per tick + 8 B per on-link vehicle-step + 48 B per
    move + 16 B per active node."""
    return (48 * n_links * counters["ticks"] + 8 * counters["agent_steps"] + 48 * counters["moves"]
            + 16 * counters["active_nodes"])


def flat_nbytes(f):
    return int(sum(v.nbytes for v in f.as_dict().values()))


def tick_cases(args):
    """The three tick workloads: BASELINE.json configs 4 (headline), 2 and 3 (stand-ins of SURVEY.md 8d: the reference
    ships neither the Berkeley nor the Butte inputs)."""
    c4 = {"key": "tick", "name": f"config 4: {args.tick_grid}x{args.tick_grid} grid", "ticks": args.tick_ticks,
          "flat": lambda: synth.grid_scenario(args.tick_grid, args.tick_agents, horizon=TICK_HORIZON, seed=1, device_routes=True),
          "flat_cpu": lambda: synth.grid_scenario(args.tick_grid, args.tick_agents, horizon=TICK_HORIZON, seed=1),
          "router": (REROUTE_FREQ, WEIGHT_FREQ), "rules": None, "closures": False, "shelter": SHELTER_AFTER,
          "window": args.cpu_window,
          "what": f"every agent rerouted each {REROUTE_FREQ} s on weights refreshed every {WEIGHT_FREQ} s (8 reverse trees per "
                  f"refresh), give-up after {SHELTER_AFTER} s on a link; all on the device",
          "cpu_note": "routes fixed at t=0, no rerouting"}
    c2 = {"key": "tick_c2", "name": "config 2: Berkeley stand-in (OSM-like network, signal + priority nodes)", "ticks": 1401,
          "flat": lambda: synth.berkeley_standin(), "flat_cpu": None,
          "router": None, "rules": None, "closures": False, "shelter": SHELTER_AFTER, "window": 1401,
          "what": "benchmark/profile_berkeley.py:144: ticks 0..1400, routes fixed at t=0, no rerouting",
          "cpu_note": "the same run (no rerouting in this configuration)"}
    c3 = {"key": "tick_c3", "name": "config 3: Butte stand-in (evacuation to six exits)", "ticks": 18001,
          "flat": lambda: synth.butte_standin(device_routes=True), "flat_cpu": lambda: synth.butte_standin(device_routes=False),
          "router": (REROUTE_FREQ, WEIGHT_FREQ), "rules": (2.0, 10), "closures": True, "shelter": 9000, "window": 1800,
          "what": f"model/dta_meso_butte.py: rerouting every {REROUTE_FREQ} s on weights refreshed every {WEIGHT_FREQ} s, shelter_a2 "
                  "when a reroute more than doubles the trip, reroute upon closure (stagger 10), closure schedule of :56-115 "
                  "(fire closures every 300 s with p = 0.15, reopening after 3,600 s, three timed arterial closures) as a device "
                  "event table, give-up after 9,000 s; all on the device, no host call per tick",
          "cpu_note": "routes fixed at t=0, no rerouting, no closures"}
    return [c4] if args.no_configs23 else [c4, c2, c3]


def make_tick_sim(flat, case, stream=None):
    """A workload as the drivers run it, all on the device (see tick_cases)."""
    from spatial_queue_b200.tick import DeviceSim
    sim = DeviceSim(flat, seed=7, routes_from_router=case["router"] is not None)
    if stream is not None:
        sim.set_stream(stream)
    if case["router"]:
        sim.attach_router(synth.routing_graph(flat), flat, reroute_freq=case["router"][0], weight_freq=case["router"][1])
    if case["rules"]:
        sim.set_reroute_rules(*case["rules"])
    if case["closures"]:
        ev = synth.butte_closure_schedule(flat)
        sim.set_closure_schedule([e[0] for e in ev], [e[1] for e in ev], [e[2] for e in ev])
    sim.set_shelter_after(case["shelter"])
    return sim


def run_tick_gpu(case, args, torch, peak):
    t0 = time.perf_counter()
    flat = case["flat"]()
    build_s = time.perf_counter() - t0
    L = len(flat.link_end_node)
    ticks, window = case["ticks"], min(case["window"], case["ticks"])
    stream = torch.cuda.current_stream().cuda_stream
    res = {}
    # device-resident run: state already in HBM when the clock starts
    warm = make_tick_sim(flat, case, stream)             # throwaway instance: first-launch costs, cold caches
    warm.run(0, 256); warm.sync(); warm.close()
    best = None
    for rep in range(max(1, args.tick_reps)):
        sim = make_tick_sim(flat, case, stream)
        sim.sync()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ev[0].record()
        sim.run(0, window)
        ev[1].record()
        cw = sim.counters()                              # the device's own counters at the window boundary (syncs)
        ev[2].record()
        if ticks > window:
            sim.run(window, ticks - window)
        ev[3].record()
        sim.sync()
        ms_window = ev[0].elapsed_time(ev[1])
        ms = ms_window + ev[2].elapsed_time(ev[3])
        c = sim.counters()
        if best is None or ms < best[0]:
            best = (ms, ms_window, c, cw)
        if rep + 1 < max(1, args.tick_reps):
            sim.close()
    ms, ms_window, c, cw = best
    res["value"] = c["agent_steps"] / (ms / 1e3)
    res["ms_per_tick"] = ms / ticks
    res["counters"] = c
    nbytes = tick_bytes(c, L)
    tr = measured_traffic(case["key"])
    res["roofline"] = {"bound": "hbm", "achieved": nbytes / (ms / 1e3) / 1e9, "peak": peak, "unit": "GB/s",
                       "traffic": tr["bytes_per_tick"] if tr else None, "traffic_source": tr["source"] if tr else None,
                       "bytes_per_tick": nbytes / max(c["ticks"], 1)}
    res["roofline"]["frac"] = res["roofline"]["achieved"] / peak
    arrive = sim.agent_state()["arrive_tick"]
    res["stopped"] = {k: v for k, v in sim.summary().items() if k in ("arrive", "shelter_p", "shelter_a1", "shelter_a2", "closed_links")}
    sim.close()
    # end to end: host arrays in, full run, arrival ticks + link state out
    t0 = time.perf_counter()
    sim = make_tick_sim(flat, case)
    sim.run(0, ticks)
    ag = sim.agent_state()
    ls = sim.link_state()
    wall = time.perf_counter() - t0
    assert (ag["arrive_tick"] == arrive).all(), "two runs of the same scenario differ"
    c2 = sim.counters()
    sim.close()
    res["e2e"] = {"value": c2["agent_steps"] / wall, "unit": "agent-steps/s", "wall_s": wall,
                  "h2d_bytes_per_step": flat_nbytes(flat),
                  "d2h_bytes_per_step": int(sum(v.nbytes for v in ag.values()) + sum(v.nbytes for v in ls.values()))}
    res["window_ticks"] = window
    res["window_ms"] = ms_window
    res["window_agent_steps"] = cw["agent_steps"]
    res["build_s"] = build_s
    res["arrived"] = int((arrive >= 0).sum())
    return res, flat


def python_reference_note():
    """The Python reference itself, timed in the build container (tools/time_reference_tick.py; it cannot travel to the
    GPU box): one line for the cpu_baseline sample text."""
    try:
        rows = [json.loads(l) for l in open(os.path.join(ROOT, "profiles", "r2_reference_tick_timing.jsonl")) if l.startswith("{")]
        g = [r for r in rows if r["case"].startswith("50x50")][0]
        return (f"; the unmodified Python reference on one core of the build container: {g['on_link_agent_steps_per_s']:.3g} "
                f"on-link agent-steps/s ({g['s_per_tick']:.3f} s/tick) on a 50x50 grid with 20k agents "
                f"(profiles/r2_reference_tick_timing.jsonl)")
    except Exception:
        return ""


def run_tick_cpu(flat, ticks):
    from oracle.tick import OracleSim
    sim = OracleSim(flat, seed=7)
    t0 = time.perf_counter()
    sim.run(0, ticks)
    wall = time.perf_counter() - t0
    c = sim.counters()
    sim.close()
    return c["agent_steps"] / wall, wall, c


# ---------------------------------------------------------------------------------------------------------------
def reference_arm(args, rank, world):
    """CPU implementation of both paths on the host cores; rank 0 only."""
    if rank != 0:
        return
    from oracle import sssp as osssp
    from oracle import tick as _  # noqa: F401  (builds nothing; import check)
    s, e, w, V = sssp_workload(args.sssp_grid)
    cores = os.cpu_count() or 1
    kind = "reference" if osssp.ref_available() else "port"
    per_core, per_tree = calibrate_ref_trees(s, e, w, V, target_s=args.ref_seconds) if kind == "reference" else (64, None)
    times = []
    for step in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        if kind == "reference":
            tps = sssp_reference_trees_per_s(s, e, w, V, per_core, cores)
        else:
            og = osssp.OracleGraph(s, e, w)
            t1 = time.perf_counter()
            og.sssp_batch(origins_for(0, per_core, V, step=99))
            tps = per_core / (time.perf_counter() - t1)
            cores = 1
        if step >= args.warmup:
            times.append(tps)
    value = float(np.mean(times))
    sample = f"{per_core} full trees per core x {cores} processes per step on the {args.sssp_grid}x{args.sssp_grid} grid"
    line = {"impl": "reference", "metric": "sssp_trees_per_s", "value": value, "unit": "trees/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * cores * per_core / value,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": sssp_workload_name(args.sssp_grid, len(s), V), "sample": sample},
            "cpu_baseline": {"value": value, "unit": "trees/s", "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "trees/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    if not args.no_tick:
        def router(rs, re_, rw, roots):
            og = osssp.OracleGraph(rs, re_, rw)
            return og.sssp_batch(roots.astype(np.int32), want_dist=False)[1]
        flat = synth.grid_scenario(args.tick_grid, args.tick_agents, horizon=TICK_HORIZON, seed=1, router=router)
        rate, wall, c = run_tick_cpu(flat, args.cpu_window)
        line["tick"] = {"metric": "agent_steps_per_s", "value": rate, "unit": "agent-steps/s",
                        "cpu_baseline": {"value": rate, "unit": "agent-steps/s", "cores": 1, "kind": "port",
                                         "sample": f"ticks 0..{args.cpu_window - 1} of config 4 ({wall:.1f} s)"}}
    print(json.dumps(line), flush=True)


def run_sweep(args, torch, dist, g, V, E, rank, world, peak):
    """BASELINE config 5 as the sweep it is named as: B = 1k ... 64k TOTAL origins, split over the ranks (strong scaling:
    rank r computes shard_bounds(B, r, world)), full trees + predecessors per step, tables gathered (every rank, or
    --gather-to).  One point per B: trees/s over all ranks (max over ranks of the device time), which kernel ran."""
    from spatial_queue_b200 import shard
    points = []
    bytes_per_tree = 20 * E + 24 * V
    compact = world > 1 and not args.no_gather
    g.set_parent_format(1 if compact else 0)
    pdt = torch.uint8 if compact else torch.int32
    for Btot in (1024, 2048, 4096, 8192, 16384, 32768, 65536):
        lo, hi = shard.shard_bounds(Btot, rank, world)
        n = hi - lo
        rows = max(shard.shard_bounds(Btot, r, world)[1] - shard.shard_bounds(Btot, r, world)[0] for r in range(world))
        perm = np.random.default_rng(4321).permutation(V)
        o_dev = torch.from_numpy(perm[(np.arange(lo, hi)) % V].astype(np.int32)).cuda()
        dist_dev = torch.empty((n, V), dtype=torch.float64, device="cuda")
        par_dev = torch.empty((rows, V), dtype=pdt, device="cuda")
        gathered = None
        if compact:
            gathered = torch.empty((world * rows, V), dtype=pdt, device="cuda") if (args.gather_to < 0 or rank == args.gather_to) else None

        def step():
            g.sssp_batch_device(o_dev, dist_dev, par_dev[:n])
            if compact:
                if args.gather_to < 0:
                    dist.all_gather_into_tensor(gathered, par_dev)
                else:
                    parts = list(gathered.split(rows)) if rank == args.gather_to else None
                    dist.gather(par_dev, parts, dst=args.gather_to)

        for _ in range(max(3, args.warmup)):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        if world > 1:
            tm = torch.tensor([ms], device="cuda")
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            ms = float(tm.item())
        tps = Btot / (ms / 1e3)
        n_sm = torch.cuda.get_device_properties(0).multi_processor_count
        kernel = "sssp_warp_kernel" if n >= 13 * n_sm else "sssp_ell_kernel"
        points.append({"total_origins": Btot, "origins_per_gpu": n, "ms_per_step": ms, "trees_per_s": tps,
                       "frac_of_one_gpu_roofline": tps * bytes_per_tree / (peak * 1e9) / world, "kernel": kernel})
        del dist_dev, par_dev, gathered
        torch.cuda.empty_cache()
    g.set_parent_format(0)
    return points


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=18944, help="SSSP origins per GPU per step (default: 8 waves of 16 warps on 148 SMs)")
    ap.add_argument("--sssp-grid", type=int, default=SSSP_GRID)
    ap.add_argument("--tick-grid", type=int, default=TICK_GRID)
    ap.add_argument("--tick-agents", type=int, default=TICK_AGENTS)
    ap.add_argument("--tick-ticks", type=int, default=TICK_TICKS)
    ap.add_argument("--tick-reps", type=int, default=1)
    ap.add_argument("--cpu-window", type=int, default=1800, help="ticks of config 4 the CPU tick baseline runs")
    ap.add_argument("--ref-seconds", type=float, default=10.0, help="CPU seconds per reference-arm step")
    ap.add_argument("--no-tick", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-gather", action="store_true")
    ap.add_argument("--nccl-gather", action="store_true", help="N > 1: NCCL all-gather instead of copy-engine pushes")
    ap.add_argument("--chunks", type=int, default=4, help="N > 1: pieces the batch is cut into (gather of one overlaps the trees of the next)")
    ap.add_argument("--gather-to", type=int, default=-1, help="N > 1: push the tables to this rank only (the simulating GPU) instead of to every rank")
    ap.add_argument("--sweep", action="store_true", help="config 5 as a sweep: 1k..64k total origins split over the ranks (strong scaling)")
    ap.add_argument("--no-configs23", action="store_true", help="skip the config 2 / 3 tick objects")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))

    if args.impl == "reference":
        reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    numa_note = bind_to_gpu_numa_node(torch, local) if world > 1 else None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from spatial_queue_b200 import _lib, sp_interface as spi
    peak, peak_src = measured_peaks()

    # ---- HP-2 -------------------------------------------------------------------------------------------------
    s, e, w, V = sssp_workload(args.sssp_grid)
    E = len(s)
    g = spi.from_arrays(s, e, w, max_vertex_id=V - 1)
    g.sync()
    B = args.batch
    bytes_per_tree = 20 * E + 24 * V
    if args.sweep:
        pts = run_sweep(args, torch, dist, g, V, E, rank, world, peak)
        if rank == 0:
            print(json.dumps({"metric": "sssp_trees_per_s", "unit": "trees/s", "n_gpus": world, "scaling": "strong",
                              "config": {"workload": sssp_workload_name(args.sssp_grid, E, V),
                                         "gather": "none" if world == 1 or args.no_gather else
                                         ("NCCL gather to rank %d" % args.gather_to if args.gather_to >= 0 else "NCCL all-gather on every rank")},
                              "steps": args.steps, "sweep": pts}), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return
    dist_dev = torch.empty((B, V), dtype=torch.float64, device="cuda")
    # N > 1: the int32 predecessor (next-hop) tables of every rank are all-gathered to every rank.  The batch is
    # cut into chunks; the gather of chunk k runs on a side stream while chunk k+1 is computed, and the tables
    # are double-buffered across steps so that a step never waits for the previous step's last gather.
    n_chunks = args.chunks if (world > 1 and not args.no_gather) else 1
    assert B % n_chunks == 0, "--batch must be a multiple of --chunks for N > 1"
    cb = B // n_chunks
    # N > 1 ships the tables as one-byte in-edge codes (sp_set_parent_format 1): every rank holds the graph, so
    # the code of the predecessor edge is all the simulating GPU needs; N = 1 writes plain int32 vertex ids
    compact = world > 1 and not args.no_gather
    pdt = torch.uint8 if compact else torch.int32
    if compact:
        g.set_parent_format(1)
    par_bufs = [torch.empty((B, V), dtype=pdt, device="cuda") for _ in range(2 if n_chunks > 1 else 1)]
    gathered = None
    comm_stream = None
    peer = None
    gather_kind = None
    if world > 1 and not args.no_gather:
        if not args.nccl_gather:
            try:
                from spatial_queue_b200.shard import PeerTables
                peer = PeerTables(n_chunks, cb, V, pdt)
                gather_kind = "copy-engine pushes into symmetric memory" + (f", to rank {args.gather_to} only" if args.gather_to >= 0 else "")
            except Exception as ex:       # no symmetric memory on this box: NCCL collective instead
                if rank == 0:
                    print(f"symmetric memory unavailable ({type(ex).__name__}: {ex}); using the NCCL all-gather", file=sys.stderr)
                peer = None
        if peer is None:
            gathered = [torch.empty((world * cb, V), dtype=pdt, device="cuda") for _ in range(n_chunks)]
            comm_stream = torch.cuda.Stream()
            gather_kind = "NCCL all_gather_into_tensor"
        buf_free = [None, None]      # event: all gathers reading par_bufs[k] are done
    origins_dev = [torch.from_numpy(origins_for(rank, B, V, step=i)).cuda() for i in range(args.warmup + args.steps)]
    launches0 = _lib.lib().sp_kernel_launches()

    def step(i, ev=None):
        par_dev = par_bufs[i % len(par_bufs)]
        cur = torch.cuda.current_stream()
        gathering = gathered is not None or peer is not None
        if gathering and buf_free[i % 2] is not None:
            cur.wait_event(buf_free[i % 2])
        if ev:
            ev[0].record()
        for c in range(n_chunks):
            sl = slice(c * cb, (c + 1) * cb if c + 1 < n_chunks else B)
            g.sssp_batch_device(origins_dev[i][sl], dist_dev[sl], par_dev[sl])
            if gathering:
                done = torch.cuda.Event()
                done.record(cur)
                if peer is not None:
                    peer.push(c, par_dev[sl], after=done, to=args.gather_to if args.gather_to >= 0 else None)
                else:
                    comm_stream.wait_event(done)
                    with torch.cuda.stream(comm_stream):
                        dist.all_gather_into_tensor(gathered[c], par_dev[sl])
        if ev:
            ev[1].record()
        if gathering:
            if peer is not None:
                e = peer.fence()         # every rank's rows of this step are in every rank's table
            else:
                e = torch.cuda.Event()
                e.record(comm_stream)
            buf_free[i % 2] = e
            if i == args.warmup + args.steps - 1 or i == args.warmup - 1:
                cur.wait_event(e)      # the timed region ends when the last rows have landed

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    t_ev0, t_ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_ev0.record()
    for i in range(args.steps):
        step(args.warmup + i, kev[i])
    t_ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms_total = t_ev0.elapsed_time(t_ev1)
    if world > 1:
        tmax = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms_total = float(tmax.item())
    kernel_ms = float(np.mean([a.elapsed_time(b) for a, b in kev]))
    launches = _lib.lib().sp_kernel_launches() - launches0 - 0
    value = world * B * args.steps / (ms_total / 1e3)

    gather_ok = None
    if gather_kind is not None:
        # outside the timed region: every rank's block of the gathered table must be what that rank computed
        last = par_bufs[(args.warmup + args.steps - 1) % len(par_bufs)]
        mine = torch.stack([last[c * cb:(c + 1) * cb].to(torch.int64).sum() for c in range(n_chunks)])
        sums = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(sums, mine)
        ok = True
        if args.gather_to < 0 or peer is None or rank == args.gather_to:
            for c in range(n_chunks):
                tbl = peer.table(c) if peer is not None else gathered[c]
                for r in range(world):
                    ok = ok and int(tbl[r * cb:(r + 1) * cb].to(torch.int64).sum().item()) == int(sums[r][c].item())
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        gather_ok = bool(flag.item())
    if compact:
        g.set_parent_format(0)
    # ---- e2e through the host-buffer C ABI: host origins in, host next-hop tables out ---------------------------------
    # `e2e` = sp_sssp_batch_codes4: the whole batch, predecessors as four-bit in-edge codes in page-locked host memory
    # (25 KB per tree; a consumer holding the graph follows the codes directly, sp_decode_parents_host4 gives vertex ids).
    # `e2e_codes8` = sp_sssp_batch_codes, the one-byte form (50 KB per tree; any in-degree up to 254).
    # `e2e_full` = sp_sssp_batch: f64 distances + int32 predecessors (602 KB per tree: PCIe-bound), 4,096 trees.
    e2e = e2e_full = e2e_codes8 = None
    if rank == 0 or world > 1:
        oh = origins_for(rank, B, V, step=7)
        reps = 2

        def timed_codes(call, width, what):
            ch = spi.pinned_empty((B, width), np.uint8)
            call(oh[:2368], out=(None, ch[:2368]))
            call(oh, out=(None, ch))
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            for _ in range(reps):
                call(oh, out=(None, ch))
            wall = time.perf_counter() - t0
            if world > 1:
                tw = torch.tensor([wall], device="cuda")
                dist.all_reduce(tw, op=dist.ReduceOp.MAX)
                wall = float(tw.item())
            return {"value": world * B * reps / wall, "unit": "trees/s", "h2d_bytes_per_step": int(oh.nbytes),
                    "d2h_bytes_per_step": int(ch.nbytes), "trees_per_call": B, "call": what}

        e2e = timed_codes(g.sssp_batch_codes4, (V + 1) // 2,
                          "sp_sssp_batch_codes4 (host origins in, four-bit predecessor codes in page-locked host memory out)")
        e2e_codes8 = timed_codes(g.sssp_batch_codes, V,
                                 "sp_sssp_batch_codes (host origins in, one-byte predecessor codes in page-locked host memory out)")
        eb = min(B, 4096)
        dh, ph = spi.pinned_empty((eb, V), np.float64), spi.pinned_empty((eb, V), np.int32)
        g.sssp_batch(oh[:256], out=(dh[:256], ph[:256]))
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            g.sssp_batch(oh[:eb], out=(dh, ph))
        wall = time.perf_counter() - t0
        if world > 1:
            tw = torch.tensor([wall], device="cuda")
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
            wall = float(tw.item())
        e2e_full = {"value": world * eb * reps / wall, "unit": "trees/s", "h2d_bytes_per_step": int(oh[:eb].nbytes),
                    "d2h_bytes_per_step": int(dh.nbytes + ph.nbytes), "trees_per_call": eb,
                    "call": "sp_sssp_batch (f64 distances + int32 predecessors, 602 KB per tree over PCIe)"}
        del dh, ph

    # ---- e2e of the call the reference's agents make: get_path = distance(destin) + route(destin) per agent ----------
    # (sp_route_batch: host origins + destinations in, host distances + paths out; one tree per origin on the device)
    e2e_routes = None
    if rank == 0 or world > 1:
        rb = B
        orr = origins_for(rank, rb, V, step=11)
        drr = np.random.default_rng(100 + rank).integers(0, V, rb).astype(np.int32)
        g.route_batch(orr[:512], drr[:512])
        g.route_batch(orr, drr)                       # sizes the chunk buffers and the page-locked path pool
        t0 = time.perf_counter()
        reps = 2
        for _ in range(reps):
            rd, roff, rlen, rpool = g.route_batch(orr, drr)
        wall = time.perf_counter() - t0
        if world > 1:
            tw = torch.tensor([wall], device="cuda")
            dist.all_reduce(tw, op=dist.ReduceOp.MAX)
            wall = float(tw.item())
        e2e_routes = {"value": world * rb * reps / wall, "unit": "trees/s", "call": "sp_route_batch (one tree per agent, "
                      "distance + path returned)", "h2d_bytes_per_step": int(orr.nbytes + drr.nbytes),
                      "d2h_bytes_per_step": int(rd.nbytes + roff.nbytes + rlen.nbytes + rpool.nbytes),
                      "trees_per_call": rb, "mean_path_vertices": float(rlen.mean())}

    clocks = sampler.stop() if rank == 0 else None
    line = None
    if rank == 0:
        achieved = bytes_per_tree * B / (kernel_ms / 1e3) / 1e9
        tr = measured_traffic("sssp")
        line = {"metric": "sssp_trees_per_s", "value": value, "unit": "trees/s", "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": sssp_workload_name(args.sssp_grid, E, V),
                           "batch_per_gpu": B,
                           "gather": (f"one-byte next-hop tables in {n_chunks} chunks ({gather_kind}), overlapped with the next "
                                      f"chunk's trees") if gather_kind else None,
                           "gather_verified": gather_ok,
                           "e2e_call": "sp_sssp_batch_codes4: host origins in, four-bit predecessor codes (25 KB per tree) in "
                                       "page-locked host memory out; e2e_codes8 = one-byte codes; e2e_full = f64 distances + "
                                       "int32 predecessors",
                           "host_numa": numa_note,
                           "l2": f"outputs of one step ({(12 * B * V) >> 20} MiB) exceed L2; no flush needed",
                           "metric_of": "BASELINE.json: agent-steps/s per sim tick (object 'tick') + reroute SSSP trees/s (top level)"},
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": tr["bytes_per_tree"] * B if tr else None,
                             "traffic_source": tr["source"] if tr else None,
                             "kernel": "tree kernel + predecessor / un-interleave pass, both inside kernel_ms", "kernel_ms": kernel_ms,
                             "bytes_per_tree": bytes_per_tree, "peak_source": peak_src},
                "e2e": e2e, "e2e_codes8": e2e_codes8, "e2e_full": e2e_full, "e2e_routes": e2e_routes, "gpu_launches": int(launches), "clocks": clocks}

    # ---- HP-1 + CPU baselines: single GPU, rank 0 ----------------------------------------------------------------
    if rank == 0 and world == 1:
        del dist_dev, par_bufs
        torch.cuda.empty_cache()
        if not args.no_cpu:
            from oracle import sssp as osssp
            cores = os.cpu_count() or 1
            if osssp.ref_available():
                per_core, per_tree = calibrate_ref_trees(s, e, w, V, target_s=args.ref_seconds)
                tps = sssp_reference_trees_per_s(s, e, w, V, per_core, cores)
                line["cpu_baseline"] = {"value": tps, "unit": "trees/s", "cores": cores, "kind": "reference",
                                        "sample": f"{per_core} full trees per core x {cores} processes, same graph "
                                                  f"(oracle/_ref/liblsp.so; one core: {1 / per_tree:.1f} trees/s)"}
            else:
                og = osssp.OracleGraph(s, e, w)
                t1 = time.perf_counter()
                og.sssp_batch(origins_for(0, 256, V, step=99))
                line["cpu_baseline"] = {"value": 256 / (time.perf_counter() - t1), "unit": "trees/s", "cores": 1,
                                        "kind": "port", "sample": "256 full trees, oracle/sssp_oracle.c"}
        if not args.no_tick:
            for case in tick_cases(args):
                l0 = _lib.lib().sq_kernel_launches()
                tick, flat = run_tick_gpu(case, args, torch, peak)
                tick_line = {"metric": "agent_steps_per_s", "value": tick["value"], "unit": "agent-steps/s",
                             "ms_per_tick": tick["ms_per_tick"], "dtype": "i32 state + f64 capacities",
                             "config": {"workload": f"{case['name']}, {len(flat.link_end_node)} links (source links included), "
                                                    f"{len(flat.agent_key)} agents, {case['ticks']} ticks; {case['what']}"},
                             "counters": tick["counters"], "arrived": tick["arrived"], "stopped": tick["stopped"],
                             "roofline": tick["roofline"], "e2e": tick["e2e"],
                             "gpu_launches": int(_lib.lib().sq_kernel_launches() - l0)}
                if not args.no_cpu:
                    # the C port replays the tick only (the reference's per-agent Dijkstra would come on top)
                    flat_cpu = case["flat_cpu"]() if case["flat_cpu"] else flat
                    rate, wall, c = run_tick_cpu(flat_cpu, tick["window_ticks"])
                    tick_line["cpu_baseline"] = {
                        "value": rate, "unit": "agent-steps/s", "cores": 1, "kind": "port",
                        "sample": f"ticks 0..{tick['window_ticks'] - 1} of the same network and demand, {case['cpu_note']} "
                                  f"({wall:.1f} s, oracle/tick_oracle.c){python_reference_note()}",
                        # the device's own agent-steps over the same ticks / its own time for them
                        "gpu_same_window": tick["window_agent_steps"] / (tick["window_ms"] / 1e3),
                        "gpu_window_agent_steps": tick["window_agent_steps"], "cpu_window_agent_steps": c["agent_steps"]}
                line[case["key"]] = tick_line
                line["gpu_launches"] += tick_line["gpu_launches"]
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
