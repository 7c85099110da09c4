#!/usr/bin/env python3
"""Benchmark of the daily transmission cycle: agent-days/s on synthetic populations (BASELINE.json).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" is DAYS_PER_STEP consecutive simulated days (Community::tick + daily importation) of the
Miami-Dade-scale configuration: ~2.7 M synthetic agents, wild type + alpha + delta entering inside
the window, the active-vac parameter set (covid-abm_b200/params.py). Every rank prints nothing; rank 0
prints ONE JSON line.

  value      device-timed whole-job agent-days/s with all state resident in HBM (CUDA-graph replay of the
             day pipeline, CUDA events on the library's stream, max over ranks)
  e2e        the same days driven the way simulator.h drives them: one cabm_tick + cabm_import +
             ledger-row read per day through the C ABI with host buffers (H2D of the day context,
             D2H of the ledger row and import count inside the timed region)
  roofline   dominant kernel of the pipeline: algorithmic bytes (SURVEY.md 8d byte model, counted on
             device) / its CUDA-event time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the unmodified reference (oracle/_ref/ref_driver) on one host core: warmed untimed to the
             first timed day of the GPU arm, then DAYS_PER_STEP days of the same workload timed
  dropin     (N = 1) the reference's unmodified simulate_epidemic (simulator.h:265-533) on the C++ facade over
             the CUDA library (covid-abm_b200/facade/_build/facade_driver_cuda), same population: wall-clock
             ms per simulated day without and with a vaccination campaign -- what a maintainer who swaps
             Community.cpp for the facade gets

With --gpus N > 1 (torchrun, one rank per GPU) the headline stays what it was: each rank advances an
independent replicate of the same population (seed 1 + rank) -- the reference's own scaling mode (SLURM
arrays of replicates); no collective sits on the data path ("scaling": "weak"). Two more records are added
for the multi-GPU cases BASELINE.json names:
  sharded    configs[2]: ONE Florida-scale community (21.5 M agents) advanced by the N ranks together through
             cabm_shard_run_days (NCCL + peer memory inside the library): us per day, speed-up over one handle
             on rank 0's GPU, whether the summed ledger equals the single handle's, per-exchange device time
  batch      configs[4]: 256 communities of 100 k agents, 256 / N per GPU through cabm_batch_run_days
--impl reference times the unmodified reference on the host cores over THE SAME days as the GPU arm:
W x DAYS_PER_STEP warm-up days untimed, then K steps of DAYS_PER_STEP days ("same_config": true).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))

DAYS_PER_STEP = 10
N_AGENTS = int(os.environ.get("CABM_BENCH_AGENTS", 2_700_000))
MUTANTS = ("2020-03-21", "2020-05-20", "2030-01-01")          # alpha ~day 40, delta ~day 100 (start 2020-02-10)
REF_DRIVER = os.path.join(ROOT, "oracle", "_ref", "ref_driver")
BYTES = dict(agents=0, active=40, flags=12, patrons=16.6, visits=4, hot_visits=20, hot_locs=28, nbr_edges=8, trials=10, infections=64)
# kernels of the day pipeline and the byte-model terms each one moves (DESIGN.md "kernels and byte model"); the gate and the
# two apply launches share the per-infection term
KERNEL_GROUPS = [
    ("k_scan+k_active", ("k_scan", "k_active"), ("active", "flags")),
    ("k_visit", ("k_visit",), ("patrons", "visits", "hot_visits")),
    ("k_hazard", ("k_hazard",), ("hot_locs", "nbr_edges")),
    ("k_trials+k_hosp_import", ("k_trials", "k_hosp_import"), ("trials",)),
    ("k_gate+k_apply_status", ("k_gate", "k_apply_status", "k_apply(phase1)"), ("infections",)),
]


def rank_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler(threading.Thread):
    """nvidia-smi SM clock / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True); self.index = index; self.rows = []; self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        import statistics
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def make_workload(total_days, seed):
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(N_AGENTS)
    P = cabm.params.default_params(run_length=total_days, seed=seed, mutant_intro_dates=MUTANTS)
    T = cabm.params.build_tables(P)
    return cabm, pop, P, T


def algorithmic_bytes(work: dict, terms=None) -> float:
    return sum(BYTES[k] * work[k] for k in (terms or BYTES))


def run_reference_sample(pop, P, T, warm_days, timed_days, procs=1):
    """Time the unmodified reference on the same population: `procs` concurrent single-threaded replicates."""
    import covid_abm_b200 as cabm
    import synthpop
    if not os.path.exists(REF_DRIVER):
        return None
    tmp = tempfile.mkdtemp(prefix="cabm_ref_")
    synthpop.write_reference_files(pop, tmp, "bench")
    jobs = []
    for r in range(procs):
        Pr = dict(P); Pr["seed"] = int(P["seed"]) + r
        pfile = os.path.join(tmp, f"params{r}.txt")
        cabm.params.write_text(Pr, T, pfile)
        jobs.append(subprocess.Popen([REF_DRIVER, "run", pfile, tmp, "bench", os.path.join(tmp, f"out{r}.csv"),
                                      str(warm_days + timed_days), str(warm_days)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True))
    outs = []
    for j in jobs:
        o, _ = j.communicate()
        outs.append(json.loads(o.strip().splitlines()[-1]))
    agent_days = sum(o["agents"] * o["days"] for o in outs)
    secs = max(o["tick_import_s"] for o in outs)
    subprocess.run(["rm", "-rf", tmp])
    return {"value": agent_days / secs, "seconds": secs, "agent_days": agent_days, "procs": procs}


FACADE_DRIVER = os.path.join(ROOT, "covid-abm_b200", "facade", "_build", "facade_driver_cuda")
DROPIN_DAYS = 120
SHARD_AGENTS = int(os.environ.get("CABM_BENCH_SHARD_AGENTS", 21_500_000))
SHARD_WARM, SHARD_DAYS = 30, 60
BATCH_REPLICATES, BATCH_AGENTS, BATCH_DAYS = 256, 100_000, 200


def dropin_record(pop, P0):
    """The reference's unmodified simulate_epidemic on the facade over the CUDA library: ms per simulated day, wall clock."""
    import covid_abm_b200 as cabm
    import synthpop
    if not os.path.exists(FACADE_DRIVER):
        return {"unavailable": "covid-abm_b200/facade/_build/facade_driver_cuda not built (needs the reference sources at build time)"}
    P = cabm.params.default_params(run_length=DROPIN_DAYS, seed=int(P0["seed"]), mutant_intro_dates=MUTANTS)
    T = cabm.params.build_tables(P)
    tmp = tempfile.mkdtemp(prefix="cabm_dropin_")
    out = {"boundary": "unmodified simulator.h simulate_epidemic -> facade Community -> C ABI -> libcabm.so", "days": DROPIN_DAYS,
           "agents": int(pop["n_people"])}
    try:
        synthpop.write_reference_files(pop, tmp, "b")
        cabm.params.write_text(P, T, os.path.join(tmp, "params.txt"))
        for campaign, key in ((0, "ms_per_day"), (1, "ms_per_day_campaign"), (2, "ms_per_day_campaign_ring")):
            r = subprocess.run([FACADE_DRIVER, "sim", os.path.join(tmp, "params.txt"), tmp, "b", os.path.join(tmp, f"o{campaign}"), str(campaign)],
                               capture_output=True, text=True, timeout=900)
            if r.returncode != 0:
                out[key] = None; out["error"] = r.stderr[-300:]
                continue
            j = json.loads(r.stdout.strip().splitlines()[-1])
            out[key] = j["simulate_epidemic_s"] * 1e3 / j["days"]
            if campaign == 0:
                out.update(load_s=j["load_s"], upload_s=j.get("upload_s"), agent_days_per_s=j["agent_days_per_s"], infections=j["infections"],
                           note="ms_per_day = wall clock of simulate_epidemic / days; text-file load and the one-off device upload are load_s and upload_s")
    except Exception as e:                                  # a record, not the headline: never fail the bench line for it
        out["error"] = repr(e)[:300]
    subprocess.run(["rm", "-rf", tmp])
    return out


def sharded_record(dist, dev, rank, world):
    """configs[2]: one Florida-scale community over the ranks of the job (cabm_shard_run_days), against one handle on rank 0."""
    import numpy as np
    import torch
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(SHARD_AGENTS)
    n = int(pop["n_people"]); total = SHARD_WARM + SHARD_DAYS
    P = cabm.params.default_params(run_length=total, seed=1, mutant_intro_dates=MUTANTS, prob_daily_exposure=2e-4)   # exp/active-vac/main.cpp:248-249 at this scale
    T = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(n, P["seed"])
    cuts = cabm.sharded.household_cuts(pop, world)
    shard = cabm.sharded.Shard(P, pop, T, rank, world, cuts, attrs=attrs, device=dev)
    shard.nccl_init(dist)
    shard.run_days_nccl(SHARD_WARM)
    snap = shard.snapshot()
    shard.device_time_ms()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    shard.run_days_nccl(SHARD_DAYS)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    t = torch.tensor([wall, shard.device_time_ms()[0]], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    led = torch.from_numpy(shard.ledger_table().astype(np.int64)).to(dev); dist.all_reduce(led)
    out = {"workload": f"configs[2]: one community of {n} synthetic agents, WT+alpha+delta, days {SHARD_WARM}-{SHARD_WARM + SHARD_DAYS - 1}, sharded over {world} GPUs",
           "agents": n, "days": SHARD_DAYS, "exchange_a": "peer memory (NVLink loads)" if shard.peer_memory() else "nccl all-gather",
           "us_per_day": float(t[0]) * 1e6 / SHARD_DAYS, "us_per_day_device": float(t[1]) * 1e3 / SHARD_DAYS,
           "agent_days_per_s": n * SHARD_DAYS / float(t[0]), "timing": "wall clock around cabm_shard_run_days between barriers, max over ranks",
           "infections": int(led[: SHARD_WARM + SHARD_DAYS, 0].sum())}
    shard.restore(snap)                                                # per-kernel / per-exchange device time of rank 0: the same days again, untimed
    shard.set_profile(True); shard.kernel_times()
    shard.run_days_nccl(SHARD_DAYS)
    kt = shard.kernel_times(); shard.set_profile(False)
    out["rank0_us_per_day"] = {k: round(v[0] * 1e3 / SHARD_DAYS, 2) for k, v in kt.items() if v[1]}
    led2 = torch.from_numpy(shard.ledger_table().astype(np.int64)).to(dev); dist.all_reduce(led2)
    out["profiled_pass_equals_timed_pass"] = bool(torch.equal(led, led2))
    shard.snap_free(snap); shard.close()
    if rank == 0:
        single = cabm.Community(P, pop, T, attrs=attrs)
        single.run_days(SHARD_WARM); single.device_time_ms()
        torch.cuda.synchronize(); t1 = time.perf_counter()
        single.run_days(SHARD_DAYS)
        torch.cuda.synchronize(); w1 = time.perf_counter() - t1
        ls = single.ledger_table().astype(np.int64)[: SHARD_WARM + SHARD_DAYS]
        out["single_handle_us_per_day"] = w1 * 1e6 / SHARD_DAYS
        out["single_handle_us_per_day_device"] = single.device_time_ms()[0] * 1e3 / SHARD_DAYS
        out["speedup"] = out["single_handle_us_per_day"] / out["us_per_day"]
        out["equals_single_handle"] = bool(np.array_equal(ls, led.cpu().numpy()[: SHARD_WARM + SHARD_DAYS]))
        single.close()
    dist.barrier()
    return out


def batch_record(dist, dev, rank, world):
    """configs[4]: 256 communities of 100 k agents, 256 / world per GPU, one launch per kernel per batch (cabm_batch_run_days)."""
    import numpy as np
    import torch
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(BATCH_AGENTS)
    n = int(pop["n_people"])
    mine = cabm.replicates.replicate_slice(BATCH_REPLICATES, rank, world)
    rng = np.random.default_rng(2024)
    Ts = rng.uniform(0.023, 0.043, size=BATCH_REPLICATES)              # the prior of the ABC sweeps (exp/abc/abc_covid.json:14-18)
    comms = []
    for r in mine:
        P = cabm.params.default_params(run_length=BATCH_DAYS, seed=1 + r, T=float(Ts[r]), mutant_intro_dates=("2020-03-01", "2020-03-21", "2020-04-10"))
        comms.append(cabm.Community(P, pop, cabm.params.build_tables(P)))
    cabm.replicates.batch_run_days(comms, 20)
    comms[0].device_time_ms()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    cabm.replicates.batch_run_days(comms, BATCH_DAYS - 20)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    inf = sum(int(c.ledger_table()[:, 0].sum()) for c in comms)
    t = torch.tensor([wall, comms[0].device_time_ms()[0] * 1e-3, float(inf)], dtype=torch.float64, device=dev)
    tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX); dist.all_reduce(t, op=dist.ReduceOp.SUM)
    for c in comms:
        c.close()
    days = BATCH_DAYS - 20
    return {"workload": f"configs[4]: {BATCH_REPLICATES} communities of {n} agents (seeds 1..{BATCH_REPLICATES}, T ~ U(0.023, 0.043)), {len(mine)} per GPU, days 20-{BATCH_DAYS - 1}",
            "replicates": BATCH_REPLICATES, "per_gpu": len(mine), "days": days, "seconds": float(tmax[0]), "device_seconds": float(tmax[1]),
            "agent_days_per_s": BATCH_REPLICATES * n * days / float(tmax[0]), "us_per_replicate_day": float(tmax[0]) * 1e6 / (len(mine) * days),
            "infections": int(t[2]), "timing": "wall clock around cabm_batch_run_days between barriers, max over ranks"}


def bench_ours(args):
    rank, local_rank, world = rank_env()
    import numpy as np
    import torch
    dist = None
    dev_index = local_rank if world > 1 else 0
    dev = torch.device("cuda", dev_index)
    torch.cuda.set_device(dev)            # every GPU stays visible (the sharded record maps peer memory); cabm_create binds a handle to the current device
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, args.warmup
    total_days = (K + W) * DAYS_PER_STEP
    import covid_abm_b200 as _c
    cabm, pop, P, T = make_workload(total_days, seed=_c.replicates.replicate_seed(1, rank, 0, world))
    n = int(pop["n_people"])
    attrs = cabm.abi.immunity_attributes(n, P["seed"])
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident run: CUDA-graph replay, DAYS_PER_STEP days per call ----
    A = cabm.Community(P, pop, T, attrs=attrs)
    for _ in range(W):
        A.run_days(DAYS_PER_STEP)
    snap = A.snapshot(); work0 = A.work_counters(); A.device_time_ms()
    sampler = ClockSampler(dev_index); sampler.start()
    barrier(); t0 = time.perf_counter(); dev_ms = 0.0; launches = 0
    for _ in range(K):
        flush.fill_(1); torch.cuda.synchronize()
        A.run_days(DAYS_PER_STEP)
        ms, l = A.device_time_ms(); dev_ms += ms; launches += l
    barrier(); wall = time.perf_counter() - t0
    sampler.stop_flag = True; sampler.join(timeout=2)
    work1 = A.work_counters()
    work = {k: work1[k] - work0[k] for k in work1}
    ledger = A.ledger_table()

    # ---- per-kernel CUDA-event timing of the same days (stream launches, one event per kernel) ----
    A.restore(snap); A.set_profile(True); A.kernel_times()
    A.run_days(K * DAYS_PER_STEP)
    ktimes = A.kernel_times(); A.set_profile(False)
    assert np.array_equal(ledger, A.ledger_table()), "profiled replay diverged from the timed run"
    A.snap_free(snap); A.close()

    # ---- end to end through the per-day C-ABI calls with host buffers ----
    B = cabm.Community(P, pop, T, attrs=attrs)
    row = np.zeros(cabm.abi.N_COUNTERS, dtype=np.uint64)
    import ctypes as C

    def day_by_day(days):
        for _ in range(days):
            B.step()
            B._check(B.fn.ledger_row(B._h, B.day - 1, row.ctypes.data_as(C.POINTER(C.c_uint64))))
    day_by_day(W * DAYS_PER_STEP)
    barrier(); t1 = time.perf_counter()
    day_by_day(K * DAYS_PER_STEP)
    barrier(); e2e_wall = time.perf_counter() - t1
    assert np.array_equal(ledger, B.ledger_table()), "per-day path diverged from run_days"
    B.close()

    # ---- reduce over ranks ----
    agent_days = float(n) * K * DAYS_PER_STEP
    agent_days_all, dev_s_max, _ = cabm.replicates.job_throughput(dist, dev, agent_days, dev_ms * 1e-3)
    _, wall_max, _ = cabm.replicates.job_throughput(dist, dev, agent_days, wall)
    _, e2e_max, _ = cabm.replicates.job_throughput(dist, dev, agent_days, e2e_wall)
    dev_ms_max = dev_s_max * 1e3
    # ---- the multi-GPU cases BASELINE.json names beside the replica headline: one sharded community, the calibration batch ----
    sharded = batch = None
    if world > 1 and not args.no_multi:
        del flush; torch.cuda.empty_cache()
        try:
            sharded = sharded_record(dist, dev, rank, world)
        except Exception as e:
            sharded = {"error": repr(e)[:400]}
        try:
            batch = batch_record(dist, dev, rank, world)
        except Exception as e:
            batch = {"error": repr(e)[:400]}
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    # dominant kernel group: largest total CUDA-event time among the groups the byte model covers
    def group_ms(g):
        return sum(ktimes.get(k, (0.0, 0))[0] for k in g[1])
    # the dominant KERNEL: the single launch with the largest CUDA-event time per day (a group of several small launches is
    # not a kernel); its byte-model terms are those of the group it belongs to
    top_kernel = max(ktimes, key=lambda k: ktimes[k][0] / max(ktimes[k][1], 1))
    dom_g = next(g for g in KERNEL_GROUPS if top_kernel in g[1])
    if len(dom_g[1]) > 1:
        dom_g = (top_kernel, (top_kernel,), dom_g[2])
    dom = dom_g[0]
    dom_ms = group_ms(dom_g); dom_n = max(ktimes.get(dom_g[1][0], (0.0, 0))[1], 1)
    dom_bytes = algorithmic_bytes(work, dom_g[2])
    achieved = dom_bytes / (dom_ms * 1e-3) / 1e9 if dom_ms > 0 else 0.0
    groups = {g[0]: {"us_per_day": round(group_ms(g) * 1e3 / (K * DAYS_PER_STEP), 3),
                     "GBps": round(algorithmic_bytes(work, g[2]) / (group_ms(g) * 1e-3) / 1e9, 1) if group_ms(g) > 0 else None}
              for g in KERNEL_GROUPS}
    total_k_ms = sum(v[0] for v in ktimes.values())
    traffic = traffic_note = None
    try:                                  # dram bytes per launch of the dominant group from the committed ncu --set full capture
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            cap = json.load(f)
        traffic = float(sum(cap["dram_bytes_per_launch"][k] for k in dom_g[1]))
        cap_alg = algorithmic_bytes(cap["work"], dom_g[2])
        traffic_note = f"{cap['source']}; algorithmic bytes of that launch {cap_alg:.4g} (traffic/algorithmic {traffic / cap_alg:.2f})"
    except Exception:
        pass
    day_bytes = algorithmic_bytes(work)
    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        cpu_warm, cpu_days = W * DAYS_PER_STEP, DAYS_PER_STEP
        r = run_reference_sample(pop, P, T, cpu_warm, cpu_days, procs=1)
        if r:
            cpu = {"value": r["value"], "unit": "agent-days/s", "cores": 1, "kind": "reference",
                   "sample": f"unmodified reference (oracle/_ref/ref_driver, g++ -O2 + GSL shim), same population and parameters, warmed untimed to day {cpu_warm} "
                             f"(the GPU arm's first timed day), then days {cpu_warm}-{cpu_warm + cpu_days - 1} timed, tick+import only"}
    out = {
        "metric": "agent-days/sec", "value": agent_days_all / (dev_ms_max * 1e-3), "unit": "agent-days/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": dev_ms_max / K, "wall_ms_per_step": wall_max * 1e3 / K, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64 fixed-point pressure / f64 hazard / i16 timelines", "data": "synthetic",
        "config": {"workload": f"configs[1]: {n} synthetic agents (tools/synthpop.py seed 20200210), WT+alpha+delta, active-vac parameter set",
                   "days_per_step": DAYS_PER_STEP, "agents": n, "locations": int(pop["n_loc"]), "parallelism": "1 community per GPU" + ("" if world == 1 else f" x {world} replicates"),
                   "l2": "256 MiB flush between steps; consecutive days inside a step reuse L2 as a real run does"},
        # e2e bytes per day: H2D one day context (cabm_tick stages the day, cabm_import of the same day completes it and launches ONE
        # day graph) + the day-index reset; D2H what the block that closes the day writes into pinned host memory: the day's ledger
        # row (read by cabm_ledger_row without another device round trip) and four status words (overflow mask, new infections,
        # day, imported cases)
        "e2e": {"value": agent_days_all / e2e_max, "unit": "agent-days/s", "h2d_bytes_per_step": DAYS_PER_STEP * (88 + 4),
                "d2h_bytes_per_step": DAYS_PER_STEP * (cabm.abi.N_COUNTERS * 8 + 4 * 4), "calls_per_step": DAYS_PER_STEP * 3},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note, "kernel_avg_us": dom_ms * 1e3 / max(dom_n, 1),
                     "kernel_share_of_day": dom_ms / total_k_ms if total_k_ms else None,
                     "pipeline_achieved": day_bytes / (dev_ms_max * 1e-3) / 1e9, "pipeline_frac": day_bytes / (dev_ms_max * 1e-3) / 1e9 / peak,
                     "bytes_per_agent_day": day_bytes / (n * K * DAYS_PER_STEP),
                     "kernels_us_per_day": {k: round(v[0] * 1e3 / (K * DAYS_PER_STEP), 3) for k, v in ktimes.items() if v[1]}, "groups": groups},
        "cpu_baseline": cpu,
        "dropin": dropin_record(pop, P) if (world == 1 and not args.no_dropin) else None,
        "sharded": sharded, "batch": batch,
        "clocks": sampler.summary(),
        "work": work, "epidemic": {"infections": int(ledger[:, 0].sum()), "deaths": int(ledger[:, 4].sum())},
    }
    print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


def bench_reference(args):
    rank, _, world = rank_env()
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    if not os.path.exists(REF_DRIVER):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/ref_driver not built (reference sources absent at build time)"}))
        return
    ref_days_per_step = DAYS_PER_STEP            # the GPU arm's step, over the GPU arm's days
    total_days = (K + W) * DAYS_PER_STEP
    cabm, pop, P, T = make_workload(total_days, seed=1)
    n = int(pop["n_people"])
    cores = os.cpu_count() or 1
    mem_gb = os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES") / 2**30
    procs = int(max(1, min(cores, (mem_gb * 0.5) // (1.0 + 0.8 * n / 1e6), 64)))
    r = run_reference_sample(pop, P, T, W * ref_days_per_step, K * ref_days_per_step, procs=procs)
    sample = (f"{procs} concurrent single-threaded replicates of the unmodified reference (seeds 1..{procs}) on the same population and parameters: "
              f"{W} warm-up steps untimed, then {K} steps of {ref_days_per_step} days (days {W * ref_days_per_step}-{(W + K) * ref_days_per_step - 1}, "
              f"the GPU arm's timed window); tick+import only")
    out = {"impl": "reference", "metric": "agent-days/sec", "value": r["value"], "unit": "agent-days/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
           "ms_per_step": r["seconds"] * 1e3 / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": {"workload": f"configs[1]: {n} synthetic agents (tools/synthpop.py seed 20200210), WT+alpha+delta, active-vac parameter set",
                      "days_per_step": ref_days_per_step, "agents": n, "locations": int(pop["n_loc"]), "parallelism": f"{procs} host processes"},
           "cpu_baseline": {"value": r["value"], "unit": "agent-days/s", "cores": procs, "kind": "reference", "sample": sample},
           "e2e": {"value": r["value"], "unit": "agent-days/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "host_cores": cores, "same_config": True,
           "window": {"first_timed_day": W * ref_days_per_step, "timed_days": K * ref_days_per_step}}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dropin", action="store_true", help="skip the facade (drop-in) record")
    ap.add_argument("--no-multi", action="store_true", help="N > 1: skip the sharded and batch records")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        bench_reference(args)
    else:
        bench_ours(args)


if __name__ == "__main__":
    main()
