#!/usr/bin/env python3
"""One community sharded over the ranks of a torchrun job (one rank per GPU, NCCL), checked against a single-handle run.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/shard_run.py \
        [--agents 2700000] [--days 60] [--check 1]

Every rank loads the whole synthetic population, owns 1/N of the agents (cut at household boundaries) and advances the day in
four parts with the three exchanges of covid-abm_b200/sharded.py over NCCL. Rank 0 prints one JSON line: agent-days/s by wall
clock (barrier + synchronize on both sides, max over ranks), device ms summed per rank, and -- with --check -- whether the
summed ledger equals the same community advanced by one handle on rank 0's GPU (it must, bit for bit).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--agents", type=int, default=2_700_000)
    ap.add_argument("--days", type=int, default=60)
    ap.add_argument("--warm", type=int, default=30)
    ap.add_argument("--check", type=int, default=1)
    ap.add_argument("--profile", type=int, default=0, help="1: also print per-kernel / per-exchange device time of rank 0 (separate untimed pass)")
    ap.add_argument("--native", type=int, default=1, help="1: exchanges inside the library over NCCL (cabm_shard_run_days); 0: torch.distributed between the parts")
    args = ap.parse_args()
    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    import numpy as np
    import torch
    import torch.distributed as dist
    dev = torch.device("cuda", local); torch.cuda.set_device(dev)      # every GPU stays visible: the library maps peer memory
    dist.init_process_group("nccl", device_id=dev)
    import covid_abm_b200 as cabm
    import synthpop
    import bench
    pop = synthpop.generate(args.agents)
    n = int(pop["n_people"]); total = args.warm + args.days
    total += args.days if args.profile else 0
    P = cabm.params.default_params(run_length=total, seed=1, mutant_intro_dates=bench.MUTANTS)
    T = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(n, P["seed"])
    cuts = cabm.sharded.household_cuts(pop, world)
    shard = cabm.sharded.Shard(P, pop, T, rank, world, cuts, attrs=attrs, device=dev)
    c = cabm.sharded.ShardedCommunity([shard], cabm.sharded.DistExchange(shard, dist))
    if args.native:
        shard.nccl_init(dist)
        c.run_days = lambda k: shard.run_days_nccl(k)
    c.run_days(args.warm)
    shard.device_time_ms()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    c.run_days(args.days)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    dev_ms = shard.device_time_ms()[0]
    t = torch.tensor([wall, dev_ms], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    led = torch.from_numpy(c.ledger_table().astype(np.int64)).to(dev); dist.all_reduce(led)
    out = {"world": world, "native_nccl": bool(args.native), "exchange_a": ("peer memory" if shard.peer_memory() else "nccl all-gather") if args.native else "torch all-gather", "agents": n, "days": args.days, "wall_s": float(t[0]), "device_ms_max": float(t[1]),
           "agent_days_per_s_wall": n * args.days / float(t[0]), "agent_days_per_s_device": n * args.days / (float(t[1]) * 1e-3),
           "us_per_day_wall": float(t[0]) * 1e6 / args.days, "infections": int(led[:, 0].sum())}
    if args.profile and args.native:
        shard.set_profile(True); shard.kernel_times()
        c.run_days(args.days)
        kt = shard.kernel_times(); shard.set_profile(False)
        out["rank0_us_per_day"] = {k: round(v[0] * 1e3 / args.days, 2) for k, v in kt.items() if v[1]}
    if rank == 0 and args.check:
        single = cabm.Community(P, pop, T, attrs=attrs)
        single.run_days(args.warm)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        single.run_days(args.days)
        torch.cuda.synchronize(); out["single_handle_us_per_day_wall"] = (time.perf_counter() - t1) * 1e6 / args.days
        out["equals_single_handle"] = bool(np.array_equal(single.ledger_table().astype(np.int64), led.cpu().numpy()))
        single.device_time_ms(); single.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
