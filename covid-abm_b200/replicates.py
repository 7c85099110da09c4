"""Replicate-parallel runs: one independent community per rank, no collective on the data path.

This is the reference's own scaling mode (one replicate per SLURM array task, exp/active-vac/covid-simulate.slurm:16;
SURVEY.md 8e row 1). Ranks never exchange agent or location state; the only communication is the final gather of the
per-replicate ledgers and the max-over-ranks of the device time used for throughput. Works over NCCL (one rank per GPU)
and over gloo (CPU tests).
"""
from __future__ import annotations

import numpy as np


def replicate_seed(base_seed: int, rank: int, replicate: int = 0, world: int = 1) -> int:
    """Seed of replicate number `replicate` of `rank`: replicates are striped over ranks, seeds are base + global index."""
    return int(base_seed) + int(replicate) * int(world) + int(rank)


def replicate_slice(n_replicates: int, rank: int, world: int) -> list[int]:
    """Global replicate indices owned by `rank` (striped, so every rank gets floor or ceil of n/world)."""
    return list(range(rank, n_replicates, world))


def job_throughput(dist, device, units_local: float, seconds_local: float):
    """(total units over all ranks, max seconds over ranks, units/s): the whole-job figure bench.py reports."""
    import torch
    t = torch.tensor([seconds_local], dtype=torch.float64, device=device)
    u = torch.tensor([units_local], dtype=torch.float64, device=device)
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(u.item()), float(t.item()), float(u.item()) / float(t.item())


def gather_ledgers(dist, device, ledgers: np.ndarray, n_replicates: int):
    """Collect per-replicate ledgers [r_local][run_length][C] (uint64) on rank 0 in global replicate order.

    Every rank must pass the same run_length and C; ranks owning fewer replicates are padded for the collective.
    Returns the [n_replicates][run_length][C] array on rank 0 and None elsewhere."""
    import torch
    world = dist.get_world_size() if dist is not None and dist.is_initialized() else 1
    rank = dist.get_rank() if world > 1 else 0
    if world == 1:
        return np.asarray(ledgers)
    per = -(-n_replicates // world)
    shape = (per,) + tuple(ledgers.shape[1:])
    buf = np.zeros(shape, dtype=np.int64)
    buf[:ledgers.shape[0]] = ledgers.astype(np.int64)
    mine = torch.from_numpy(buf).to(device)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine)
    if rank != 0:
        return None
    full = np.zeros((n_replicates,) + tuple(ledgers.shape[1:]), dtype=np.uint64)
    for r in range(world):
        idx = replicate_slice(n_replicates, r, world)
        full[idx] = out[r].cpu().numpy()[:len(idx)].astype(np.uint64)
    return full


def batch_run_days(comms, n_days: int):
    """`n_days` of every community in `comms` (same population shape, same calendar position) with ONE launch per kernel for the
    whole batch (cabm_batch_run_days, include/cabm.h): the calibration-batch form of Community.run_days. Each community keeps
    its own parameters, seed, state and ledger; results equal run_days on each of them."""
    import ctypes as C
    c0 = comms[0]
    first = c0.day
    if any(c.day != first or c.n_people != c0.n_people or c.backend is not c0.backend for c in comms):
        raise ValueError("a batch advances communities of one shape from the same day on one backend")
    strain_w = np.zeros((n_days, 4)); exp = np.zeros(n_days)
    for i in range(n_days):
        for c in comms:
            c.update_strain_mix(first + i)
        strain_w[i] = c0.strains; exp[i] = c0.expected_imports(first + i)
        # cabm_batch_run_days takes ONE importation schedule and strain urn for the whole batch (include/cabm.h): a sweep that
        # varies prob_daily_exposure or the variant calendar across replicates must not be batched silently with comms[0]'s
        for c in comms[1:]:
            if list(c.strains) != list(c0.strains) or c.expected_imports(first + i) != exp[i]:
                raise ValueError("cabm_batch_run_days shares the importation schedule and the strain urn: every community of a batch "
                                 "must have the same prob_daily_exposure, population size and mutant_intro_dates")
    jd = np.ascontiguousarray(c0.julian_day[first:first + n_days]); wd = np.ascontiguousarray(c0.is_weekday[first:first + n_days])
    hs = (C.c_void_p * len(comms))(*[c._h for c in comms])
    c0._check(c0.fn.batch_run_days(hs, len(comms), first, n_days, jd.ctypes.data_as(C.POINTER(C.c_int32)), wd.ctypes.data_as(C.POINTER(C.c_int32)),
                                   exp.ctypes.data_as(C.POINTER(C.c_double)), strain_w.ctypes.data_as(C.POINTER(C.c_double))))
    for c in comms:
        c.day += n_days
