"""The N>1 path of bench.py is replicate-parallel: world_size-2 gloo test of its host logic.

Each rank advances its own replicates (here on the CPU oracle, which exports the same ABI), ledgers are gathered to
rank 0 in global replicate order and must equal the same replicates run serially in one process; the whole-job
throughput reduction takes the max time and the sum of units.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
N_REPL, DAYS, AGENTS = 5, 25, 3000


def _run_replicate(seed):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import covid_abm_b200 as cabm
    import synthpop
    import oracle_lib
    pop = synthpop.generate(AGENTS, seed=3)
    P = cabm.params.default_params(run_length=DAYS, seed=seed, T=0.15, prob_daily_exposure=2e-3)
    c = oracle_lib.oracle_community(P, pop, cabm.params.build_tables(P))
    c.run_days(DAYS)
    out = c.ledger_table()
    c.close()
    return out, int(pop["n_people"])


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    import covid_abm_b200 as cabm
    R = cabm.replicates
    mine = R.replicate_slice(N_REPL, rank, world)
    ledgers, n = [], 0
    for k, g in enumerate(mine):
        seed = R.replicate_seed(100, rank, k, world)
        assert seed == 100 + g
        led, n = _run_replicate(seed)
        ledgers.append(led)
    full = R.gather_ledgers(dist, torch.device("cpu"), np.stack(ledgers), N_REPL)
    units, secs, rate = R.job_throughput(dist, torch.device("cpu"), float(n) * DAYS * len(mine), 1.0 + rank)
    if rank == 0:
        q.put((full, units, secs, rate, n))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_replicates_match_serial():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    full, units, secs, rate, n = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60); assert p.exitcode == 0
    assert full.shape[0] == N_REPL
    for g in range(N_REPL):
        want, _ = _run_replicate(100 + g)
        assert np.array_equal(full[g], want), f"replicate {g}"
    assert len({full[g].tobytes() for g in range(N_REPL)}) == N_REPL            # replicates are distinct epidemics
    assert units == float(n) * DAYS * N_REPL and secs == 2.0 and rate == units / 2.0


def test_replicate_striping():
    import covid_abm_b200 as cabm
    R = cabm.replicates
    for n, w in ((256, 8), (5, 2), (3, 4), (1, 1)):
        owned = [R.replicate_slice(n, r, w) for r in range(w)]
        assert sorted(sum(owned, [])) == list(range(n))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
        seeds = {R.replicate_seed(7, r, k, w) for r in range(w) for k in range(len(owned[r]))}
        assert seeds == {7 + g for g in range(n)}


def _batch_case(make, R=3, n=4000, days=40):
    import covid_abm_b200 as cabm
    import synthpop
    pop = synthpop.generate(n, seed=77)
    comms, solo = [], []
    for r in range(R):
        P = cabm.params.default_params(run_length=days, seed=100 + r, T=0.15 + 0.02 * r, prob_daily_exposure=1e-3,
                                       mutant_intro_dates=("2020-02-20", "2020-03-05", "2020-03-20"))
        tab = cabm.params.build_tables(P)
        comms.append(make(P, pop, tab)); solo.append(make(P, pop, tab))
    cabm.replicates.batch_run_days(comms, 15)
    pids = np.arange(0, n, 3, dtype=np.uint32)
    for c, s in zip(comms, solo):
        s.run_days(15)
        c.vaccinate(pids, 15); s.vaccinate(pids, 15)
    cabm.replicates.batch_run_days(comms, days - 15)
    totals = []
    for c, s in zip(comms, solo):
        s.run_days(days - 15)
        assert c.day == s.day == days
        assert np.array_equal(c.ledger_table(), s.ledger_table())
        assert np.array_equal(c.current_infections(), s.current_infections())
        totals.append(int(c.ledger_table()[:, 0].sum()))
    assert len(set(totals)) == R and min(totals) > 100          # different parameters and seeds: different epidemics
    for c in comms + solo:
        c.close()


def test_batched_replicates_wrapper_on_the_oracle():
    """Host side of cabm_batch_run_days (strain urn, calendar slices, handle array): batch == one community at a time."""
    import oracle_lib
    _batch_case(lambda P, pop, tab: oracle_lib.oracle_community(P, pop, tab))


@pytest.mark.gpu
def test_batched_replicates_one_launch_per_kernel_equals_run_days():
    """CUDA: every kernel of the day launched once for the whole batch (replicate = blockIdx.y) == cabm_run_days per handle."""
    import covid_abm_b200 as cabm
    _batch_case(lambda P, pop, tab: cabm.Community(P, pop, tab), R=5, n=30000, days=60)
