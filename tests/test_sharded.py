"""One community over several handles: the sharded run is bit-identical to the single-handle run.

CPU: the oracle backend, shards in one process (LocalExchange) and one shard per process over gloo (DistExchange).
GPU: the CUDA library, several shards on one device (LocalExchange) against the single-handle CUDA run and the oracle;
the in-library NCCL loop (cabm_shard_run_days) with a one-rank communicator on any box and, where the box has two GPUs, with
two ranks under torchrun (tools/shard_run.py) against the single-handle run.
"""
import os
import socket
import sys

import numpy as np
import pytest

import covid_abm_b200 as cabm
import synthpop
import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MUT = ("2020-02-25", "2020-03-10", "2020-03-25")


def _case(n=9000, days=70, seed=13, T=0.16, pde=2e-3):
    pop = synthpop.generate(n, seed=21)
    P = cabm.params.default_params(run_length=days, seed=seed, T=T, prob_daily_exposure=pde, mutant_intro_dates=MUT)
    tab = cabm.params.build_tables(P)
    attrs = cabm.abi.immunity_attributes(int(pop["n_people"]), P["seed"])
    return pop, P, tab, attrs


def _script(c, pop, days, check=None):
    """The same sequence of calls on a single community or a sharded one."""
    n = int(pop["n_people"]); rng = np.random.default_rng(3)
    for pid in range(0, 1500, 11):
        c.infect(pid, 0, 0)
    for d in range(days):
        if d in (12, 33):
            pids = rng.choice(n, size=n // 3, replace=False)
            c.vaccinate(pids, d)
        if d == 25:
            c.quarantine(rng.choice(n, size=n // 8, replace=False), d, d + 12)
        c.step()
        if check is not None:
            check(d)


def _sorted_hist(c):
    h = c.infection_histories()
    return h[np.lexsort((h["t"][:, 0], h["person"]))]


def _assert_same(single, sharded, label):
    assert np.array_equal(single.ledger_table(), sharded.ledger_table()), f"{label}: ledger"
    assert np.array_equal(single.getCumulIncidenceByOutcome(), sharded.getCumulIncidenceByOutcome()), f"{label}: outcomes"
    a, b = single.current_infections(), sharded.current_infections()
    bad = np.flatnonzero(a != b)
    assert bad.size == 0, f"{label}: {bad.size} agents differ, first {bad[:3]}"


def test_household_cuts_never_split_a_home():
    pop, P, tab, attrs = _case(n=5000)
    for world in (2, 3, 8):
        cuts = cabm.sharded.household_cuts(pop, world)
        assert cuts[0] == 0 and cuts[-1] == int(pop["n_people"]) and all(a <= b for a, b in zip(cuts, cuts[1:]))
        for p in cuts[1:-1]:
            assert pop["home"][p] != pop["home"][p - 1]
        sizes = np.diff(cuts)
        assert sizes.max() - sizes.min() < 20


@pytest.mark.parametrize("world", [2, 3])
def test_oracle_sharded_equals_single(world):
    pop, P, tab, attrs = _case()
    single = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    sh = cabm.sharded.local_sharded(P, pop, tab, world, backend=oracle_lib.oracle_backend(), attrs=attrs)
    single.foi(); [s.foi() for s in sh.shards]
    _script(single, pop, 70)

    def check(d):
        if d in (5, 30, 69):
            f = sh.foi()
            for s in sh.shards[1:]:
                assert np.array_equal(f, s.foi()), f"day {d}: shards disagree on the force of infection"
    _script(sh, pop, 70, check)
    _assert_same(single, sh, f"world {world}")
    hs, hh = _sorted_hist(single), _sorted_hist(sh)
    assert hs.size == hh.size > 1500 and np.array_equal(hs, hh)
    assert np.array_equal(single.foi(), sh.foi())
    w1, w2 = single.work_counters(), sh.work_counters()
    assert w1 == w2, {k: (w1[k], w2[k]) for k in w1 if w1[k] != w2[k]}
    owned = [int(s.ledger("new_infected").sum()) for s in sh.shards]
    assert all(o > 0 for o in owned)                        # every shard carried part of the epidemic
    strains = single.ledger_table()[:, 16:20].sum(axis=0)
    assert (strains > 0).sum() >= 3


def _gloo_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import covid_abm_b200 as cabm2
    import oracle_lib as ol
    pop, P, tab, attrs = _case(n=6000, days=45)
    cuts = cabm2.sharded.household_cuts(pop, world)
    shard = cabm2.sharded.Shard(P, pop, tab, rank, world, cuts, backend=ol.oracle_backend(), attrs=attrs)
    c = cabm2.sharded.ShardedCommunity([shard], cabm2.sharded.DistExchange(shard, dist))
    for pid in range(0, 1500, 11):
        c.infect(pid, 0, 0)
    c.run_days(45)
    led = torch.from_numpy(c.ledger_table().astype(np.int64))
    dist.all_reduce(led)
    cur = c.current_infections()
    gathered = [None] * world
    dist.all_gather_object(gathered, cur)
    if rank == 0:
        q.put((led.numpy().astype(np.uint64), np.concatenate(gathered)))
    dist.barrier(); dist.destroy_process_group()


def test_gloo_two_ranks_equal_single():
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    ctx = mp.get_context("spawn"); q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    led, cur = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60); assert p.exitcode == 0
    pop, P, tab, attrs = _case(n=6000, days=45)
    single = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    for pid in range(0, 1500, 11):
        single.infect(pid, 0, 0)
    single.run_days(45)
    assert np.array_equal(single.ledger_table(), led)
    assert np.array_equal(single.current_infections(), cur)
    assert led[:, 0].sum() > 500


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4])
def test_cuda_sharded_on_one_device_equals_single(world):
    import torch
    pop, P, tab, attrs = _case(n=30000, days=70)
    single = cabm.Community(P, pop, tab, attrs=attrs)
    oracle = oracle_lib.oracle_community(P, pop, tab, attrs=attrs)
    sh = cabm.sharded.local_sharded(P, pop, tab, world, attrs=attrs, device=torch.device("cuda", 0))
    _script(single, pop, 70); _script(oracle, pop, 70); _script(sh, pop, 70)
    _assert_same(single, sh, f"cuda world {world} vs single")
    _assert_same(oracle, sh, f"cuda world {world} vs oracle")
    hs, hh = _sorted_hist(single), _sorted_hist(sh)
    assert hs.size == hh.size > 3000 and np.array_equal(hs, hh)


@pytest.mark.gpu
def test_nccl_day_loop_with_one_rank_equals_run_days():
    """cabm_shard_run_days (NCCL dlopen()ed, collectives on the library's stream, gathers sized between day groups) == cabm_run_days."""
    import torch
    pop, P, tab, attrs = _case(n=30000, days=75)
    single = cabm.Community(P, pop, tab, attrs=attrs)
    cuts = cabm.sharded.household_cuts(pop, 1)
    sh = cabm.sharded.Shard(P, pop, tab, 0, 1, cuts, attrs=attrs, device=torch.device("cuda", 0))
    sh.nccl_init(None)
    for c in (single, sh):
        for pid in range(0, 1500, 11):
            c.infect(pid, 0, 0)
    single.run_days(30); sh.run_days_nccl(30)
    pids = np.random.default_rng(3).choice(int(pop["n_people"]), size=9000, replace=False)
    single.vaccinate(pids, 30); sh.vaccinate(pids, 30)
    single.run_days(45); sh.run_days_nccl(45)                      # 45 days = 11 groups of four + one
    assert np.array_equal(single.ledger_table(), sh.ledger_table())
    assert np.array_equal(single.current_infections(), sh.current_infections())
    assert single.ledger_table()[:, 0].sum() > 3000
    single.close(); sh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4])
def test_nccl_day_loop_on_several_gpus_equals_single_handle(world):
    """One rank per GPU under torchrun (tools/shard_run.py): NCCL + peer-memory exchanges inside the library, summed ledger equal
    to one handle's. Runs with 2 ranks wherever the box has two GPUs and with 4 wherever it has four."""
    import json
    import subprocess
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip(f"{torch.cuda.device_count()} GPU(s) on this box")
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tools", "shard_run.py"), "--agents", "200000", "--days", "40", "--warm", "20"],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    out = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert out["world"] == world and out["native_nccl"] and out["equals_single_handle"] and out["infections"] > 0
