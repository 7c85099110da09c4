"""One community advanced by several handles (one per GPU): host side of the `cabm_shard_*` entry points of include/cabm.h.

SURVEY.md 8e row 2. Every shard holds the whole static population; shard r owns a contiguous range of agents cut at
household boundaries (an agent is owned by the shard of its home). A day is four device parts with three exchanges:

    part 0 -> all-gather of the flag records -> part 1 -> all-reduce of the visitor census -> part 2
           -> all-reduce of the hospital admissions delta -> part 3

`Exchange` is the plumbing only: `LocalExchange` moves the buffers between handles that live in one process (tests; several
shards on one GPU or on the CPU oracle), `DistExchange` uses torch.distributed (NCCL between GPUs, gloo for the CPU tests).
No agent or location arithmetic happens here. Results are bit-identical to the same community on a single handle
(tests/test_sharded.py).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi
from .community import Community


def household_cuts(pop: dict, world: int) -> list[int]:
    """`world`+1 agent-id cut points, as even as possible, never inside a household (members of a home are contiguous ids)."""
    n = int(pop["n_people"]); home = np.asarray(pop["home"])
    cuts = [0]
    for r in range(1, world):
        p = (n * r) // world
        while 0 < p < n and home[p] == home[p - 1]:
            p += 1
        cuts.append(max(p, cuts[-1]))
    cuts.append(n)
    return cuts


class _Buf:
    """A backend buffer (pointer + element count + dtype) viewed as numpy (host backends) or torch (CUDA) without copying."""

    def __init__(self, ptr: int, n: int, dtype, device):
        self.ptr, self.n, self.dtype, self.device = ptr, n, np.dtype(dtype), device

    def numpy(self) -> np.ndarray:
        if self.n == 0:
            return np.zeros(0, dtype=self.dtype)
        return np.ctypeslib.as_array(C.cast(self.ptr, C.POINTER(C.c_uint8)), shape=(self.n * self.dtype.itemsize,)).view(self.dtype)

    def torch(self):
        import torch
        if self.device is None:
            return torch.from_numpy(self.numpy())
        if self.n == 0:
            return torch.zeros(0, dtype=getattr(torch, self.dtype.name), device=self.device)

        class _W:                                     # zero-copy view of device memory owned by the library
            pass
        w = _W()
        w.__cuda_array_interface__ = {"shape": (self.n,), "typestr": self.dtype.str, "data": (self.ptr, False), "version": 2}
        return torch.as_tensor(w, device=self.device)


class Shard(Community):
    """One handle of a sharded community; same interface as Community plus the four day parts."""

    def __init__(self, P, pop, tables, rank: int, world: int, cuts, backend=None, attrs=None, device=None):
        super().__init__(P, pop, tables, backend=backend, attrs=attrs)
        self.rank, self.world, self.lo, self.hi = rank, world, int(cuts[rank]), int(cuts[rank + 1])
        self.device = device                          # torch device of the buffers, None for host backends
        self._check(self.fn.shard_config(self._h, rank, world, self.lo, self.hi))

    def part(self, k: int, day: int, with_import: bool = True) -> int:
        out = C.c_int32(0)
        w = (C.c_double * 4)(*self.strains) if (with_import and k == 3) else None
        self._check(self.fn.shard_day_part(self._h, k, day, int(self.julian_day[day]), int(self.is_weekday[day]),
                                           self.expected_imports(day) if w is not None else 0.0, w, C.byref(out)))
        return out.value

    def _buf(self, fn, dtype) -> _Buf:
        p = C.c_void_p(); n = C.c_size_t(0)
        self._check(fn(self._h, C.byref(p), C.byref(n)))
        return _Buf(p.value or 0, n.value, dtype, self.device)

    # ---- the same protocol inside the library over NCCL (cabm_shard_nccl_init / cabm_shard_run_days) ----
    def nccl_init(self, dist, group=None):
        """Collective: rank 0 makes the NCCL id, torch.distributed (any backend) hands it round, every rank joins."""
        if self.fn.shard_nccl_init is None:
            raise RuntimeError("this backend has no NCCL path")
        ident = [None]
        if self.rank == 0:
            buf = (C.c_uint8 * 128)()
            self._check(self.fn.shard_nccl_id(C.cast(buf, C.c_void_p)))
            ident = [bytes(buf)]
        if self.world > 1:
            dist.broadcast_object_list(ident, src=0, group=group)
        buf = (C.c_uint8 * 128).from_buffer_copy(ident[0])
        self._check(self.fn.shard_nccl_init(self._h, C.cast(buf, C.c_void_p), self.rank, self.world))

    def peer_memory(self) -> bool:
        m = C.c_int(0); self._check(self.fn.shard_exchange_mode(self._h, C.byref(m)))
        return bool(m.value)

    def run_days_nccl(self, n_days: int):
        first = self.day
        strain_w = np.zeros((n_days, 4)); exp = np.zeros(n_days)
        for i in range(n_days):
            self.update_strain_mix(first + i)
            strain_w[i] = self.strains; exp[i] = self.expected_imports(first + i)
        jd = np.ascontiguousarray(self.julian_day[first:first + n_days]); wd = np.ascontiguousarray(self.is_weekday[first:first + n_days])
        self._check(self.fn.shard_run_days(self._h, first, n_days, jd.ctypes.data_as(C.POINTER(C.c_int32)), wd.ctypes.data_as(C.POINTER(C.c_int32)),
                                           exp.ctypes.data_as(C.POINTER(C.c_double)), strain_w.ctypes.data_as(C.POINTER(C.c_double))))
        self.day += n_days

    def flags_out(self) -> _Buf:
        b = self._buf(self.fn.shard_flags_out, np.uint8)
        b.n *= abi.SHARD_FLAG_BYTES
        return b

    def flags_in(self, ptr: int, n_records: int):
        if n_records:
            self._check(self.fn.shard_flags_in(self._h, C.c_void_p(ptr), n_records))

    def census(self) -> _Buf:
        return self._buf(self.fn.shard_census, np.uint32)

    def patients(self) -> _Buf:
        return self._buf(self.fn.shard_patients, np.int32)


class LocalExchange:
    """All shards live in this process: buffers are copied handle to handle (device-to-device for the CUDA library)."""

    def __init__(self, shards):
        self.shards = shards

    def flags(self):
        outs = [s.flags_out() for s in self.shards]
        for i, s in enumerate(self.shards):
            for j, o in enumerate(outs):
                if i != j:
                    s.flags_in(o.ptr, o.n // abi.SHARD_FLAG_BYTES)

    def _sum(self, bufs):
        if bufs[0].device is None:
            views = [b.numpy() for b in bufs]
            tot = np.sum(np.stack(views), axis=0, dtype=views[0].dtype)
            for v in views:
                v[...] = tot
        else:
            import torch
            views = [b.torch() for b in bufs]
            if views[0].numel():
                tot = torch.stack([v.to(torch.int64) for v in views]).sum(dim=0).to(views[0].dtype)
                for v in views:
                    v.copy_(tot)
                torch.cuda.synchronize()

    def census(self):
        self._sum([s.census() for s in self.shards])

    def patients(self):
        self._sum([s.patients() for s in self.shards])


class DistExchange:
    """One shard per process; exchanges are torch.distributed collectives (NCCL on GPUs, gloo on CPU)."""

    def __init__(self, shard, dist, group=None):
        self.s, self.dist, self.group = shard, dist, group

    def flags(self):
        import torch
        s, dist = self.s, self.dist
        out = s.flags_out()
        dev = s.device if s.device is not None else "cpu"
        cnt = torch.tensor([out.n], dtype=torch.int64, device=dev)
        cnts = [torch.zeros_like(cnt) for _ in range(s.world)]
        dist.all_gather(cnts, cnt, group=self.group)
        cnts = [int(c.item()) for c in cnts]
        m = max(cnts)
        if m == 0:
            return
        send = torch.zeros(m, dtype=torch.uint8, device=dev)
        if out.n:
            send[:out.n] = out.torch()
        recv = torch.empty(m * s.world, dtype=torch.uint8, device=dev)
        dist.all_gather_into_tensor(recv, send, group=self.group)
        if s.device is not None:
            torch.cuda.current_stream().synchronize()
        for r, c in enumerate(cnts):
            if r != s.rank and c:
                s.flags_in(recv.data_ptr() + r * m, c // abi.SHARD_FLAG_BYTES)
        self._keep = recv                             # alive until the library has consumed it (next part is stream-ordered after the copy)

    def _sum(self, buf):
        import torch
        t = buf.torch()
        if t.numel() == 0:
            return
        if t.dtype == torch.uint32:
            t = t.view(torch.int32)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        if self.s.device is not None:
            torch.cuda.current_stream().synchronize()

    def census(self):
        self._sum(self.s.census())

    def patients(self):
        self._sum(self.s.patients())


class ShardedCommunity:
    """Drives the shards of this process through the day; with LocalExchange it holds all of them, with DistExchange one."""

    def __init__(self, shards, exchange):
        self.shards, self.x = shards, exchange
        self.day = 0

    def step(self):
        d = self.day
        for s in self.shards:
            s.part(0, d)
        self.x.flags()
        for s in self.shards:
            s.part(1, d)
        self.x.census()
        for s in self.shards:
            s.part(2, d)
        self.x.patients()
        n = 0
        for s in self.shards:
            s.update_strain_mix(d)
            n += s.part(3, d)
            s.day = d + 1
        self.day = d + 1
        return n

    def run_days(self, n: int):
        for _ in range(n):
            self.step()

    def infect(self, pid: int, strain: int = 0, day: int | None = None) -> bool:
        return any([s.infect(pid, strain, self.day if day is None else day) for s in self.shards])   # every shard counts the call

    def owner(self, pid: int):
        for s in self.shards:
            if s.lo <= pid < s.hi:
                return s
        return None

    def vaccinate(self, pids, day: int | None = None):
        pids = np.asarray(pids, dtype=np.uint32); ok = np.zeros(pids.size, dtype=bool)
        for s in self.shards:
            m = (pids >= s.lo) & (pids < s.hi)
            if m.any():
                ok[m] = s.vaccinate(pids[m], self.day if day is None else day)
        return ok

    def quarantine(self, pids, from_day: int, until_day: int):
        pids = np.asarray(pids, dtype=np.uint32)
        for s in self.shards:
            m = (pids >= s.lo) & (pids < s.hi)
            if m.any():
                s.quarantine(pids[m], from_day, until_day)

    # local sums over the shards of this process (with DistExchange the caller all-reduces them)
    def ledger_table(self) -> np.ndarray:
        return np.sum(np.stack([s.ledger_table() for s in self.shards]), axis=0, dtype=np.uint64)

    def getCumulIncidenceByOutcome(self) -> np.ndarray:
        return np.sum(np.stack([s.getCumulIncidenceByOutcome() for s in self.shards]), axis=0, dtype=np.uint64)

    def current_infections(self) -> np.ndarray:
        return np.concatenate([s.current_infections(np.arange(s.lo, s.hi, dtype=np.uint32)) for s in self.shards])

    def infection_histories(self) -> np.ndarray:
        return np.concatenate([s.infection_histories() for s in self.shards])

    def work_counters(self) -> dict:
        out = {}
        for s in self.shards:
            for k, v in s.work_counters().items():
                out[k] = out.get(k, 0) + v
        return out

    def foi(self) -> np.ndarray:
        return self.shards[0].foi()

    def close(self):
        for s in self.shards:
            s.close()


def local_sharded(P, pop, tables, world: int, backend=None, attrs=None, device=None) -> ShardedCommunity:
    cuts = household_cuts(pop, world)
    shards = [Shard(P, pop, tables, r, world, cuts, backend=backend, attrs=attrs, device=device) for r in range(world)]
    return ShardedCommunity(shards, LocalExchange(shards))
