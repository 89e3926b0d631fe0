"""A small executable model of the ring / flag protocol of csrc/kernels.cuh (run_pipeline) and
csrc/ll.cuh, for checking it on the CPU under randomised schedules.

What is modelled: N ranks x B rings, DEPTH slots per ring; monotonic per-ring
tile counters; the three flag channels READY / MID / DONE; the slot guard (wait for
DONE(g - DEPTH) from all ranks); one flag round per tile with IN(t+1)+OUT(t-1) and MID(t)
running concurrently on two warp teams; per-collective slot regions and who writes / reads
them. Every stage is split into a begin and an end event so that other CTAs can be
scheduled in between, and flags only become visible when posted (writes of a stage are
complete before the post, which is what the system fence guarantees).

What is checked:
  * progress: every CTA of every rank terminates (no deadlock) under any schedule tried;
  * every read of a slot region sees exactly the data of the tile it belongs to, written by
    the rank it expects (no stale data, no data of a later tile);
  * no write lands in a region while a read of an earlier tile's data is still in flight,
    or before every expected reader of that earlier data has read it.

`pipeline="v4"` models the round-2 kernels: one agent per stage that has work on the rank, a
publisher that raises every flag late (behind its fence) and raises the flags of idle stages
on behalf of the stage before them, and the call-boundary guard of the first DEPTH tiles.
(The model keeps DEPTH fixed; the kernels choose it per call, which only changes how far IN
may run ahead.)

It mirrors the protocol, not the CUDA source line by line: it validates the design (guards,
flag semantics, team overlap, interleaving of different collectives on the same rings), which
is the part that cannot be unit-tested on a GPU deterministically.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import Callable, Dict, Generator, List, Optional, Tuple

DEPTH = 4
READY, MID, DONE = 0, 1, 2

# An access: (rank whose block is touched, region name). Regions are symbolic:
# ("A", q) partial/tile from rank q, ("B", o) reduced chunk o, ("T",) the whole tile,
# ("LL", q) LL packets from rank q.
Region = Tuple
Access = Tuple[int, Region]


@dataclass
class Stage:
    writes: List[Access] = field(default_factory=list)
    reads: List[Tuple[int, Region, int]] = field(default_factory=list)  # (rank, region, expected writer)


@dataclass
class Collective:
    """Per-rank description: which stages exist and what they touch."""
    name: str
    has_out: bool
    in_: Callable[[int], Optional[Stage]]    # rank -> Stage or None (no work on this rank)
    mid: Callable[[int], Optional[Stage]]
    out: Callable[[int], Optional[Stage]]
    ready_to: Callable[[int], List[int]]
    ready_from: Callable[[int], List[int]]
    ll: bool = False  # LL kernels: no READY/MID flags, data carries its own flag


def _all(n):
    return list(range(n))


def all_reduce_oneshot(n: int) -> Collective:
    return Collective(
        "all_reduce_oneshot", False,
        in_=lambda r: Stage(writes=[(p, ("A", r)) for p in range(n) if p != r]),
        mid=lambda r: Stage(reads=[(r, ("A", q), q) for q in range(n) if q != r]),
        out=lambda r: None,
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def all_reduce_twoshot(n: int) -> Collective:
    return Collective(
        "all_reduce_twoshot", True,
        in_=lambda r: Stage(writes=[(p, ("A", r)) for p in range(n) if p != r]),
        mid=lambda r: Stage(reads=[(r, ("A", q), q) for q in range(n) if q != r],
                            writes=[(p, ("B", r)) for p in range(n) if p != r]),
        out=lambda r: Stage(reads=[(r, ("B", j), j) for j in range(n) if j != r]),
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def all_reduce_nvls(n: int) -> Collective:
    # copy-in own tile; ld_reduce reads every rank's tile copy; multimem.st writes the reduced
    # shard into every rank's tile; copy-out reads own tile
    return Collective(
        "all_reduce_nvls", True,
        in_=lambda r: Stage(writes=[(r, ("T", "in"))]),
        mid=lambda r: Stage(reads=[(p, ("T", "in"), p) for p in range(n)],
                            writes=[(p, ("S", r)) for p in range(n)]),
        out=lambda r: Stage(reads=[(r, ("S", j), j) for j in range(n)]),
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def reduce_rooted_twophase(n: int, root: int) -> Collective:
    owners = [r for r in range(n) if r != root]
    return Collective(
        f"reduce_2p(root={root})", True,
        in_=lambda r: Stage(writes=[(p, ("A", r)) for p in owners if p != r]),
        mid=lambda r: None if r == root else Stage(
            reads=[(r, ("A", q), q) for q in range(n) if q != r], writes=[(root, ("B", r))]),
        out=lambda r: Stage(reads=[(root, ("B", j), j) for j in owners]) if r == root else None,
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def reduce_oneshot(n: int, root: int) -> Collective:
    return Collective(
        f"reduce_1s(root={root})", False,
        in_=lambda r: None if r == root else Stage(writes=[(root, ("A", r))]),
        mid=lambda r: Stage(reads=[(root, ("A", q), q) for q in range(n) if q != root]) if r == root else None,
        out=lambda r: None,
        ready_to=lambda r: [root], ready_from=lambda r: _all(n) if r == root else [])


def broadcast_push(n: int, root: int) -> Collective:
    return Collective(
        f"broadcast(root={root})", False,
        in_=lambda r: Stage(writes=[(p, ("T", "bc")) for p in range(n) if p != root]) if r == root else None,
        mid=lambda r: None if r == root else Stage(reads=[(r, ("T", "bc"), root)]),
        out=lambda r: None,
        ready_to=lambda r: _all(n) if r == root else [], ready_from=lambda r: [] if r == root else [root])


def broadcast_twophase(n: int, root: int) -> Collective:
    owners = [r for r in range(n) if r != root]
    return Collective(
        f"broadcast_2p(root={root})", True,
        in_=lambda r: Stage(writes=[(p, ("C", p)) for p in owners]) if r == root else None,
        mid=lambda r: None if r == root else Stage(
            reads=[(r, ("C", r), root)], writes=[(p, ("C2", r)) for p in owners if p != r]),
        out=lambda r: None if r == root else Stage(
            reads=[(r, ("C", r), root)] + [(r, ("C2", j), j) for j in owners if j != r]),
        ready_to=lambda r: _all(n) if r == root else [], ready_from=lambda r: [root])


def scatter(n: int, root: int) -> Collective:
    return Collective(
        f"scatter(root={root})", False,
        in_=lambda r: Stage(writes=[(p, ("T", "sc")) for p in range(n) if p != root]) if r == root else None,
        mid=lambda r: None if r == root else Stage(reads=[(r, ("T", "sc"), root)]),
        out=lambda r: None,
        ready_to=lambda r: _all(n) if r == root else [], ready_from=lambda r: [] if r == root else [root])


def gather(n: int, root: int) -> Collective:
    return Collective(
        f"gather(root={root})", False,
        in_=lambda r: None if r == root else Stage(writes=[(root, ("A", r))]),
        mid=lambda r: Stage(reads=[(root, ("A", q), q) for q in range(n) if q != root]) if r == root else None,
        out=lambda r: None,
        ready_to=lambda r: [root], ready_from=lambda r: _all(n) if r == root else [])


def all_gather_push(n: int) -> Collective:
    return Collective(
        "all_gather_push", False,
        in_=lambda r: Stage(writes=[(p, ("A", r)) for p in range(n) if p != r]),
        mid=lambda r: Stage(reads=[(r, ("A", q), q) for q in range(n) if q != r]),
        out=lambda r: None,
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def all_gather_pull(n: int) -> Collective:
    return Collective(
        "all_gather_pull", False,
        in_=lambda r: Stage(writes=[(r, ("T", "ag"))]),
        mid=lambda r: Stage(reads=[(p, ("T", "ag"), p) for p in range(n) if p != r]),
        out=lambda r: None,
        ready_to=lambda r: _all(n), ready_from=lambda r: _all(n))


def ll_all_reduce(n: int) -> Collective:
    return Collective(
        "ll_all_reduce", False,
        in_=lambda r: Stage(writes=[(p, ("LL", r)) for p in range(n) if p != r]),
        mid=lambda r: Stage(reads=[(r, ("LL", q), q) for q in range(n) if q != r]),
        out=lambda r: None, ready_to=lambda r: [], ready_from=lambda r: [], ll=True)


class ProtocolError(AssertionError):
    pass


class Machine:
    """State shared by all CTAs: flags, slot contents, in-flight reads."""

    def __init__(self, n: int, rings: int):
        self.n, self.rings = n, rings
        self.flag = {}     # (chan, dst_rank, ring, src_rank) -> value
        self.epoch = {}    # (rank, ring) -> tiles processed
        # (rank, ring, pos, region) -> dict(tile=g, writer=w, readers_left=set, reading=int, writing=bool)
        self.mem: Dict[Tuple, dict] = {}
        self.no_guard = False  # fault injection for the tests: skip the DONE(g - DEPTH) slot guard

    def get_flag(self, chan, dst, ring, src):
        return self.flag.get((chan, dst, ring, src), 0)

    def post(self, chan, dst, ring, src, val):
        key = (chan, dst, ring, src)
        if val < self.flag.get(key, 0):
            raise ProtocolError(f"flag {key} would go backwards: {self.flag[key]} -> {val}")
        self.flag[key] = val

    # ---- memory model -------------------------------------------------------------
    def begin_write(self, who, rank, ring, g, region, readers):
        key = (rank, ring, g % DEPTH, region)
        # Regions of different collectives alias physically (they are all carved out of the
        # same slot), so a write for tile g must find NO older tile's data in this slot that
        # is still being read or still has readers to come, whatever its region is called.
        for other_key, other in self.mem.items():
            if other_key[:3] == key[:3] and other["tile"] != g and (other["reading"] or other["readers_left"]):
                raise ProtocolError(f"rank {who} writes slot {key[:3]} for tile {g} while region "
                                    f"{other_key[3]} of tile {other['tile']} is unread / being read "
                                    f"(readers left {sorted(other['readers_left'])})")
        cell = self.mem.get(key)
        if cell is not None:
            if cell["reading"]:
                raise ProtocolError(f"rank {who} writes {key} for tile {g} while it is being read "
                                    f"(holds tile {cell['tile']})")
            if cell["writing"]:
                raise ProtocolError(f"rank {who} writes {key} for tile {g}: concurrent writer")
            if cell["readers_left"] and cell["tile"] != g:
                raise ProtocolError(f"rank {who} overwrites {key} (tile {cell['tile']}) with tile {g} "
                                    f"before ranks {sorted(cell['readers_left'])} read it")
        self.mem[key] = {"tile": g, "writer": who, "readers_left": set(readers), "reading": 0,
                         "writing": True}

    def end_write(self, rank, ring, g, region):
        self.mem[(rank, ring, g % DEPTH, region)]["writing"] = False

    def begin_read(self, who, rank, ring, g, region, writer):
        key = (rank, ring, g % DEPTH, region)
        cell = self.mem.get(key)
        if cell is None or cell["tile"] != g or cell["writer"] != writer or cell["writing"]:
            raise ProtocolError(f"rank {who} reads {key} for tile {g} from writer {writer} but finds {cell}")
        cell["reading"] += 1

    def end_read(self, who, rank, ring, g, region):
        cell = self.mem[(rank, ring, g % DEPTH, region)]
        if cell["tile"] != g:
            raise ProtocolError(f"data of tile {g} in {(rank, ring, region)} was replaced (tile "
                                f"{cell['tile']}) while rank {who} was reading it")
        cell["reading"] -= 1
        cell["readers_left"].discard(who)


def _expected_readers(coll: Collective, n: int):
    """region written at (rank) -> set of ranks that will read it, derived from the stages."""
    readers: Dict[Tuple[int, Region], set] = {}
    for r in range(n):
        for st in (coll.mid(r), coll.out(r)):
            if st is None:
                continue
            for (rank, region, _w) in st.reads:
                readers.setdefault((rank, region), set()).add(r)
    return readers


def cta(machine: Machine, coll: Collective, rank: int, ring: int, ntiles: int, teams: bool = True,
        early_ready: bool = False) -> Generator[None, None, None]:
    """One CTA of one rank running one collective call (mirrors run_pipeline / run_ll).
    Yields whenever another CTA may be scheduled; a wait yields until its flags are there."""
    n = machine.n
    g0 = machine.epoch.get((rank, ring), 0)
    readers = _expected_readers(coll, n)
    everyone = _all(n)

    def wait(chan, val, srcs):
        if chan == DONE and machine.no_guard:
            return
        while any(machine.get_flag(chan, rank, ring, s) < val for s in srcs):
            yield

    def run_stage(st: Optional[Stage], g: int):
        """begin/end events of one stage, as a sub-generator"""
        if st is None:
            return
        for (rk, region, writer) in st.reads:
            machine.begin_read(rank, rk, ring, g, region, writer)
        for (rk, region) in st.writes:
            machine.begin_write(rank, rk, ring, g, region, readers.get((rk, region), set()))
        yield
        for (rk, region, _w) in st.reads:
            machine.end_read(rank, rk, ring, g, region)
        for (rk, region) in st.writes:
            machine.end_write(rk, ring, g, region)

    def interleave(gens):
        gens = [x for x in gens if x is not None]
        while gens:
            for x in list(gens):
                try:
                    next(x)
                except StopIteration:
                    gens.remove(x)
            yield

    def post(chan, val, dsts):
        for d in dsts:
            machine.post(chan, d, ring, rank, val)

    if coll.ll:
        for t in range(ntiles):
            g = g0 + t
            yield from wait(DONE, g - DEPTH + 1, everyone)
            # push own packets, then poll the peers' packets: a read only starts once the
            # packet of THIS tile is there (the flag travels inside the data)
            yield from run_stage(coll.in_(rank), g)
            st = coll.mid(rank)
            if st is not None:
                for (rk, region, writer) in st.reads:
                    while True:
                        cell = machine.mem.get((rk, ring, g % DEPTH, region))
                        if cell is not None and cell["tile"] == g and not cell["writing"]:
                            break
                        yield
                yield from run_stage(st, g)
            post(DONE, g + 1, everyone)
            yield
        machine.epoch[(rank, ring)] = g0 + ntiles
        return

    has_in, has_mid = coll.in_(rank) is not None, coll.mid(rank) is not None
    has_out = coll.has_out and coll.out(rank) is not None
    yield from wait(DONE, g0 - DEPTH + 1, everyone)
    if has_in:
        yield from run_stage(coll.in_(rank), g0)
    post(READY, g0 + 1, coll.ready_to(rank))
    yield
    for t in range(ntiles):
        g = g0 + t
        more = t + 1 < ntiles
        if more:
            yield from wait(DONE, g + 1 - DEPTH + 1, everyone)
        yield from wait(READY, g + 1, coll.ready_from(rank))
        if coll.has_out and t >= 1:
            yield from wait(MID, g, everyone)
        do_in, do_out = more and has_in, has_out and t >= 1
        def in_then_ready():
            # candidate for round 2: team A publishes READY(t+1) as soon as its own stores are
            # out, instead of at the end of the iteration (after MID(t) of team B)
            yield from run_stage(coll.in_(rank), g + 1)
            post(READY, g + 2, coll.ready_to(rank))

        a = ([in_then_ready()] if early_ready else [run_stage(coll.in_(rank), g + 1)]) if do_in else []
        b = [run_stage(coll.mid(rank), g)] if has_mid else []
        o = [run_stage(coll.out(rank), g - 1)] if do_out else []
        if teams and (len(a) + len(b) + len(o)) >= 2:
            def chain(gs):
                for x in gs:
                    yield from x
            if do_in and has_mid:
                yield from interleave([chain(a + o), chain(b)])
            elif do_in:
                yield from interleave([chain(a), chain(o)])
            else:
                yield from interleave([chain(o), chain(b)])
        else:
            for x in a + b + o:
                yield from x
        if more and not (early_ready and do_in):
            post(READY, g + 2, coll.ready_to(rank))
        if coll.has_out:
            post(MID, g + 1, everyone)
            if t >= 1:
                post(DONE, g, everyone)
        else:
            post(DONE, g + 1, everyone)
        yield
    if coll.has_out:
        yield from wait(MID, g0 + ntiles, everyone)
        if has_out:
            yield from run_stage(coll.out(rank), g0 + ntiles - 1)
        post(DONE, g0 + ntiles, everyone)
        yield
    machine.epoch[(rank, ring)] = g0 + ntiles


def cta_v4(machine: Machine, coll: Collective, rank: int, ring: int, ntiles: int,
           rng: random.Random) -> Generator[None, None, None]:
    """One CTA of pipeline v4 (round 2, csrc/kernels.cuh): every stage that has work on this
    rank is an agent of its own that only waits for its input flags; a publisher agent raises
    the flags some time AFTER the stage finished (its system fence takes 10-25 us), for all
    stages at once, and raises the flag of a stage without work on behalf of the nearest
    earlier stage that has work. The first DEPTH tiles of a call wait until every rank has
    finished everything that came before on the ring."""
    if coll.ll:
        yield from cta(machine, coll, rank, ring, ntiles)
        return
    n = machine.n
    g0 = machine.epoch.get((rank, ring), 0)
    readers = _expected_readers(coll, n)
    everyone = _all(n)
    fns = [coll.in_, coll.mid, coll.out if coll.has_out else (lambda r: None)]
    has = [f(rank) is not None for f in fns]
    if not any(has):
        has[0] = True  # a rank with no work still takes part in the flag chain
    finished = [0, 0, 0]
    # v4 rule (kernels.cuh, ready_from()): a rank whose IN stage touches no slot raises no
    # READY, so nobody lists it (in the kernels the root of a gather does have an IN stage, a
    # local copy outside the slots; leaving it out here is the same thing)
    ready_from = [s for s in coll.ready_from(rank) if coll.in_(s) is not None]

    def wait(chan, val, srcs):
        if chan == DONE and machine.no_guard:
            return
        while any(machine.get_flag(chan, rank, ring, s) < val for s in srcs):
            yield

    def run_stage(st, g):
        if st is None:
            return
        for (rk, region, writer) in st.reads:
            machine.begin_read(rank, rk, ring, g, region, writer)
        for (rk, region) in st.writes:
            machine.begin_write(rank, rk, ring, g, region, readers.get((rk, region), set()))
        yield
        for (rk, region, _w) in st.reads:
            machine.end_read(rank, rk, ring, g, region)
        for (rk, region) in st.writes:
            machine.end_write(rk, ring, g, region)

    def agent(stage):
        for t in range(ntiles):
            g = g0 + t
            if stage == 0:
                yield from wait(DONE, g0 if t < DEPTH else g - DEPTH + 1, everyone)
            elif stage == 1:
                yield from wait(READY, g + 1, ready_from)
            else:
                yield from wait(MID, g + 1, everyone)
            yield from run_stage(fns[stage](rank), g)
            finished[stage] = t + 1
            yield

    def publisher():
        pub = [0, 0, 0]
        out_on = coll.has_out and has[2]
        while any(has[s] and pub[s] < ntiles for s in range(3)):
            snap = list(finished)
            for _ in range(rng.randint(0, 3)):  # the fence: flags lag the data
                yield
            def post(chan, val, dsts):
                for d in dsts:
                    machine.post(chan, d, ring, rank, val)
            if has[0] and snap[0] > pub[0]:
                v = g0 + snap[0]
                post(READY, v, coll.ready_to(rank))
                if coll.has_out and not has[1]:
                    post(MID, v, everyone)
                if not has[1] and not out_on:
                    post(DONE, v, everyone)
                pub[0] = snap[0]
            if has[1] and snap[1] > pub[1]:
                v = g0 + snap[1]
                if coll.has_out:
                    post(MID, v, everyone)
                if not out_on:
                    post(DONE, v, everyone)
                pub[1] = snap[1]
            if out_on and snap[2] > pub[2]:
                post(DONE, g0 + snap[2], everyone)
                pub[2] = snap[2]
            yield

    gens = [agent(s) for s in range(3) if has[s] and (s < 2 or coll.has_out)] + [publisher()]
    while gens:
        x = rng.choice(gens)
        try:
            next(x)
        except StopIteration:
            gens.remove(x)
        yield
    machine.epoch[(rank, ring)] = g0 + ntiles


def run_sequence(n: int, rings: int, calls: List[Tuple[Collective, int, int]], seed: int,
                 teams: bool = True, max_steps: int = 2_000_000, no_guard: bool = False,
                 early_ready: bool = False, pipeline: str = "v2") -> int:
    """Runs `calls` = [(collective, ntiles, nrings_used)] back to back on every rank with a random
    scheduler. A rank starts call k+1 on a ring as soon as ITS call k on that ring is done (stream
    order), regardless of the other ranks. Returns the number of scheduler steps."""
    rng = random.Random(seed)
    m = Machine(n, rings)
    m.no_guard = no_guard
    # one queue of generators per (rank, ring): stream order within a rank
    queues: Dict[Tuple[int, int], List] = {(r, b): [] for r in range(n) for b in range(rings)}
    for coll, ntiles, used in calls:
        for r in range(n):
            for b in range(used):
                queues[(r, b)].append((coll, ntiles))
    active: Dict[Tuple[int, int], Optional[Generator]] = {k: None for k in queues}
    steps = 0
    idle_rounds = 0
    while True:
        keys = [k for k in queues if active[k] is not None or queues[k]]
        if not keys:
            return steps
        k = rng.choice(keys)
        if active[k] is None:
            coll, ntiles = queues[k].pop(0)
            active[k] = (cta_v4(m, coll, k[0], k[1], ntiles, rng) if pipeline == "v4"
                         else cta(m, coll, k[0], k[1], ntiles, teams, early_ready))
        before = (dict(m.flag), {kk: (v["tile"], v["writing"], v["reading"]) for kk, v in m.mem.items()})
        try:
            next(active[k])
        except StopIteration:
            active[k] = None
        after = (m.flag, {kk: (v["tile"], v["writing"], v["reading"]) for kk, v in m.mem.items()})
        steps += 1
        idle_rounds = idle_rounds + 1 if before == after and active[k] is not None else 0
        if idle_rounds > 50 * len(queues) + 1000:
            raise ProtocolError(f"no progress for {idle_rounds} steps: deadlock "
                                f"(pending: {[kk for kk in keys][:8]})")
        if steps > max_steps:
            raise ProtocolError("step budget exceeded")


def run_sym_barriers(n: int, ncalls: int, seed: int) -> int:
    """Model of the two arrival-counter barriers of the zero-copy kernels (csrc/sym.cuh) for
    ONE CTA index: per call every rank adds 1 to the `enter` counter of all ranks
    (multimem.red.add), waits until its own copy reaches (epoch + 1) * n, works, does the same
    on the `exit` counter, then bumps its local epoch. Counters never reset. Checked under a
    random schedule: nobody works in call k before every rank has entered call k, and nobody
    leaves call k before every rank has finished working in it. Returns the steps taken."""
    rng = random.Random(seed)
    counter = [[0, 0] for _ in range(n)]          # [rank][phase]
    entered = [0] * n                             # calls rank r has entered (arrived at enter)
    worked = [0] * n                              # calls rank r has finished working in

    def rank_prog(r):
        for k in range(ncalls):
            for q in range(n):
                counter[q][0] += 1
            entered[r] = k + 1
            yield
            while counter[r][0] < (k + 1) * n:
                yield
            if min(entered) < k + 1:
                raise ProtocolError(f"rank {r} works in call {k} before everyone entered it")
            yield                                  # the work
            worked[r] = k + 1
            for q in range(n):
                counter[q][1] += 1
            yield
            while counter[r][1] < (k + 1) * n:
                yield
            if min(worked) < k + 1:
                raise ProtocolError(f"rank {r} leaves call {k} before everyone finished it")
    gens = [rank_prog(r) for r in range(n)]
    steps = 0
    while gens:
        g = rng.choice(gens)
        try:
            next(g)
        except StopIteration:
            gens.remove(g)
        steps += 1
        if steps > 1_000_000:
            raise ProtocolError("sym barrier model: no progress")
    return steps
