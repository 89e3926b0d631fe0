"""Randomised-schedule check of the ring / flag protocol on the CPU:
the executable model in pdcc_b200/utils/protocol_model.py mirrors run_pipeline() and the LL
kernels; here every collective, alone and in mixed back-to-back sequences with different tile
counts, must terminate and never read stale data or overwrite unread data. A fault-injection
case (slot guard removed) shows that the checker does catch such a bug."""
import random

import pytest

from pdcc_b200.utils import protocol_model as pm


def _collectives(n):
    return [pm.all_reduce_oneshot(n), pm.all_reduce_twoshot(n), pm.all_reduce_nvls(n),
            pm.reduce_rooted_twophase(n, 0), pm.reduce_rooted_twophase(n, n - 1),
            pm.reduce_oneshot(n, 1 % n), pm.broadcast_push(n, 0), pm.broadcast_twophase(n, n - 1),
            pm.scatter(n, 0), pm.gather(n, n - 1), pm.all_gather_push(n), pm.all_gather_pull(n),
            pm.ll_all_reduce(n)]


@pytest.mark.parametrize("n", [2, 3, 4])
def test_each_collective_alone_many_tiles(n):
    for coll in _collectives(n):
        for ntiles in (1, 2, 5, 9):  # 9 > 2 x DEPTH: every slot is recycled twice
            for seed in range(3):
                pm.run_sequence(n, 1, [(coll, ntiles, 1)], seed)


@pytest.mark.parametrize("teams", [True, False])
def test_mixed_sequences_share_the_rings(teams):
    n = 3
    rng = random.Random(7)
    colls = _collectives(n)
    for seed in range(12):
        calls = [(rng.choice(colls), rng.choice([1, 1, 2, 3, 6]), rng.choice([1, 2])) for _ in range(10)]
        pm.run_sequence(n, 2, calls, seed, teams=teams)


def test_checker_catches_a_missing_slot_guard():
    n = 3
    caught = 0
    for seed in range(40):
        try:
            pm.run_sequence(n, 1, [(pm.all_reduce_twoshot(n), 9, 1), (pm.broadcast_push(n, 0), 9, 1)],
                            seed, no_guard=True)
        except pm.ProtocolError:
            caught += 1
    assert caught > 0, "removing the DONE(g - DEPTH) guard must produce a detectable hazard"


def test_early_ready_variant_is_also_safe():
    """Design candidate for round 2: the pushing team posts READY(t+1) right
    after its own stores, not at the end of the iteration. Checked here before it is built."""
    n = 4
    rng = random.Random(11)
    colls = _collectives(n)
    for seed in range(8):
        calls = [(rng.choice(colls), rng.choice([1, 2, 4, 7]), 1) for _ in range(8)]
        pm.run_sequence(n, 1, calls, seed, early_ready=True)


@pytest.mark.parametrize("n", [2, 3, 4])
def test_pipeline_v4_independent_stages_and_late_flags(n):
    """The round-2 kernels (csrc/kernels.cuh, pipeline v4): one agent per stage, flags raised
    late by a publisher, idle stages' flags raised by the stage before them. Alone with slot
    reuse, and in mixed sequences that also contain LL calls."""
    colls = _collectives(n)
    for coll in colls:
        for ntiles in (1, 3, 9):
            for seed in range(3):
                pm.run_sequence(n, 1, [(coll, ntiles, 1)], seed, pipeline="v4")
    rng = random.Random(5)
    for seed in range(10):
        calls = [(rng.choice(colls), rng.choice([1, 2, 3, 6, 9]), rng.choice([1, 2])) for _ in range(10)]
        pm.run_sequence(n, 2, calls, seed, pipeline="v4")


def test_pipeline_v4_checker_catches_a_missing_slot_guard():
    n, caught = 3, 0
    for seed in range(40):
        try:
            pm.run_sequence(n, 1, [(pm.all_reduce_twoshot(n), 9, 1), (pm.broadcast_push(n, 0), 9, 1)],
                            seed, no_guard=True, pipeline="v4")
        except pm.ProtocolError:
            caught += 1
    assert caught > 0


@pytest.mark.parametrize("n", [2, 4, 8])
def test_zero_copy_arrival_counter_barriers(n):
    """csrc/sym.cuh: never-reset arrival counters with target (epoch + 1) * n keep calls apart."""
    for seed in range(20):
        pm.run_sym_barriers(n, 6, seed)
