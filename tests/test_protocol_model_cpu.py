# -*- coding: utf-8 -*-
"""Model of the epoch / parity protocol of the fused collective kernels, run under random
schedules: the reference serialises every op with a host MPI barrier (functions.py:62, 121, 179); here nothing but
sequence numbers, two workspace parities and ``done`` counters keeps a fast rank from overwriting what a slow rank
still reads.  The model has the same rules as the kernels:

* call ``k`` uses workspace parity ``k & 1``;
* a rank pushes (gathered rows, partial sums) into a peer's workspace only after that peer's ``done >= k - 2``;
* a gathering call finishes when the pushes of call ``k`` of every peer have landed, a reducing call when every
  peer's partial sums of call ``k`` have landed; finishing publishes ``done = k``;
* gather reuse (ops/attention.py::gather_reusable): the row-side backward at ``f + 1`` reads the gather of the forward
  ``f`` in the OTHER parity's workspace and pushes nothing, the column-side backward at ``f + 2`` reads it in its own
  parity and pushes only partial sums.

The checked invariant: whenever a rank reads a workspace region, every peer's slot still holds the data of the call
the reader expects.  A negative control (reuse by the row-side kernel alone, followed by a pushing call) must be
caught.
"""
import random

import pytest

AG, ROWS_REUSE, COLS, COLS_REUSE = 'ag', 'rows_reuse', 'cols', 'cols_reuse'


class Violation(Exception):
    pass


class Rank(object):
    def __init__(self, r, world):
        self.r = r
        self.k = 0                 # call in flight (0: none yet)
        self.started = False
        self.pushed = set()        # peers served by the call in flight
        self.done = 0
        # slot contents: [parity][source rank] -> sequence number of the call that wrote it
        self.gather = [[0] * world for _ in range(2)]
        self.part = [[0] * world for _ in range(2)]


def simulate(calls, world, seed, wait_lag=2, max_events=200000):
    """``calls[k - 1]`` = kind of call ``k`` (same program on every rank)."""
    rng = random.Random(seed)
    ranks = [Rank(r, world) for r in range(world)]
    n = len(calls)

    def gather_source(k):
        """(parity, sequence number) of the gather that call k reads"""
        kind = calls[k - 1]
        if kind in (AG, COLS):
            return k & 1, k
        if kind == ROWS_REUSE:
            return (k - 1) & 1, k - 1
        return (k - 2) & 1, k - 2          # COLS_REUSE (same parity as k)

    def check_reads(me, k):
        par, want = gather_source(k)
        for q in range(world):
            if me.gather[par][q] != want:
                raise Violation('rank %d, call %d (%s): gathered rows of rank %d are those of call %d, not %d'
                                % (me.r, k, calls[k - 1], q, me.gather[par][q], want))

    for _ in range(max_events):
        live = [x for x in ranks if x.done < n]
        if not live:
            return True
        me = rng.choice(live)
        if not me.started:
            me.k = me.done + 1
            me.started = True
            me.pushed = set()
            kind = calls[me.k - 1]
            if kind in (AG, COLS):
                me.gather[me.k & 1][me.r] = me.k          # own rows are read in place: model them as a local push
            if kind in (COLS, COLS_REUSE):
                me.part[me.k & 1][me.r] = me.k
            if kind in (ROWS_REUSE, COLS_REUSE):
                check_reads(me, me.k)                     # reads start right away ...
            continue
        k, kind, par = me.k, calls[me.k - 1], me.k & 1
        # ---- pushes, one peer per event, each behind the done-counter wait
        todo = [q for q in range(world) if q != me.r and q not in me.pushed]
        if todo:
            q = rng.choice(todo)
            if ranks[q].done >= k - wait_lag:
                if kind in (AG, COLS):
                    ranks[q].gather[par][me.r] = k
                if kind in (COLS, COLS_REUSE):
                    ranks[q].part[par][me.r] = k
                me.pushed.add(q)
            continue
        # ---- completion: everything this call consumes has landed
        par_g, want = gather_source(k)
        if kind in (AG, COLS) and any(me.gather[par_g][q] < want for q in range(world)):
            continue
        if kind in (COLS, COLS_REUSE) and any(me.part[par][q] < k for q in range(world)):
            continue
        check_reads(me, k)                                # ... and last until the call finishes
        if kind in (COLS, COLS_REUSE):
            for q in range(world):
                if me.part[par][q] != k:
                    raise Violation('rank %d, call %d: partial sums of rank %d are those of call %d'
                                    % (me.r, k, q, me.part[par][q]))
        me.done = k
        me.started = False
    raise AssertionError('schedule did not terminate (deadlock in the model)')


def training_steps(steps, reuse):
    """forward, row-side backward, column-side backward per step, as the flagship layer issues them"""
    one = [AG, ROWS_REUSE, COLS_REUSE] if reuse else [AG, AG, COLS]
    return one * steps


@pytest.mark.parametrize('world', [2, 3, 4, 8])
@pytest.mark.parametrize('reuse', [False, True])
def test_protocol_holds_under_random_schedules(world, reuse):
    for seed in range(60):
        assert simulate(training_steps(5, reuse), world, seed)


def test_mixed_programs_hold():
    """two layers: only the last forward's gather is reused (the heap has moved on for the first one)"""
    program = [AG, AG, ROWS_REUSE, COLS_REUSE, AG, COLS] * 3 + [AG] * 3 + [COLS] * 2
    for world in (2, 4):
        for seed in range(60):
            assert simulate(program, world, seed)


def test_model_catches_reuse_by_the_row_side_kernel_alone():
    """Why ``gather_reusable`` insists on both kernels: a row-side call reading the forward's parity at f + 1,
    followed by a call that pushes into that parity at f + 2, only waits for the peers' f."""
    program = [AG, ROWS_REUSE, AG] * 4
    caught = 0
    for seed in range(200):
        try:
            simulate(program, 3, seed)
        except Violation:
            caught += 1
    assert caught > 0


def test_done_wait_is_what_protects_the_readers():
    """The same program is safe once the pushing call waits for the peers' PREVIOUS call (k - 1): the violation
    above comes from the k - 2 wait, not from the model.  (In programs made of gathering / reducing calls only,
    every call needs every peer's push, so no rank ever runs more than one call ahead and the wait never binds.)"""
    program = [AG, ROWS_REUSE, AG] * 4
    for seed in range(100):
        assert simulate(program, 3, seed, wait_lag=1)
    for seed in range(50):
        assert simulate([AG] * 9, 3, seed, wait_lag=9)
