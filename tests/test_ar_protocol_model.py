"""Model check of the buffer / flag protocol of the fused decode all-reduce (csrc/allreduce_norm.cu).

The kernel itself needs GPUs; what is checked here, on the CPU, is the PROTOCOL it implements: receive
slots double-buffered by call parity, flags that carry a monotonically increasing epoch and are never
reset, one-shot (push to all) and two-shot (row r reduced by rank r % tp, result pushed back) modes.
Ranks are little state machines stepped by a random scheduler (any interleaving of pushes, flag stores,
polls and reads across ranks; on one rank call c+1 starts only after call c has finished, which is what
stream order gives the kernel).  Every slot carries a tag (call, source); a reader that sees its epoch
must find exactly the tag it expects -- an early overwrite or a stale read fails the test.  Memory is
sequentially consistent here; ordering of a push before its flag is the kernel's st.release / CTA barrier.
"""
import random

import pytest


class Rank:
    def __init__(self, me, tp, rows):
        self.me, self.tp, self.rows = me, tp, rows
        self.data = {}     # (parity, src, row) -> tag
        self.flag = {}     # (parity, src, row) -> epoch


def cta_program(world, me, call, row, two_shot, result, n_par=2):
    """Generator: one CTA (rank me, row) of all-reduce call `call`; yields after every memory operation."""
    tp = len(world)
    parity, epoch = call % n_par, call + 1
    mine = world[me]
    if not two_shot:
        for r in range(tp):
            world[r].data[(parity, me, row)] = ("partial", call, me)
            yield
        for r in range(tp):
            world[r].flag[(parity, me, row)] = epoch
            yield
        for src in range(tp):
            while mine.flag.get((parity, src, row)) != epoch:
                yield
        got = []
        for src in range(tp):
            got.append(mine.data[(parity, src, row)])
            yield
        assert got == [("partial", call, s) for s in range(tp)], (me, call, row, got)
        result[(me, call, row)] = tuple(got)
        return
    owner = row % tp
    if me != owner:
        world[owner].data[(parity, me, row)] = ("partial", call, me)
        yield
        world[owner].flag[(parity, me, row)] = epoch
        yield
        while mine.flag.get((parity, owner, row)) != epoch:
            yield
        got = mine.data[(parity, owner, row)]
        assert got == ("sum", call, owner), (me, call, row, got)
        result[(me, call, row)] = got
        return
    for src in range(tp):
        if src != me:
            while mine.flag.get((parity, src, row)) != epoch:
                yield
    for src in range(tp):
        if src != me:
            got = mine.data[(parity, src, row)]
            assert got == ("partial", call, src), (me, call, row, src, got)
            yield
    for r in range(tp):
        if r != me:
            world[r].data[(parity, me, row)] = ("sum", call, me)
            yield
    for r in range(tp):
        if r != me:
            world[r].flag[(parity, me, row)] = epoch
            yield
    result[(me, call, row)] = ("sum", call, me)


def rank_program(world, me, n_calls, rows, two_shot, result, rng, n_par=2):
    """All calls of one rank: the CTAs of a call run interleaved; the next call starts when all are done."""
    for call in range(n_calls):
        mode = two_shot(call) if callable(two_shot) else two_shot
        ctas = [cta_program(world, me, call, row, mode, result, n_par) for row in range(rows)]
        while ctas:
            g = rng.choice(ctas)
            try:
                next(g)
            except StopIteration:
                ctas.remove(g)
            yield


@pytest.mark.parametrize("two_shot", [False, True])
@pytest.mark.parametrize("tp,rows", [(2, 3), (4, 5), (8, 9)])
def test_random_interleavings_never_overwrite_or_read_stale(tp, rows, two_shot):
    for seed in range(12):
        rng = random.Random(seed * 100 + tp)
        world = [Rank(r, tp, rows) for r in range(tp)]
        result = {}
        n_calls = 7
        progs = {r: rank_program(world, r, n_calls, rows, two_shot, result, rng) for r in range(tp)}
        # biased scheduler: one rank is starved in bursts, another races ahead as far as the protocol lets it
        steps = 0
        while progs:
            weights = [1 + 9 * ((steps // 50 + r) % tp == 0) for r in progs]
            r = rng.choices(list(progs), weights)[0]
            try:
                next(progs[r])
            except StopIteration:
                del progs[r]
            steps += 1
            assert steps < 2_000_000, "protocol dead-locked"
        assert len(result) == tp * n_calls * rows


def test_modes_may_alternate_from_call_to_call():
    """The mode depends on the batch size, so consecutive calls may differ; they share slots and flags."""
    tp, rows, n_calls = 4, 5, 9
    for pattern in (lambda c: c % 2 == 0, lambda c: c % 3 == 0, lambda c: c in (1, 2, 5)):
        for seed in range(8):
            rng = random.Random(seed)
            world = [Rank(r, tp, rows) for r in range(tp)]
            result = {}
            progs = {r: rank_program(world, r, n_calls, rows, pattern, result, rng) for r in range(tp)}
            while progs:
                r = rng.choice(list(progs))
                try:
                    next(progs[r])
                except StopIteration:
                    del progs[r]
            assert len(result) == tp * n_calls * rows


def test_a_rank_is_never_more_than_one_call_ahead():
    """The invariant two parities rely on: rank A can only finish call c once every rank has STARTED call c."""
    tp, rows, n_calls = 4, 2, 6
    for two_shot in (False, True):
        rng = random.Random(7)
        world = [Rank(r, tp, rows) for r in range(tp)]
        result = {}
        progs = {r: rank_program(world, r, n_calls, rows, two_shot, result, rng) for r in range(tp)}

        def calls_done(r):
            return min((max((c for (rr, c, row) in result if rr == r and row == x), default=-1) for x in range(rows)))

        while progs:
            r = 0 if (0 in progs and rng.random() < 0.9) else rng.choice(list(progs))   # rank 0 races ahead
            try:
                next(progs[r])
            except StopIteration:
                del progs[r]
            done = [calls_done(x) for x in range(tp)]
            assert max(done) - min(done) <= 1, done


def _violations_single_buffered(two_shot):
    tp, rows, n_calls = 4, 3, 6
    caught = 0
    for seed in range(20):
        rng = random.Random(seed)
        world = [Rank(r, tp, rows) for r in range(tp)]
        result = {}
        progs = {r: rank_program(world, r, n_calls, rows, two_shot, result, rng, n_par=1) for r in range(tp)}
        try:
            steps = 0
            while progs and steps < 200_000:
                r = rng.choice(list(progs))
                try:
                    next(progs[r])
                except StopIteration:
                    del progs[r]
                steps += 1
        except AssertionError:
            caught += 1
    return caught


def test_the_checker_catches_a_single_buffered_one_shot():
    """Negative control: WITHOUT the parity double buffer a fast rank overwrites a slot a slow rank has not read
    yet (one-shot: finishing call c only needs everybody's PUSH of call c, not everybody's read)."""
    assert _violations_single_buffered(two_shot=False) > 0


def test_two_shot_would_not_even_need_the_second_buffer():
    """In the two-shot form a rank leaves call c only after the owner has consumed its partial row and pushed the
    sum back, so the same schedules find no hazard even single-buffered; the kernel keeps both parities because
    the two modes share buffers and may alternate from call to call."""
    assert _violations_single_buffered(two_shot=True) == 0
