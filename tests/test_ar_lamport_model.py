"""Model check of the FLAG-FREE decode all-reduce protocol (csrc/allreduce_norm.cu: ar_lamport_add_norm_kernel).

What the kernel does per (rank, row): write its folded row, word by word, into slot [parity][me][row] of every
rank (one-shot) or of the row's owner (two-shot); then, word by word, poll the words it needs until none holds
the SENTINEL, add them in rank order, and write the sentinel back.  No flags: a word either is the sentinel or
is the value of exactly one call.  Checked here with random interleavings of single-WORD operations across
ranks (so a row may be half arrived, and different rows / ranks may be calls apart as far as the protocol lets
them): every word a reader takes must carry the tag of the call it is in (no stale word, no word of a later
call), nothing may be overwritten before it was consumed, and nothing may dead-lock.  The two-parity argument
of the kernel's header is the invariant `a rank is never more than one call ahead`; the negative control shows
the checker catches a single-buffered variant.  Memory is sequentially consistent here; on the GPU every word is
self-validating, so no ordering between different words is assumed by the protocol in the first place.
"""
import random

import pytest

SENTINEL = None
WORDS = 3          # words per row: enough to see a row partly arrived


class Rank:
    def __init__(self, tp, rows, n_par):
        self.slot = {(p, s, r, w): SENTINEL for p in range(n_par) for s in range(tp) for r in range(rows) for w in range(WORDS)}


def put(world, at, key, tag):
    assert world[at].slot[key] is SENTINEL, f"overwrote an unconsumed word {key} at rank {at}: {world[at].slot[key]} <- {tag}"
    world[at].slot[key] = tag


def take(world, me, key, want):
    """Generator: poll one word until it is not the sentinel, check it, clear it."""
    while world[me].slot[key] is SENTINEL:
        yield
    got = world[me].slot[key]
    assert got == want, f"rank {me} took {got} at {key}, expected {want}"
    world[me].slot[key] = SENTINEL
    yield


def cta_program(world, me, call, row, two_shot, result, n_par):
    tp = len(world)
    par = call % n_par
    if not two_shot:
        for w in range(WORDS):                      # the row leaves word by word, rank by rank
            for r in range(tp):
                if r != me:
                    put(world, r, (par, me, row, w), ("partial", call, me, w))
                    yield
        for w in range(WORDS):
            for src in range(tp):
                if src != me:
                    yield from take(world, me, (par, src, row, w), ("partial", call, src, w))
        result[(me, call, row)] = True
        return
    owner = row % tp
    if me != owner:
        for w in range(WORDS):
            put(world, owner, (par, me, row, w), ("partial", call, me, w))
            yield
        for w in range(WORDS):
            yield from take(world, me, (par, owner, row, w), ("sum", call, owner, w))
        result[(me, call, row)] = True
        return
    for w in range(WORDS):                          # the owner reduces word by word and sends each sum back at once
        for src in range(tp):
            if src != me:
                yield from take(world, me, (par, src, row, w), ("partial", call, src, w))
        for r in range(tp):
            if r != me:
                put(world, r, (par, me, row, w), ("sum", call, me, w))
                yield
    result[(me, call, row)] = True


def rank_program(world, me, n_calls, rows, two_shot, result, rng, n_par):
    for call in range(n_calls):                     # stream order: call c+1 starts when every row of call c is done
        mode = two_shot(call) if callable(two_shot) else two_shot
        ctas = [cta_program(world, me, call, row, mode, result, n_par) for row in range(rows)]
        while ctas:
            g = rng.choice(ctas)
            try:
                next(g)
            except StopIteration:
                ctas.remove(g)
            yield


def run(tp, rows, n_calls, two_shot, seed, n_par=2, bias=True):
    rng = random.Random(seed)
    world = [Rank(tp, rows, n_par) for _ in range(tp)]
    result = {}
    progs = {r: rank_program(world, r, n_calls, rows, two_shot, result, rng, n_par) for r in range(tp)}
    steps = 0
    while progs:
        weights = [1 + (9 if bias and (steps // 40 + r) % tp == 0 else 0) for r in progs]
        r = rng.choices(list(progs), weights)[0]
        try:
            next(progs[r])
        except StopIteration:
            del progs[r]
        steps += 1
        assert steps < 3_000_000, "protocol dead-locked"
    assert len(result) == tp * n_calls * rows
    for rk in world:                                # every slot is back to the sentinel: nothing left unconsumed
        assert all(v is SENTINEL for v in rk.slot.values())


@pytest.mark.parametrize("two_shot", [False, True])
@pytest.mark.parametrize("tp,rows", [(2, 3), (4, 5), (8, 9), (4, 2)])
def test_random_interleavings(tp, rows, two_shot):
    for seed in range(10):
        run(tp, rows, 7, two_shot, seed * 31 + tp)


def test_modes_may_alternate_from_call_to_call():
    for pattern in (lambda c: c % 2 == 0, lambda c: c % 3 == 0, lambda c: c in (1, 2, 5)):
        for seed in range(6):
            run(4, 5, 9, pattern, seed)


def test_single_buffered_variant_is_caught():
    """Negative control: with ONE parity a fast rank's call c+1 overwrites words a slow rank has not consumed."""
    caught = 0
    for seed in range(20):
        try:
            run(4, 3, 6, False, seed, n_par=1)
        except AssertionError:
            caught += 1
    assert caught > 0
