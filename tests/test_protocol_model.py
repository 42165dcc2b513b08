"""A small model checker for the gradient-buffer protocol of the fused allreduce+Adam kernels (csrc/comm.cu).

compute-sanitizer is closed on the GPU pool, so the race argument is checked three ways: the skewed-rank stress on real
NVLink (tests/workers/gpu_multi_worker.py), the emulated-rank stress on one GPU (test_gpu_comm_emulated.py), and — here,
on the CPU — an interleaving model of the protocol itself:

    step t, parity p = t & 1, on every rank r:
      1. backward accumulates this step's gradient into grad[r][p]        (buffer must be clean)
      2. barrier(t): store epoch t+1 into every peer's flag slot, spin until every peer's epoch has arrived
      3. reduce: read grad[q][p] of every rank q                           (must hold exactly q's gradient of step t)
      4. zero grad[r][1 - p]  — the buffer step t-1 used, which step t+1 will accumulate into

ONE barrier per step is enough only because of the double buffering: passing barrier(t) proves every peer has finished
step t-1's reads, so the old buffer may be cleared, while peers that are already a step ahead write the other parity.
The model runs the ranks as coroutines under a random scheduler (thousands of interleavings, ranks drifting up to a full
step apart) and checks the invariants; the mutants below (single buffer, clearing before the barrier, no barrier) must
each be caught, which is what shows the checker can see a race at all.
"""
import random

import pytest


class Violation(AssertionError):
    pass


def rank_proc(r, n, steps, mem, variant):
    """Coroutine: yields after every shared-memory micro-operation so the scheduler can interleave ranks anywhere."""
    nbuf = 1 if variant == "single_buffer" else 2
    for t in range(steps):
        p = t % nbuf
        if variant == "zero_before_barrier" and t > 0:
            mem["grad"][r][(t - 1) % nbuf] = None                     # BUG under test: clears while peers may still read
            yield
        # 1. backward: accumulate into a clean buffer (two partial writes, like several wgrad kernels)
        if mem["grad"][r][p] is not None:
            raise Violation(f"rank {r} step {t}: accumulating into a dirty buffer {mem['grad'][r][p]}")
        mem["grad"][r][p] = ("partial", r, t)
        yield
        mem["grad"][r][p] = ("full", r, t)
        yield
        # 2. barrier: release-store the epoch to every peer, then acquire-spin on our own slots
        if variant != "no_barrier":
            for q in range(n):
                mem["flag"][q][r] = t + 1
                yield
            while any(mem["flag"][r][q] < t + 1 for q in range(n)):
                yield
        # 3. reduce: every peer's buffer of this parity must hold exactly that peer's gradient of this step
        for q in range(n):
            got = mem["grad"][q][p]
            if got != ("full", q, t):
                raise Violation(f"rank {r} step {t}: read {got} from rank {q}")
            yield
        mem["reduced"][r] = t + 1
        # 4. clear the buffer of the previous step for the step after this one
        if variant not in ("zero_before_barrier",) and nbuf == 2:
            mem["grad"][r][1 - p] = None
            yield
        if nbuf == 1:
            mem["grad"][r][p] = None                                   # single buffer: must clear the one it just used
            yield


def run(n, steps, seed, variant="correct"):
    rng = random.Random(seed)
    mem = {"grad": [[None, None] for _ in range(n)], "flag": [[0] * n for _ in range(n)], "reduced": [0] * n}
    procs = {r: rank_proc(r, n, steps, mem, variant) for r in range(n)}
    max_skew = 0
    # a biased scheduler: one rank is "slow" for a while, then another, so ranks really drift apart
    slow, ticks = rng.randrange(n), 0
    while procs:
        ticks += 1
        if ticks % 37 == 0:
            slow = rng.randrange(n)
        live = list(procs)
        r = rng.choice(live)
        if r == slow and len(live) > 1 and rng.random() < 0.85:
            continue
        try:
            next(procs[r])
        except StopIteration:
            del procs[r]
        done = mem["reduced"]
        max_skew = max(max_skew, max(done) - min(done))
        if ticks > 2_000_000:
            raise RuntimeError("model deadlocked")
    return max_skew


@pytest.mark.parametrize("n", [2, 4, 8])
def test_one_barrier_with_double_buffering_is_race_free(n):
    worst = 0
    for seed in range(300):
        worst = max(worst, run(n, steps=12, seed=seed))
    assert worst >= 1            # the schedules really let ranks run a step apart (otherwise the check proves nothing)


@pytest.mark.parametrize("variant", ["single_buffer", "zero_before_barrier", "no_barrier"])
def test_the_checker_catches_broken_protocols(variant):
    caught = 0
    for seed in range(200):
        try:
            run(4, steps=8, seed=seed, variant=variant)
        except Violation:
            caught += 1
    assert caught > 0, f"mutant {variant} survived every schedule"


# ---------------------------------------------------------------------------------------------------------------------
# Two-shot bucket: owner-sharded Adam, updated bf16 weights PUSHED into every peer's working copy.
#
#   step t on rank r:
#     forward/backward READ w16[r][*]  (all slices)           — must see version t everywhere
#     barrier A(t)   (launched only after this rank's last read of those weights: the dense dgrad)
#     owner work: reduce slice r, Adam -> version t+1, store w16[q][r] = t+1 into EVERY rank q (peer stores)
#     barrier B(t)
#   then step t+1 reads version t+1.
# Hazards: a fast rank's push landing in a slow rank's working copy while that rank still reads version t (prevented by
# A), and a rank starting step t+1 before every owner's push has landed (prevented by B).
# ---------------------------------------------------------------------------------------------------------------------
def twoshot_proc(r, n, steps, mem, variant):
    def barrier(name, t):
        for q in range(n):
            mem[name][q][r] = t + 1
            yield
        while any(mem[name][r][q] < t + 1 for q in range(n)):
            yield

    for t in range(steps):
        # forward + backward: read every slice of the local working copy, twice (fwd and dgrad), with yields in between
        for _ in range(2):
            for s in range(n):
                if mem["w16"][r][s] != t:
                    raise Violation(f"rank {r} step {t}: slice {s} of its working weights is version {mem['w16'][r][s]}")
                yield
        if variant != "push_before_A":
            yield from barrier("A", t)
        for q in range(n):                                   # owner r: push its updated slice to everyone
            mem["w16"][q][r] = t + 1
            yield
        if variant == "push_before_A":
            yield from barrier("A", t)
        if variant != "no_B":
            yield from barrier("B", t)
        mem["reduced"][r] = t + 1


def run_twoshot(n, steps, seed, variant="correct"):
    rng = random.Random(seed)
    mem = {"w16": [[0] * n for _ in range(n)], "A": [[0] * n for _ in range(n)], "B": [[0] * n for _ in range(n)], "reduced": [0] * n}
    procs = {r: twoshot_proc(r, n, steps, mem, variant) for r in range(n)}
    slow, ticks = rng.randrange(n), 0
    while procs:
        ticks += 1
        if ticks % 41 == 0:
            slow = rng.randrange(n)
        r = rng.choice(list(procs))
        if r == slow and len(procs) > 1 and rng.random() < 0.85:
            continue
        try:
            next(procs[r])
        except StopIteration:
            del procs[r]
        if ticks > 2_000_000:
            raise RuntimeError("model deadlocked")


@pytest.mark.parametrize("n", [2, 4, 8])
def test_twoshot_weight_push_is_race_free_with_its_two_barriers(n):
    for seed in range(200):
        run_twoshot(n, steps=8, seed=seed)


@pytest.mark.parametrize("variant", ["push_before_A", "no_B"])
def test_twoshot_mutants_are_caught(variant):
    caught = 0
    for seed in range(200):
        try:
            run_twoshot(4, steps=6, seed=seed, variant=variant)
        except Violation:
            caught += 1
    assert caught > 0, f"mutant {variant} survived every schedule"


# ---------------------------------------------------------------------------------------------------------------------
# The fused dense step (round 2; b200dp/models/b200_engine.py::_backward_fused): two barriers per step.
#
#   step t on rank r:
#     forward  : write act[r] = (r, t)  (a2 and dz: what the peers will gather); READ every slice of w16[r]  -> version t
#     dgrad    : READ every slice of w16[r] again                                                             -> version t
#     barrier G(t)   — launched only after this rank's dgrad
#     gather   : READ act[q] of every rank q                      -> must be exactly (q, t)
#     GEMM+Adam: owner r stores w16[q][r] = t+1 into EVERY rank q (TMA stores into peer memory)
#     closing collective, push form: store my small gradient into inbox[q][t & 1][r] of every peer q, THEN raise flag E(t);
#                wait for every peer's E(t); READ inbox[r][t & 1][q] -> must be exactly (q, t)
#   then step t+1 overwrites act[r] and reads version t+1 of the weights.
# Hazards and what prevents them: a push landing while the target still reads version t (G: arrival after the dgrad);
# act[r] overwritten while a peer still gathers it, or step t+1 starting before every owner's push landed (E: arrival after
# the own gather and the own pushes); an inbox row overwritten before it was consumed (two parities + E of the step between).
# ---------------------------------------------------------------------------------------------------------------------
def fused_proc(r, n, steps, mem, variant):
    def arrive(name, t):
        for q in range(n):
            mem[name][q][r] = t + 1
            yield

    def wait(name, t):
        while any(mem[name][r][q] < t + 1 for q in range(n)):
            yield

    nin = 1 if variant == "single_inbox" else 2
    for t in range(steps):
        mem["act"][r] = ("partial", r, t)                    # conv2 forward / head write a2, dz
        yield
        mem["act"][r] = (r, t)
        yield
        for phase in range(2):                               # dense forward, then dense dgrad: both read the working weights
            if variant == "gather_before_dgrad" and phase == 1:
                yield from arrive("G", t)                    # BUG under test: tells the peers "done with the old weights" too early
            for s in range(n):
                if mem["w16"][r][s] != t:
                    raise Violation(f"rank {r} step {t}: slice {s} of its working weights is version {mem['w16'][r][s]}")
                yield
        if variant != "gather_before_dgrad":
            yield from arrive("G", t)
        yield from wait("G", t)
        for q in range(n):                                   # gather
            if mem["act"][q] != (q, t):
                raise Violation(f"rank {r} step {t}: gathered {mem['act'][q]} from rank {q}")
            yield
        for q in range(n):                                   # owner r: weight push
            mem["w16"][q][r] = t + 1
            yield
        for q in range(n):                                   # closing collective: inbox stores first ...
            if q != r:
                mem["inbox"][q][t % nin][r] = (r, t)
                yield
        if variant != "no_closing_barrier":
            yield from arrive("E", t)                        # ... then the flag
            yield from wait("E", t)
            for q in range(n):
                if q != r and mem["inbox"][r][t % nin][q] != (q, t):
                    raise Violation(f"rank {r} step {t}: inbox row of rank {q} holds {mem['inbox'][r][t % nin][q]}")
                yield
        mem["reduced"][r] = t + 1


def run_fused(n, steps, seed, variant="correct"):
    rng = random.Random(seed)
    mem = {"act": [None] * n, "w16": [[0] * n for _ in range(n)], "G": [[0] * n for _ in range(n)], "E": [[0] * n for _ in range(n)],
           "inbox": [[[None] * n for _ in range(2)] for _ in range(n)], "reduced": [0] * n}
    procs = {r: fused_proc(r, n, steps, mem, variant) for r in range(n)}
    slow, ticks = rng.randrange(n), 0
    while procs:
        ticks += 1
        if ticks % 43 == 0:
            slow = rng.randrange(n)
        r = rng.choice(list(procs))
        if r == slow and len(procs) > 1 and rng.random() < 0.85:
            continue
        try:
            next(procs[r])
        except StopIteration:
            del procs[r]
        if ticks > 3_000_000:
            raise RuntimeError("model deadlocked")


@pytest.mark.parametrize("n", [2, 4, 8])
def test_fused_dense_step_is_race_free_with_gather_and_closing_barriers(n):
    for seed in range(150):
        run_fused(n, steps=8, seed=seed)


def test_fused_step_needs_no_second_inbox_but_the_general_entry_point_does():
    """With TWO barriers per step a single inbox would do: a peer reaches its next inbox store only through G(t+1), which
    this rank joins after it has consumed inbox(t).  The inbox is double-buffered anyway because b200_comm_set_inbox serves
    any one-shot launch, and with ONE barrier per step (the plain gradient-buffer protocol at the top of this file) a
    fast rank's stores of step t+1 would land in rows a slow rank has not summed yet."""
    for seed in range(150):
        run_fused(4, steps=6, seed=seed, variant="single_inbox")


@pytest.mark.parametrize("variant", ["gather_before_dgrad", "no_closing_barrier"])
def test_fused_step_mutants_are_caught(variant):
    caught = 0
    for seed in range(300):
        try:
            run_fused(4, steps=6, seed=seed, variant=variant)
        except Violation:
            caught += 1
    assert caught > 0, f"mutant {variant} survived every schedule"
