"""CPU test: a randomized-schedule model of the flag-free ("low-latency push") protocol of the fused tensor-parallel
all-reduce + add + RMSNorm kernel (nano-vllm-rs_b200/csrc/comm.cu, allreduce_add_rmsnorm_ll_kernel).

The kernel has no barrier between ranks: the data is the signal (an unwritten word holds a reserved pattern), buffer pairs
rotate per call and a call clears the pair TWO calls ahead.  Its comment argues that this is race-free because a peer can be
at most one call ahead.  This model checks that argument, not the CUDA code: every rank runs several concurrent "CTAs", each
a sequential program of pushes / clears / polls; remote stores are delivered after arbitrary delays and in arbitrary order;
a scheduler picks the next event at random.  Invariants checked on every event:
  * a poll that succeeds sees the value written for THIS call by the expected source (never a stale one),
  * a remote store never lands on a cell that still holds something (the cell was cleared in time),
  * a clear never wipes a value nobody has consumed yet,
  * nobody deadlocks.
With three pairs every schedule passes; with two pairs (clear the pair of the NEXT call) the model finds the race."""
import random

import pytest

EMPTY = None


class Violation(Exception):
    pass


class Rank:
    def __init__(self, r, world, pairs, max_rows):
        self.r, self.world = r, world
        # stage[pair][src][j] / bcast[pair][row] -> (value, consumed)
        self.stage = [[[EMPTY] * max_rows for _ in range(world)] for _ in range(pairs)]
        self.bcast = [[EMPTY] * (max_rows * world) for _ in range(pairs)]
        self.used = [0] * pairs          # rows per source the last call on that pair touched
        self.call = 0                    # index of the call this rank is executing
        self.ctas = []                   # program counters of the running call


def run(world, pairs, tokens_per_call, n_ctas, seed, clear_ahead):
    rng = random.Random(seed)
    max_rows = max((t + world - 1) // world for t in tokens_per_call)
    ranks = [Rank(r, world, pairs, max_rows) for r in range(world)]
    inflight = []                        # (dst rank, kind, pair, index..., value)

    def program(rank, i):
        """The per-CTA programs of call i on `rank`: lists of ops, rows dealt round-robin to the CTAs (as blockIdx does)."""
        T = tokens_per_call[i]
        R = (T + world - 1) // world
        b = i % pairs
        progs = [[] for _ in range(n_ctas)]
        for row in range(T):             # phase A: push my partial of every foreign row to its owner
            owner = row // R
            if owner != rank.r:
                progs[row % n_ctas].append(("push_stage", owner, b, rank.r, row - owner * R, (i, rank.r, row)))
        bclr = (b + clear_ahead) % pairs
        for c in range(n_ctas):          # housekeeping: clear my copy of the pair `clear_ahead` calls ahead
            progs[c].append(("clear", bclr, c))
        row0 = rank.r * R
        for k, row in enumerate(range(row0, min(T, row0 + R))):   # phase B: reduce the rows I own, broadcast them
            ops = progs[k % n_ctas]
            for q in range(world):
                if q != rank.r:
                    ops.append(("poll_stage", b, q, row - row0, (i, q, row)))
            for q in range(world):
                if q != rank.r:
                    ops.append(("push_bcast", q, b, row, (i, rank.r, row)))
        for row in range(T):             # phase C: gather the rows the others own
            owner = row // R
            if owner != rank.r:
                progs[row % n_ctas].append(("poll_bcast", b, row, (i, owner, row)))
        return progs, R

    for rk in ranks:
        rk.progs, rk.R = program(rk, 0)
        rk.pc = [0] * n_ctas

    steps = 0
    while True:
        steps += 1
        if steps > 2_000_000:
            raise Violation("livelock")
        events = []
        for k in range(len(inflight)):
            events.append(("deliver", k))
        for rk in ranks:
            if rk.call >= len(tokens_per_call):
                continue
            for c in range(n_ctas):
                if rk.pc[c] >= len(rk.progs[c]):
                    continue
                op = rk.progs[c][rk.pc[c]]
                if op[0] == "poll_stage" and rk.stage[op[1]][op[2]][op[3]] is EMPTY:
                    continue             # blocked: spinning on the sentinel
                if op[0] == "poll_bcast" and rk.bcast[op[1]][op[2]] is EMPTY:
                    continue
                events.append(("step", rk.r, c))
        if not events:
            if all(rk.call >= len(tokens_per_call) for rk in ranks):
                return steps
            raise Violation(f"deadlock: calls {[rk.call for rk in ranks]}")
        ev = rng.choice(events)
        if ev[0] == "deliver":
            dst, kind, pair, a, bidx, val = inflight.pop(ev[1])
            cells = ranks[dst].stage[pair][a] if kind == "stage" else ranks[dst].bcast[pair]
            if cells[bidx] is not EMPTY:
                raise Violation(f"store of {val} lands on uncleared cell holding {cells[bidx]} (rank {dst})")
            cells[bidx] = [val, False]
            continue
        rk = ranks[ev[1]]
        c = ev[2]
        op = rk.progs[c][rk.pc[c]]
        rk.pc[c] += 1
        if op[0] == "push_stage":
            inflight.append((op[1], "stage", op[2], op[3], op[4], op[5]))
        elif op[0] == "push_bcast":
            inflight.append((op[1], "bcast", op[2], 0, op[3], op[4]))
        elif op[0] == "clear":
            pair, part = op[1], op[2]
            rows = rk.used[pair]
            for q in range(world):
                for j in range(part, rows, n_ctas):
                    cell = rk.stage[pair][q][j]
                    if cell is not EMPTY and not cell[1]:
                        raise Violation(f"clear wipes unconsumed {cell[0]} (rank {rk.r}, call {rk.call})")
                    rk.stage[pair][q][j] = EMPTY
            for j in range(part, rows * world, n_ctas):
                cell = rk.bcast[pair][j]
                if cell is not EMPTY and not cell[1]:
                    raise Violation(f"clear wipes unconsumed {cell[0]} (rank {rk.r}, call {rk.call})")
                rk.bcast[pair][j] = EMPTY
        elif op[0] == "poll_stage":
            cell = rk.stage[op[1]][op[2]][op[3]]
            if cell[0] != op[4]:
                raise Violation(f"stale read: wanted {op[4]}, saw {cell[0]}")
            cell[1] = True
        elif op[0] == "poll_bcast":
            cell = rk.bcast[op[1]][op[2]]
            if cell[0] != op[3]:
                raise Violation(f"stale read: wanted {op[3]}, saw {cell[0]}")
            cell[1] = True
        if all(rk.pc[k] >= len(rk.progs[k]) for k in range(n_ctas)):   # kernel boundary: the call is complete
            rk.used[rk.call % pairs] = rk.R
            rk.call += 1
            if rk.call < len(tokens_per_call):
                rk.progs, rk.R = program(rk, rk.call)
                rk.pc = [0] * n_ctas


@pytest.mark.parametrize("world", [2, 3, 4])
def test_three_pairs_clear_two_ahead_is_race_free(world):
    for seed in range(60):
        rng = random.Random(1000 * world + seed)
        calls = [rng.randint(1, 7) for _ in range(rng.randint(4, 9))]   # changing message sizes, fewer rows than ranks too
        run(world, pairs=3, tokens_per_call=calls, n_ctas=rng.randint(1, 3), seed=seed, clear_ahead=2)


def test_two_pairs_are_not_enough():
    """The same protocol with two pairs (each call clears the pair of the next call) must be caught: a rank one call ahead
    already pushes into the pair its slower peer is still clearing."""
    found = 0
    for seed in range(300):
        rng = random.Random(seed)
        calls = [rng.randint(1, 7) for _ in range(8)]
        try:
            run(2, pairs=2, tokens_per_call=calls, n_ctas=2, seed=seed, clear_ahead=1)
        except Violation:
            found += 1
    assert found >= 3   # 8 of these 300 schedules hit it (stores landing on uncleared cells, clears wiping fresh data)
