"""Host-side model of the cross-GPU flag / epoch protocol of the fused collective kernels.

The device code (csrc/optim.cu ``arrive_and_wait_all`` + ``wait_flag_sys`` in csrc/common.cuh) is a few
lines, but its correctness argument is about interleavings no GPU test can enumerate.  This model restates
it in Python — rank r publishes ``epoch`` into word ``[slot][r]`` of every peer's flag block, then waits
until every peer's word in ITS block satisfies ``int32(flag - epoch) >= 0`` — and explores every
interleaving of a few ranks over several calls with a breadth-first search over the joint state.  Checked:

* safety: no rank leaves phase A of call k before every rank has entered call k (so nobody reads a
  gradient buffer that is not complete), and no rank leaves phase C before every rank finished phase B;
* liveness: from every reachable state some rank can move until all are done (no deadlock), including
  when one rank arrives a whole call late;
* wrap-around: the signed-difference comparison keeps both properties when the 32-bit epoch wraps;
* the timeout path: a rank that gives up leaves its counters consistent (the next call's epoch still
  compares greater) and never writes state.
"""
import itertools
from collections import deque


def i32(x):
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & 0x80000000 else x


def reached(flag, epoch):
    """wait_flag_sys's exit test: wrap-safe 'flag >= epoch'."""
    return i32(flag - epoch) >= 0


class Model:
    """Program of each rank per call: A_pub (publish epoch in slot 0) -> A_wait -> B (work) ->
    C_pub (slot 1) -> C_wait -> next call.  State: per rank (call index, pc) + flag words."""

    PCS = ("A_pub", "A_wait", "B", "C_pub", "C_wait")

    def __init__(self, world, calls, epoch0):
        self.world, self.calls, self.epoch0 = world, calls, epoch0 & 0xFFFFFFFF

    def epoch(self, call):
        return (self.epoch0 + call) & 0xFFFFFFFF

    def initial(self):
        # flags[dst][slot][src]: the epoch before the first call is "already seen"
        start = (self.epoch0 - 1) & 0xFFFFFFFF
        flags = tuple(tuple(tuple(start for _ in range(self.world)) for _ in range(2)) for _ in range(self.world))
        return tuple((0, 0) for _ in range(self.world)), flags

    def done(self, state):
        return all(call == self.calls for call, _ in state[0])

    def step(self, state, r):
        """Successor when rank r takes its next step, or None if r is blocked / finished."""
        ranks, flags = state
        call, pc = ranks[r]
        if call == self.calls:
            return None
        name = self.PCS[pc]
        e = self.epoch(call)
        if name in ("A_pub", "C_pub"):
            slot = 0 if name == "A_pub" else 1
            fl = [[list(s) for s in dst] for dst in flags]
            for dst in range(self.world):
                if dst != r:
                    fl[dst][slot][r] = e
            flags = tuple(tuple(tuple(s) for s in dst) for dst in fl)
        elif name in ("A_wait", "C_wait"):
            slot = 0 if name == "A_wait" else 1
            if not all(reached(flags[r][slot][p], e) for p in range(self.world) if p != r):
                return None
        nxt = (call, pc + 1) if pc + 1 < len(self.PCS) else (call + 1, 0)
        ranks = tuple(nxt if i == r else v for i, v in enumerate(ranks))
        return ranks, flags

    def check_safety(self, state):
        ranks, _ = state
        for r, (call, pc) in enumerate(ranks):
            if call == self.calls:
                continue
            name = self.PCS[pc]
            for p, (pcall, ppc) in enumerate(ranks):
                if p == r:
                    continue
                if name in ("B", "C_pub", "C_wait"):
                    # r is past phase A of `call`: every peer must have ENTERED call `call` (published A)
                    assert (pcall, ppc) >= (call, 1), f"rank {r} passed A of call {call} before rank {p} arrived: {ranks}"
                if (call, pc) > (0, 0) and pc == 0:
                    # r finished call-1 entirely: every peer must have finished phase B of call-1
                    assert (pcall, ppc) >= (call - 1, 4), f"rank {r} left C of call {call - 1} early: {ranks}"

    def explore(self):
        seen = set()
        q = deque([self.initial()])
        seen.add(q[0])
        finals = 0
        while q:
            st = q.popleft()
            self.check_safety(st)
            succ = [s for s in (self.step(st, r) for r in range(self.world)) if s is not None]
            if not succ:
                assert self.done(st), f"deadlock at {st[0]}"
                finals += 1
            for s in succ:
                if s not in seen:
                    seen.add(s)
                    q.append(s)
        assert finals >= 1
        return len(seen)


def test_every_interleaving_of_two_and_three_ranks_is_safe_and_live():
    assert Model(2, 3, epoch0=1).explore() > 50
    assert Model(3, 2, epoch0=1).explore() > 100


def test_epoch_wrap_around_keeps_safety_and_liveness():
    # calls straddle 0xFFFFFFFF -> 0: the signed-difference test must still order them
    Model(2, 4, epoch0=0xFFFFFFFE).explore()
    Model(3, 2, epoch0=0xFFFFFFFF).explore()
    # the comparison itself
    assert reached(0x00000001, 0xFFFFFFFF) and not reached(0xFFFFFFFF, 0x00000001)
    assert reached(5, 5) and not reached(4, 5)


def test_a_rank_a_whole_call_late_cannot_be_overtaken():
    """Rank 1 does nothing until rank 0 has gone as far as it can: rank 0 must be stuck in call 0's phase A
    (it may never run ahead into call 1), and once rank 1 moves both finish."""
    m = Model(2, 2, epoch0=7)
    st = m.initial()
    while True:
        nxt = m.step(st, 0)
        if nxt is None:
            break
        st = nxt
    assert st[0][0] == (0, 1), st[0]          # blocked at A_wait of call 0
    # now any fair schedule completes
    for order in itertools.product((0, 1), repeat=12):
        s = st
        for r in order:
            n = m.step(s, r)
            if n is not None:
                s = n
                m.check_safety(s)
    while not m.done(st):
        moved = False
        for r in (1, 0):
            n = m.step(st, r)
            if n is not None:
                st, moved = n, True
                m.check_safety(st)
        assert moved, "deadlock after the late rank arrived"


def test_timeout_leaves_the_next_call_consistent():
    """A rank that times out in phase A returns without touching state and without publishing phase C.
    Its peers' words for that call stay below the next epoch, so the next call (epoch + 1) still waits for
    a real arrival: stale flags can never satisfy a later call."""
    m = Model(2, 2, epoch0=100)
    ranks, flags = m.initial()
    st = m.step((ranks, flags), 0)            # rank 0 publishes call 0 ... rank 1 never shows up
    assert m.step(st, 0) is None              # rank 0 would wait; on the device it gives up here (status word set)
    # rank 0 skips B / C of call 0 and enters call 1
    st = (((1, 0), st[0][1]), st[1])
    st = m.step(st, 0)                        # publishes epoch 101
    assert m.step(st, 0) is None, "a stale flag satisfied the next call"
    # rank 1 finally runs call 0's phase A: it sees rank 0's word = 101 >= 100 and proceeds - it observes a
    # peer that is AHEAD, never one that is behind, which is the property the gradient buffers need
    s1 = m.step(st, 1)
    assert s1 is not None and m.step(s1, 1) is not None
