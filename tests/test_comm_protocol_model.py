"""Model check of the flag/epoch protocol of the in-kernel collectives (csrc/comm.cu).

compute-sanitizer's racecheck cannot see cross-GPU ordering (and is closed on this pool), so the
protocol is checked as a model on the host: W ranks, each running the same sequence of collective
kernels, are stepped in random interleavings one atomic action at a time.

Per kernel and rank, as in comm.cu:
    produce   the backward pass writes this step's gradients      (stream-ordered before the kernel)
    begin     publish ready[me] = epoch into every rank's flag block, wait for ready[*] >= epoch
    read      load every peer's gradients
    end       publish done[me] = epoch everywhere, wait for done[*] >= epoch, epoch += 1
Epochs only grow (no reset), comparisons are `flag >= epoch`.

Safety: whenever a rank reads a peer's gradients they are the ones of the current epoch — not yet
produced is a missing begin barrier, already overwritten a missing end barrier.  Liveness: every
schedule finishes.  The negative controls drop one barrier each and must be caught.

Double-buffered variant (B200_COMM_NO_EXIT_WAIT): the producer of epoch e writes buffer e % 2 and
the kernel has NO end barrier.  A peer can only overwrite buffer e % 2 when producing epoch e + 2,
after its epoch e + 1 kernel passed the begin barrier, i.e. after every rank announced e + 1 —
which a rank does only once its epoch e reads are complete.  Reads are modelled one peer at a time
(they are not atomic on the device either).  Negative control: the same protocol on ONE buffer.
"""
import random

import pytest

PRODUCE, PUB_READY, WAIT_READY, READ, PUB_DONE, WAIT_DONE = range(6)


class Rank:
    def __init__(self, r, world):
        self.r, self.world = r, world
        self.ready = [0] * world      # this rank's flag block: ready[j] written by rank j
        self.done = [0] * world
        self.epoch = 0                # local, advanced at the end of each kernel
        self.data = 0                 # epoch whose gradients currently sit in the arena
        self.bufs = [0, 0]            # double-buffered arena: epoch held by buffer 0 / 1
        self.read_next = 0            # next peer to read (reads are one action per peer)
        self.pc = PRODUCE
        self.kernels_done = 0


def run_schedule(world, kernels, rng, use_begin=True, use_end=True, buffers=1):
    """Random interleaving; returns None when all invariants held, else a description.
    buffers=2 is the double-buffered arena (combine with use_end=False for the no-exit-wait kernel)."""
    ranks = [Rank(r, world) for r in range(world)]
    budget = 200 * world * kernels
    while any(k.kernels_done < kernels for k in ranks):
        budget -= 1
        if budget < 0:
            return "deadlock / livelock"
        me = rng.choice([k for k in ranks if k.kernels_done < kernels])
        e = me.epoch + 1
        if me.pc == PRODUCE:
            me.data = e                                   # overwrites the previous step's gradients
            me.bufs[e % buffers] = e                      # (the other buffer keeps epoch e - 1)
            me.pc = PUB_READY
        elif me.pc == PUB_READY:
            for peer in ranks:
                peer.ready[me.r] = e                      # st.release.sys into every flag block
            me.pc = WAIT_READY
        elif me.pc == WAIT_READY:
            if not use_begin or all(v >= e for v in me.ready):
                me.pc = READ
        elif me.pc == READ:
            peer = ranks[me.read_next]                    # one peer per action: reads are not atomic
            have = peer.bufs[e % buffers]
            if have != e:
                kind = "not yet produced" if have < e else "already overwritten"
                return f"rank {me.r} epoch {e} read rank {peer.r} data of epoch {have}: {kind}"
            me.read_next += 1
            if me.read_next == world:
                me.read_next = 0
                me.pc = PUB_DONE
        elif me.pc == PUB_DONE:
            for peer in ranks:
                peer.done[me.r] = e
            me.pc = WAIT_DONE
        elif me.pc == WAIT_DONE:
            if not use_end or all(v >= e for v in me.done):
                me.epoch = e
                me.kernels_done += 1
                me.pc = PRODUCE
    return None


@pytest.mark.parametrize("world", [2, 4, 8])
def test_protocol_is_safe_and_live_under_random_interleavings(world):
    rng = random.Random(1234 + world)
    for _ in range(300):
        assert run_schedule(world, kernels=6, rng=rng) is None


@pytest.mark.parametrize("world", [2, 4, 8])
def test_double_buffered_protocol_needs_no_end_barrier(world):
    """comm.cu no_exit_wait: gradients alternate between two buffers on step parity and the
    one-shot kernel leaves without waiting for its peers."""
    rng = random.Random(4321 + world)
    for _ in range(300):
        assert run_schedule(world, kernels=8, rng=rng, use_end=False, buffers=2) is None


def test_no_end_barrier_on_a_single_buffer_is_caught():
    """The same kernel on ONE buffer lets a fast rank overwrite what a slow peer still reads —
    which is why b200_allreduce_optimizer rejects no_exit_wait without a parity stride."""
    rng = random.Random(11)
    found = [run_schedule(4, 8, rng, use_end=False, buffers=1) for _ in range(300)]
    assert any(f and "already overwritten" in f for f in found)


def test_double_buffering_does_not_replace_the_begin_barrier():
    rng = random.Random(12)
    found = [run_schedule(4, 8, rng, use_begin=False, use_end=False, buffers=2) for _ in range(300)]
    assert any(f and "not yet produced" in f for f in found)


def test_dropping_the_end_barrier_is_caught():
    """Without "reads done" a fast rank starts the next backward and overwrites gradients a slow
    peer is still reading."""
    rng = random.Random(7)
    found = [run_schedule(4, 6, rng, use_end=False) for _ in range(300)]
    assert any(f and "already overwritten" in f for f in found)


def test_dropping_the_begin_barrier_is_caught():
    """Without "gradients ready" a fast rank reads a peer that has not produced this step yet."""
    rng = random.Random(8)
    found = [run_schedule(4, 6, rng, use_begin=False) for _ in range(300)]
    assert any(f and "not yet produced" in f for f in found)


def test_epoch_comparison_survives_wraparound():
    """comm.cu compares `(int32)(flag - epoch) >= 0`, so the 32-bit epoch may wrap."""
    def ge(flag, epoch):
        d = (flag - epoch) & 0xFFFFFFFF
        return d < 0x80000000
    assert ge(5, 5) and ge(6, 5) and not ge(4, 5)
    assert ge(2, 0xFFFFFFFE) and not ge(0xFFFFFFFE, 2)      # across the wrap
