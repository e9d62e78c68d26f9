"""Model check of the hand-overs inside the single-launch SimpleNet step (csrc/mlp_fused.cu).

One launch holds, per rank, a compute cluster and H update CTAs that run `steps` training steps
without a kernel boundary between them.  What orders them (all monotonic counters / epoch flags):

    ready[r]        += 1 by the compute cluster when step t's gradients are in memory
    entry flags     update CTA 0 of rank r publishes "gradients of step t ready" to every rank;
                    every update CTA waits for all ranks' flags              (comm_device.cuh begin_barrier)
    ticket / exit   every update CTA takes a ticket after its reduce + update + push; the last one
                    publishes "rank r has pushed step t" to every rank and waits for the same from
                    all of them                                              (end_barrier, two-shot / NVLS)
    done[r]         += 1 by every update CTA (the last one after the exit wait); the compute cluster
                    starts step t + 1 once done[r] == H * t

The model steps every agent (one compute agent and H update agents per rank) one atomic action at a
time in random interleavings and checks: an update CTA only ever reads gradients of ITS step, also
at the end of a read that started earlier (the compute cluster may be a step ahead: the gradient
arena is double-buffered on step parity); the forward pass of step t + 1 sees, on every rank, the
parameters of step t from EVERY shard owner.  Negative controls drop one wait each and must be caught.

compute-sanitizer cannot see cross-GPU ordering (and is closed on this pool); the GPU-side proof is
tests/multi_gpu_worker.py (replicas bit-identical after every step at 2 and 8 GPUs).
"""
import random

import pytest


class World:
    def __init__(self, world, helpers, buffers):
        W, H = world, helpers
        self.W, self.H, self.buffers = W, H, buffers
        self.ready = [0] * W
        self.done = [0] * W
        self.tickets = [0] * W
        self.grad = [[0] * buffers for _ in range(W)]              # step held by each gradient buffer
        self.ready_f = [[0] * W for _ in range(W)]                 # [at rank][from rank]
        self.done_f = [[0] * W for _ in range(W)]
        self.param = [[[0] * H for _ in range(W)] for _ in range(W)]   # [at rank][shard owner][portion]: step


def run(world, helpers, steps, rng, buffers=2, wait_done=True, exit_wait=True, wait_entry=True):
    w = World(world, helpers, buffers)
    W, H = world, helpers
    # agents: ("c", r) and ("u", r, h); each a small program counter machine
    agents = {}
    for r in range(W):
        agents[("c", r)] = {"t": 1, "pc": "wait"}
        for h in range(H):
            agents[("u", r, h)] = {"t": 1, "pc": "ready", "peer": 0, "seen": None, "last": False}
    budget = 400 * W * (H + 1) * steps
    live = [k for k in agents]
    while live:
        budget -= 1
        if budget < 0:
            return "deadlock / livelock"
        key = rng.choice(live)
        a = agents[key]
        t = a["t"]
        if key[0] == "c":
            r = key[1]
            if a["pc"] == "wait":
                if t == 1 or not wait_done or w.done[r] >= H * (t - 1):
                    a["pc"] = "forward"
            elif a["pc"] == "forward":
                for owner in range(W):
                    for h in range(H):
                        if w.param[r][owner][h] != t - 1:
                            return (f"rank {r} runs the forward pass of step {t} on parameters of step "
                                    f"{w.param[r][owner][h]} from shard owner {owner}")
                a["pc"] = "grad"
            elif a["pc"] == "grad":
                w.grad[r][t % buffers] = t
                a["pc"] = "signal"
            elif a["pc"] == "signal":
                w.ready[r] += 1
                a["t"], a["pc"] = t + 1, "wait"
                if a["t"] > steps:
                    live.remove(key)
            continue
        _, r, h = key
        if a["pc"] == "ready":
            if w.ready[r] >= t:
                a["pc"] = "publish" if h == 0 else "entry"
        elif a["pc"] == "publish":
            for j in range(W):
                w.ready_f[j][r] = t
            a["pc"] = "entry"
        elif a["pc"] == "entry":
            if not wait_entry or all(w.ready_f[r][j] >= t for j in range(W)):
                a["pc"], a["peer"] = "read_begin", 0
        elif a["pc"] == "read_begin":            # a read of one peer's gradients is not atomic
            a["seen"] = w.grad[a["peer"]][t % buffers]
            a["pc"] = "read_end"
        elif a["pc"] == "read_end":
            now = w.grad[a["peer"]][t % buffers]
            if a["seen"] != t or now != t:
                return (f"update CTA {h} of rank {r}, step {t}: read gradients of step {a['seen']}/{now} "
                        f"from rank {a['peer']}")
            a["peer"] += 1
            a["pc"] = "read_begin" if a["peer"] < W else "push"
        elif a["pc"] == "push":                  # update of this CTA's portion of shard r, pushed to every rank
            for j in range(W):
                w.param[j][r][h] = t
            a["pc"] = "ticket"
        elif a["pc"] == "ticket":
            w.tickets[r] += 1
            a["last"] = w.tickets[r] == H
            a["pc"] = "exit_publish" if a["last"] else "count_in"
        elif a["pc"] == "exit_publish":
            w.tickets[r] = 0
            for j in range(W):
                w.done_f[j][r] = t
            a["pc"] = "exit_wait"
        elif a["pc"] == "exit_wait":
            if not exit_wait or all(w.done_f[r][j] >= t for j in range(W)):
                a["pc"] = "count_in"
        elif a["pc"] == "count_in":
            w.done[r] += 1
            a["t"], a["pc"] = t + 1, "ready"
            if a["t"] > steps:
                live.remove(key)
    return None


@pytest.mark.parametrize("world,helpers", [(1, 3), (2, 2), (4, 3), (8, 2)])
def test_single_launch_step_handshakes_are_safe_and_live(world, helpers):
    for seed in range(60):
        assert run(world, helpers, 5, random.Random(seed)) is None, seed


def test_forward_without_the_done_counter_is_caught():
    hits = sum(run(2, 2, 4, random.Random(s), wait_done=False) is not None for s in range(40))
    assert hits >= 35


def test_skipping_the_exit_wait_is_caught():
    """Without the exit wait of the last update CTA a rank can start the next forward pass before a
    slower peer has pushed its shard of the new parameters."""
    hits = sum(run(4, 2, 4, random.Random(s), exit_wait=False) is not None for s in range(60))
    assert hits >= 20


def test_skipping_the_entry_barrier_is_caught():
    hits = sum(run(4, 2, 4, random.Random(s), wait_entry=False) is not None for s in range(40))
    assert hits >= 35


def test_ticket_chain_never_deadlocks_with_one_update_cta():
    for seed in range(30):
        assert run(3, 1, 6, random.Random(seed)) is None


# --------------------------------------------------------------------------- end-to-end ring
def run_ring(steps, slots, lookahead, loss_ring, rng):
    """b200_mlp_host_run: the host feeds pinned batches into `slots` device slots (copy, then the
    step's sequence number), at most `lookahead` steps ahead of the last loss it has read; the kernel
    waits for slot[t % slots] >= t + 1, reads the batch (a read is not atomic), publishes loss t + 1
    into ring[t % loss_ring]; the host reads losses in order.  Checks: the kernel only ever reads
    the batch of its own step, start to end; the host reads every loss before it is overwritten."""
    slot_data = [None] * slots          # step whose batch sits in the slot (None: garbage)
    slot_flag = [0] * slots             # sequence number written behind the batch
    ring = [(0, None)] * loss_ring      # (sequence number, loss of which step)
    fed = got = 0
    k = {"t": 0, "pc": "wait", "seen": None}
    host_pc = "feed"
    budget = 200 * steps
    while got < steps:
        budget -= 1
        if budget < 0:
            return "deadlock / livelock"
        if rng.random() < 0.5:                                   # host
            if host_pc == "feed" and fed < steps and fed - got < lookahead:
                slot_data[fed % slots] = fed                     # the copy
                host_pc = "flag"
            elif host_pc == "flag":
                slot_flag[fed % slots] = fed + 1                 # stream-ordered behind the copy
                fed += 1
                host_pc = "feed"
            elif ring[got % loss_ring][0] >= got + 1:            # loss of step `got` has arrived
                if ring[got % loss_ring] != (got + 1, got):
                    return f"host read loss slot {ring[got % loss_ring]} for step {got}"
                got += 1
        else:                                                    # kernel
            t = k["t"]
            if t >= steps:
                continue
            if k["pc"] == "wait":
                if slot_flag[t % slots] >= t + 1:
                    k["pc"] = "read_begin"
            elif k["pc"] == "read_begin":
                k["seen"] = slot_data[t % slots]
                k["pc"] = "read_end"
            elif k["pc"] == "read_end":
                if k["seen"] != t or slot_data[t % slots] != t:
                    return f"step {t} read the batch of step {k['seen']}/{slot_data[t % slots]}"
                k["pc"] = "publish"
            elif k["pc"] == "publish":
                ring[t % loss_ring] = (t + 1, t)
                k["t"], k["pc"] = t + 1, "wait"
    return None


@pytest.mark.parametrize("slots", [2, 3, 4])
def test_input_ring_is_safe_with_slots_minus_one_steps_in_flight(slots):
    for seed in range(40):
        assert run_ring(60, slots, slots - 1, 8, random.Random(seed)) is None, seed


def test_feeding_as_many_steps_ahead_as_there_are_slots_is_safe_too_but_one_more_is_caught():
    # `slots` in flight is still safe in this model (a slot is rewritten only after the loss of the
    # step that read it); the native loop keeps one slot of margin.  slots + 1 overwrites a batch
    # that has not been read.
    assert all(run_ring(60, 3, 3, 8, random.Random(s)) is None for s in range(40))
    hits = sum(run_ring(60, 3, 4, 8, random.Random(s)) is not None for s in range(40))
    assert hits >= 30


def test_a_loss_ring_shorter_than_the_lookahead_is_caught():
    hits = sum(run_ring(60, 4, 3, 2, random.Random(s)) is not None for s in range(40))
    assert hits >= 20
