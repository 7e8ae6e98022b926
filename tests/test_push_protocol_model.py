"""Executable model of the push form's token protocol (csrc/gather.cu push_tma_cvt_kernel +
DeviceEngine.push_step / release_slot), checked under random schedules.

The reference hands a window over with zero-byte messages (reference ddl/connection.py:153-182).  The push
form hands a *ring slot* over with three monotonic counters per slot: `free` (consumer -> every owner: "I consumed
fill number n"), `ready` (every owner -> consumer: += 1 per fill) and the kernel's own completion.  compute-sanitizer
is closed on the pool, so the protocol is stated here as a small discrete-event model -- R ranks, each
with fetch streams, a training stream and per-consumer CTA groups that run in any order -- and hypothesis picks the
interleaving.  Safety: a consumer only ever reads a slot whose R parts all belong to the batch it was promised, and
no owner writes a slot its consumer has not released.  Liveness: some action is always enabled until every batch has
been consumed (no deadlock for any ring depth > prefetch depth).
"""
from hypothesis import HealthCheck, Phase, given, settings
from hypothesis import strategies as st


class Rank:
    def __init__(self, r, R, depth, streams):
        self.r, self.R = r, R
        self.ring = [[None] * R for _ in range(depth)]     # ring[slot][owner] = batch whose part lies there
        self.ready = [0] * depth                           # += 1 by every owner per fill
        self.free = [[0] * R for _ in range(depth)]        # free[slot][consumer] = fills consumed by that consumer
        self.use = [0] * depth                             # host side: fills launched into the slot
        self.queues = [[] for _ in range(streams)]         # fetch streams: kernels in launch order
        self.consumed_event = [0] * depth                  # fills of the slot marked on the training stream
        self.done = {}                                     # batch -> its kernel has completed
        self.next_consume = 0
        self.held = None


class Kernel:
    def __init__(self, batch, slot, use, R):
        self.batch, self.slot, self.use = batch, slot, use
        self.groups_left = set(range(R))                   # consumers this kernel still has to serve
        self.started = False


def simulate(R, depth, prefetch, streams, steps, choose, honour_tokens=True):
    ranks = [Rank(r, R, depth, streams) for r in range(R)]
    launched = [0] * R

    def launch(rk):
        b = launched[rk.r]
        slot = b % depth
        rk.queues[slot % streams].append(Kernel(b, slot, rk.use[slot], R))
        rk.use[slot] += 1
        launched[rk.r] += 1

    for rk in ranks:                                       # construction: the first window's prefetches
        for _ in range(min(prefetch, steps)):
            launch(rk)

    def enabled():
        acts = []
        for rk in ranks:
            for q in rk.queues:
                if not q:
                    continue
                k = q[0]
                if not k.started:
                    # stream order: behind its predecessor (popped) and the slot's consumed event
                    if rk.consumed_event[k.slot] >= k.use or not honour_tokens:
                        acts.append(("start", rk.r, k))
                    continue
                for c in sorted(k.groups_left):            # a CTA group passes its in-kernel wait on `free`
                    if rk.free[k.slot][c] >= k.use or not honour_tokens:
                        acts.append(("push", rk.r, k, c))
                if not k.groups_left and rk.ready[k.slot] >= (k.use + 1) * R:
                    acts.append(("finish", rk.r, k))
            if rk.held is None and rk.next_consume < steps and rk.done.get(rk.next_consume):
                acts.append(("take", rk.r))
            if rk.held is not None:
                acts.append(("mark", rk.r))
        return acts

    consumed = 0
    while consumed < R * steps:
        acts = enabled()
        assert acts, "deadlock: no rank can make progress"
        a = acts[choose(len(acts))]
        rk = ranks[a[1]]
        if a[0] == "start":
            a[2].started = True
        elif a[0] == "push":
            k, c = a[2], a[3]
            dst = ranks[c]
            # safety: the consumer has released what lay in this slot (its previous fill)
            assert dst.consumed_event[k.slot] >= k.use, "an owner wrote a slot its consumer still holds"
            dst.ring[k.slot][rk.r] = k.batch
            dst.ready[k.slot] += 1                          # token written after the data (fence + red.release)
            k.groups_left.discard(c)
        elif a[0] == "finish":
            k = a[2]
            next(q for q in rk.queues if q and q[0] is k).pop(0)
            rk.done[k.batch] = True
        elif a[0] == "take":
            b = rk.next_consume
            slot = b % depth
            assert rk.ring[slot] == [b] * R, f"rank {rk.r} was handed batch {b} but the slot holds {rk.ring[slot]}"
            rk.held = b
            while launched[rk.r] < min(steps, b + prefetch + 1):   # __getitem__ prefetches ahead
                launch(rk)
        else:  # mark(END_OF_BATCH): consumed event on the training stream, then the free tokens to every owner
            b, rk.held = rk.held, None
            slot = b % depth
            rk.consumed_event[slot] += 1
            for o in ranks:
                o.free[slot][rk.r] = rk.consumed_event[slot]
            rk.next_consume += 1
            consumed += 1
    return True


@settings(max_examples=150, deadline=None, suppress_health_check=list(HealthCheck),
          phases=(Phase.explicit, Phase.reuse, Phase.generate))
@given(R=st.integers(1, 4), prefetch=st.integers(1, 3), extra=st.integers(1, 2), streams=st.integers(1, 3),
       steps=st.integers(1, 14), rnd=st.randoms(use_true_random=False))
def test_push_tokens_are_safe_and_live_under_any_schedule(R, prefetch, extra, streams, steps, rnd):
    depth = prefetch + extra                               # LoaderConfig: ring_depth > prefetch_depth
    assert simulate(R, depth, prefetch, streams, steps, lambda n: rnd.randrange(n))


def test_model_catches_an_owner_that_ignores_the_tokens():
    """Mutation check of the model itself: without the `free` / `consumed` waits some schedule lets an owner refill
    a slot that its consumer still holds (or hands out a torn slot), and the model says so."""
    import random

    import pytest

    rnd = random.Random(11)
    with pytest.raises(AssertionError, match="still holds|was handed batch"):
        for _ in range(300):
            simulate(3, 2, 1, 2, 10, lambda n: rnd.randrange(n), honour_tokens=False)
