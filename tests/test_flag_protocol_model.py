"""Schedule-fuzzing model of the epoch-flag halo protocol (no GPU).

The multi-GPU path has no blocking exchange: a rank's border work items wait for the neighbours'
epoch flags, compute, push their border into the neighbours' *output* buffer and publish a flag;
interior items never wait, and launches of different ranks are not synchronised (csrc/engine.cu
`run_epochs`, csrc/life_kernels.cu `wait_neighbors` / `push_borders` / `border_item_done`).  The
reference is race-free by construction because it blocks in MPI_Waitall before every evolve
(C:328-344); here race freedom rests on two ordering rules, which this model checks under random
interleavings of the ranks' sub-steps with version stamps instead of cells:

  1. a border item of epoch e reads ghosts only after both neighbours published flag e-1;
  2. a rank publishes flag e only after all its border items of epoch e have read their ghosts
     and its pushes of epoch e have landed.

Pushes touch ghost zones only, so with these two rules even a two-buffer ping-pong is safe; the
engine nevertheless keeps the previous input out of the rotation across processes (margin for
the apron reads of interior items).  Negative controls: dropping either rule must be caught.
"""
import random

import pytest


class Rank:
    def __init__(self, nbuf):
        # per buffer: version of the interior and of the two ghost zones (written by the neighbours)
        self.interior = [None] * nbuf
        self.ghost = [[None, None] for _ in range(nbuf)]   # [from left, from right]
        self.flag = [0, 0]          # last epoch published by the left / right neighbour
        self.epoch = 1              # epoch being executed
        self.phase = 0              # 0 interior, 1 border (needs flags), 2 push, 3 publish
        self.cur = 0                # input buffer of this epoch
        self.prev_in = -1
        self.out = None


def choose_out(nbuf, cur, prev_in, exclude_prev):
    for i in range(nbuf):
        if i != cur and (i != prev_in or not exclude_prev):
            return i
    raise AssertionError("no free buffer")


def simulate(nranks, nbuf, epochs, exclude_prev, seed, wait=True, flag_before_push=False):
    """Returns None if every read saw exactly the expected version, else a description."""
    rng = random.Random(seed)
    ranks = [Rank(nbuf) for _ in range(nranks)]
    for r in ranks:                       # the load + first halo exchange (epoch 0 data everywhere)
        r.interior[0] = 0
        r.ghost[0] = [0, 0]
    while any(r.epoch <= epochs for r in ranks):
        ready = []
        for i, r in enumerate(ranks):
            if r.epoch > epochs:
                continue
            if wait and r.phase == 1 and (r.flag[0] < r.epoch - 1 or r.flag[1] < r.epoch - 1):
                continue                  # border items spin on the neighbours' flags
            ready.append(i)
        assert ready, "deadlock"
        i = rng.choice(ready)
        r, e = ranks[i], ranks[i].epoch
        left, right = ranks[(i - 1) % nranks], ranks[(i + 1) % nranks]
        if r.phase == 0:                  # interior items: read own cells only
            r.out = choose_out(nbuf, r.cur, r.prev_in, exclude_prev)
            if r.interior[r.cur] != e - 1:
                return f"rank {i} epoch {e}: interior of input buffer has version {r.interior[r.cur]}"
            r.phase = 1
        elif r.phase == 1:                # border items: read the ghost zones of the input buffer
            if r.ghost[r.cur] != [e - 1, e - 1]:
                return f"rank {i} epoch {e}: ghosts of input buffer have versions {r.ghost[r.cur]}"
            r.interior[r.out] = e
            r.phase = 2
        elif r.phase == (3 if flag_before_push else 2):   # push borders into the neighbours' output buffer
            left.ghost[r.out][1] = e      # I am the right neighbour of my left neighbour
            right.ghost[r.out][0] = e
            r.phase += 1
        else:                             # publish the epoch flag
            left.flag[1] = e
            right.flag[0] = e
            r.phase += 1
        if r.phase == 4:                  # epoch done: rotate buffers
            r.prev_in, r.cur = r.cur, r.out
            r.epoch, r.phase = e + 1, 0
    return None


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_three_buffer_rotation_is_race_free(nranks):
    for seed in range(300):
        assert simulate(nranks, nbuf=3, epochs=12, exclude_prev=True, seed=seed) is None


def test_five_buffers_with_checkpoints_are_race_free():
    # round-1 layout (GOL_NBUF = 5): two of the buffers are checkpoints and never chosen; the rotation over the
    # remaining three is what the model above covers - more free buffers can only help
    for seed in range(100):
        assert simulate(4, nbuf=5, epochs=10, exclude_prev=True, seed=seed) is None


def test_ping_pong_across_ranks_is_also_safe_in_the_model():
    for seed in range(200):
        assert simulate(3, nbuf=2, epochs=12, exclude_prev=False, seed=seed) is None


def test_missing_wait_is_caught():
    failures = [simulate(3, nbuf=3, epochs=12, exclude_prev=True, seed=s, wait=False) for s in range(200)]
    assert any(f is not None for f in failures)


def test_flag_published_before_push_is_caught():
    failures = [simulate(3, nbuf=3, epochs=12, exclude_prev=True, seed=s, flag_before_push=True) for s in range(200)]
    assert any(f is not None for f in failures)


def test_single_rank_ping_pong_is_fine():
    # one process owning the whole torus: launches are stream-ordered, two buffers suffice
    for seed in range(20):
        assert simulate(1, nbuf=2, epochs=10, exclude_prev=False, seed=seed) is None


# ------------------------------------------------------------------ deep-halo epoch pairs
# Round-2 schedule (csrc/engine.cu run_epochs, EpochMode): halos are exchanged once per two epochs.
# Odd epochs are PAIR_FIRST: border items wait for the neighbours' flag of the previous (even) epoch,
# read the full-depth ghost zone and also write the ghost rows next to the domain into their OWN
# output buffer; nothing is pushed.  Even epochs are PAIR_SECOND: no wait, they read the locally
# computed ghost rows, then push a full-depth halo into the neighbours' output buffer and publish the
# flag.  A pusher can be two epochs ahead of the rank it writes to; the engine keeps the inputs of the
# last two epochs (besides the current one) out of the choice of the output buffer: four rotating buffers.

class PairRank:
    def __init__(self, nbuf):
        self.interior = [None] * nbuf
        self.ghost = [[None, None] for _ in range(nbuf)]   # per side: (version, "push" | "local")
        self.flag = [0, 0]
        self.epoch, self.phase, self.cur, self.hist, self.out = 1, 0, 0, [-1, -1], None


def pick_out(nbuf, cur, hist, depth):
    banned = {cur} | set(hist[:depth])
    for i in range(nbuf):
        if i not in banned:
            return i
    raise AssertionError("no free buffer")


def simulate_pairs(nranks, nbuf, epochs, seed, hist_depth=2, wait=True):
    rng = random.Random(seed)
    ranks = [PairRank(nbuf) for _ in range(nranks)]
    for r in ranks:                       # load + first (full-depth) halo exchange
        r.interior[0] = 0
        r.ghost[0] = [(0, "push"), (0, "push")]
    while any(r.epoch <= epochs for r in ranks):
        ready = []
        for i, r in enumerate(ranks):
            if r.epoch > epochs:
                continue
            first = r.epoch % 2 == 1
            if wait and first and r.phase == 1 and (r.flag[0] < r.epoch - 1 or r.flag[1] < r.epoch - 1):
                continue
            ready.append(i)
        assert ready, "deadlock"
        i = rng.choice(ready)
        r, e = ranks[i], ranks[i].epoch
        first = e % 2 == 1
        left, right = ranks[(i - 1) % nranks], ranks[(i + 1) % nranks]
        if r.phase == 0:                  # interior items (never wait)
            r.out = pick_out(nbuf, r.cur, r.hist, hist_depth)
            if r.interior[r.cur] != e - 1:
                return f"rank {i} epoch {e}: interior version {r.interior[r.cur]}"
            r.phase = 1
        elif r.phase == 1:                # border items
            want = (e - 1, "push") if first else (e - 1, "local")
            if r.ghost[r.cur] != [want, want]:
                return f"rank {i} epoch {e}: ghosts {r.ghost[r.cur]}, expected {want}"
            r.interior[r.out] = e
            if first:
                r.ghost[r.out] = [(e, "local"), (e, "local")]   # ghost rows computed locally
                r.phase = 4                                     # no push, no flag
            else:
                r.phase = 2
        elif r.phase == 2:                # push a full-depth halo into the neighbours' output buffer
            left.ghost[r.out][1] = (e, "push")
            right.ghost[r.out][0] = (e, "push")
            r.phase = 3
        else:                             # publish
            left.flag[1] = e
            right.flag[0] = e
            r.phase = 4
        if r.phase == 4:
            r.hist = [r.cur, r.hist[0]]
            r.cur = r.out
            r.epoch, r.phase = e + 1, 0
    return None


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_deep_halo_pairs_with_four_rotating_buffers_are_race_free(nranks):
    for seed in range(300):
        assert simulate_pairs(nranks, nbuf=4, epochs=16, seed=seed) is None
    for seed in range(100):               # GOL_NBUF = 6 with its two checkpoints out of the rotation: 4 rotate; more only helps
        assert simulate_pairs(nranks, nbuf=6, epochs=12, seed=seed) is None


def test_deep_halo_pairs_three_rotating_buffers_already_suffice_in_the_model():
    # keeping only the previous input out of the rotation (the round-1 rule) is race-free here too: a pusher
    # reaches its pushing epoch only after the waiting epoch before it saw the neighbour's flag, which the
    # neighbour publishes after its border items have read their ghosts.  The engine keeps a fourth buffer
    # in the rotation anyway (margin for the apron reads of interior items while PDL overlaps launches).
    for seed in range(300):
        assert simulate_pairs(3, nbuf=3, epochs=16, seed=seed, hist_depth=1) is None


def test_deep_halo_pairs_missing_wait_is_caught():
    failures = [simulate_pairs(3, nbuf=4, epochs=16, seed=s, wait=False) for s in range(300)]
    assert any(f is not None for f in failures)
