"""mbarrier protocol of the persistent conv kernel (conv_igemm_persist_kernel, csrc/conv_igemm.cu), executed
as a randomly scheduled model on the host.

The kernel cannot run here, but its synchronisation is a small protocol between three actors -- the TMA
producer, the MMA issuer and the epilogue -- over `full[STAGES]`, `empty[STAGES]`, `tmem_full[2]` and
`tmem_empty[2]` barriers with parity waits.  The model below follows the kernel line by line (same loop
structure, same stage / phase / accumulator-use arithmetic) and checks, under thousands of random
interleavings, that it never deadlocks and never lets an actor touch a buffer another actor still owns:
a producer overwriting a stage the tensor core has not consumed, an MMA writing into an accumulator the
epilogue is still draining, an epilogue reading a tile that is not finished.  Asynchrony of TMA and of the
tensor core is part of the model: loads and MMAs complete at arbitrary later points.
"""
import random

import pytest


class MBar:
    """mbarrier with `count` expected arrivals; try_wait.parity(P) succeeds iff the phase of parity P has
    completed, i.e. the barrier currently sits in a phase of the other parity."""

    def __init__(self, count):
        self.count, self.pending, self.phase = count, count, 0

    def arrive(self):
        self.pending -= 1
        assert self.pending >= 0, "more arrivals than the barrier expects in one phase"
        if self.pending == 0:
            self.phase += 1
            self.pending = self.count

    def passed(self, parity):
        return (self.phase & 1) != parity


def simulate(total_tiles, grid, block, total_kb, stages, seed, epilogue_warps=4, bug=None):
    rng = random.Random(seed)
    full = [MBar(1) for _ in range(stages)]
    empty = [MBar(1) for _ in range(stages)]
    tmem_full = [MBar(1) for _ in range(2)]
    tmem_empty = [MBar(1 if bug == "early_release" else epilogue_warps) for _ in range(2)]
    my_tiles = list(range(block, total_tiles, grid))

    stage_state = ["free"] * stages        # free -> loading -> loaded -> (consumed by MMA) -> free
    stage_tag = [None] * stages            # (tile, kb) the stage holds
    acc_state = ["free", "free"]           # free -> accumulating -> complete -> draining -> free
    acc_tile = [None, None]
    acc_kb_done = [0, 0]
    tma_q, mma_q = [], []                  # asynchronous completions, in issue order per engine
    done_tiles = []

    def producer():
        stage, phase = 0, 0
        for t in my_tiles:
            for kb in range(total_kb):
                while not empty[stage].passed(phase ^ 1):
                    yield
                assert stage_state[stage] == "free", "TMA overwrites a stage the tensor core has not released"
                stage_state[stage] = "loading"
                stage_tag[stage] = (t, kb)

                def land(s=stage):
                    assert stage_state[s] == "loading"
                    stage_state[s] = "loaded"
                    full[s].arrive()       # complete_tx of the expected bytes
                tma_q.append(land)
                stage += 1
                if stage == stages:
                    stage, phase = 0, phase ^ 1
                yield

    def mma():
        stage, phase = 0, 0
        for it, t in enumerate(my_tiles):
            buf, use = it & 1, (it >> 1) & 1
            while not tmem_empty[buf].passed(use if bug == "empty_parity" else use ^ 1):
                yield
            assert acc_state[buf] == "free", "MMA starts on an accumulator the epilogue has not drained"
            acc_state[buf], acc_tile[buf], acc_kb_done[buf] = "accumulating", t, 0
            for kb in range(total_kb):
                while not full[stage].passed(phase):
                    yield
                assert stage_state[stage] == "loaded" and stage_tag[stage] == (t, kb), "MMA reads the wrong operands"

                # tcgen05.mma is asynchronous: the commit arrives on empty[stage] once the MMAs issued so far
                # have completed, in issue order
                def complete(s=stage, b=buf, k=kb, tile=t):
                    assert acc_tile[b] == tile and acc_kb_done[b] == k, "MMAs complete out of order"
                    acc_kb_done[b] += 1
                    stage_state[s] = "free"
                    empty[s].arrive()
                mma_q.append(complete)
                stage += 1
                if stage == stages:
                    stage, phase = 0, phase ^ 1
                yield

            def tile_done(b=buf, tile=t):
                assert acc_kb_done[b] == total_kb
                acc_state[b] = "complete"
                tmem_full[b].arrive()
            mma_q.append(tile_done)
            yield

    def epilogue_warp(w):
        for it, t in enumerate(my_tiles):
            buf, use = it & 1, (it >> 1) & 1
            while not tmem_full[buf].passed(use):
                yield
            assert acc_state[buf] in ("complete", "draining") and acc_tile[buf] == t, "epilogue reads an unfinished tile"
            acc_state[buf] = "draining"
            yield                          # tcgen05.ld of all column chunks, wait::ld
            if bug != "early_release" or w != 0:
                drained[buf] += 1
                if drained[buf] == epilogue_warps:     # the arrival below completes the barrier phase
                    acc_state[buf], drained[buf] = "free", 0
            tmem_empty[buf].arrive()       # lane 0 of the warp, after tcgen05.fence::before_thread_sync
            if w == 0:
                done_tiles.append(t)
            yield                          # stores / statistics of this tile

    drained = [0, 0]
    actors = [producer(), mma()] + [epilogue_warp(w) for w in range(epilogue_warps)]
    alive = list(actors)
    idle_rounds = 0
    while alive:
        progressed = False
        # asynchronous completions respect issue order per engine (TMA, tensor core), not across engines
        q = rng.choice([tma_q, mma_q])
        if q and rng.random() < 0.7:
            q.pop(0)()
            progressed = True
        a = rng.choice(alive)
        before = (tuple(b.phase for b in full + empty + tmem_full + tmem_empty), tuple(stage_state), tuple(acc_state),
                  len(tma_q) + len(mma_q), len(done_tiles))
        try:
            next(a)
        except StopIteration:
            alive.remove(a)
            progressed = True
        after = (tuple(b.phase for b in full + empty + tmem_full + tmem_empty), tuple(stage_state), tuple(acc_state),
                 len(tma_q) + len(mma_q), len(done_tiles))
        progressed = progressed or before != after
        idle_rounds = 0 if progressed else idle_rounds + 1
        assert idle_rounds < 20000, "deadlock: no actor can make progress"
    for q in (tma_q, mma_q):
        while q:
            q.pop(0)()
    assert done_tiles == my_tiles
    assert all(s == "free" for s in stage_state)


@pytest.mark.parametrize("cfg", [
    # (total_tiles, grid, total_kb, stages)
    (1, 1, 1, 4), (2, 1, 1, 4), (3, 1, 9, 6), (7, 2, 9, 6), (5, 1, 18, 4), (9, 4, 2, 8), (6, 1, 1, 8), (4, 1, 72, 4),
])
def test_persistent_kernel_protocol_never_deadlocks_or_races(cfg):
    total_tiles, grid, total_kb, stages = cfg
    for block in range(grid):
        for seed in range(60):
            simulate(total_tiles, grid, block, total_kb, stages, seed)


@pytest.mark.parametrize("bug", ["empty_parity", "early_release"])
def test_the_model_catches_protocol_mistakes(bug):
    """Sanity of the checker itself.  `empty_parity`: the MMA thread waits on tmem_empty with the
    un-inverted parity -> nothing ever arrives on a fresh barrier (deadlock on the first tile).
    `early_release`: tmem_empty initialised for one arrival instead of one per epilogue warp -> the MMA
    thread re-uses an accumulator while other warps are still draining it."""
    with pytest.raises(AssertionError):
        for seed in range(40):
            simulate(6, 1, 0, 4, 4, seed, bug=bug)


# ---------------------------------------------------------------------------------------------------
# bn_bwd_apply_pool_tma_kernel (csrc/elementwise.cu): bulk loads into a 3-stage ring, two output buffers
# drained by bulk stores that are tracked with bulk groups
# ---------------------------------------------------------------------------------------------------
def simulate_bn_tma(num_chunks, grid, block, stages, seed, bug=None):
    rng = random.Random(seed)
    full = [MBar(1) for _ in range(stages)]
    mine = list(range(block, num_chunks, grid))
    stage_state = ["free"] * stages          # free -> loading -> loaded -> free (after the block-wide barrier)
    stage_chunk = [None] * stages
    out_state = ["free", "free"]             # free -> written -> draining -> free
    loads, stores = [], []                   # in-flight asynchronous copies (complete in any order across queues)
    written = []

    def issue(j, stage):
        if j >= len(mine):
            return
        assert stage_state[stage] == "free", "bulk load into a stage that is still being read"
        stage_state[stage], stage_chunk[stage] = "loading", mine[j]

        def land(s=stage):
            stage_state[s] = "loaded"
            full[s].arrive()
        loads.append(land)

    def drain_some():
        for q in (loads, stores):
            while q and rng.random() < 0.5:
                q.pop(rng.randrange(len(q)) if q is loads else 0)()     # bulk groups complete in order

    for s in range(stages):
        issue(s, s)
    stage, phase = 0, 0
    for j, chunk in enumerate(mine):
        ob = j & 1
        # thread 0: cp.async.bulk.wait_group.read 1 -> every store but the newest has finished reading smem
        keep = 2 if bug == "no_wait" else 1
        while len(stores) > keep:
            stores.pop(0)()
        spins = 0
        while not full[stage].passed(phase):
            drain_some()
            if loads and rng.random() < 0.7:
                loads.pop(rng.randrange(len(loads)))()
            spins += 1
            assert spins < 10000, "deadlock waiting for a stage"
        # __syncthreads; compute from the stage into the output buffer
        assert stage_state[stage] == "loaded" and stage_chunk[stage] == chunk, "wrong data in the stage"
        assert out_state[ob] == "free", "output buffer overwritten while its bulk store is still reading it"
        out_state[ob] = "written"
        # fence.proxy.async; __syncthreads; thread 0: bulk store + commit, refill the stage just consumed
        stage_state[stage] = "free"
        out_state[ob] = "draining"

        def stored(b=ob, c=chunk):
            out_state[b] = "free"
            written.append(c)
        stores.append(stored)
        issue(j + stages, stage)
        drain_some()
        stage += 1
        if stage == stages:
            stage, phase = 0, phase ^ 1
    while stores:
        stores.pop(0)()
    assert sorted(written) == mine and not loads


@pytest.mark.parametrize("cfg", [(1, 1, 3), (2, 1, 3), (3, 1, 3), (10, 1, 3), (17, 4, 3), (9, 2, 2), (40, 3, 4)])
def test_bn_tma_pipeline_protocol(cfg):
    num_chunks, grid, stages = cfg
    for block in range(grid):
        for seed in range(80):
            simulate_bn_tma(num_chunks, grid, block, stages, seed)


def test_bn_tma_model_catches_a_missing_store_wait():
    with pytest.raises(AssertionError):
        for seed in range(200):
            simulate_bn_tma(12, 1, 0, 3, seed, bug="no_wait")
