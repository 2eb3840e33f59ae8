"""mbarrier protocol of the bulk-copy BatchNorm pipelines (bn_bwd_apply_pool_tma_kernel in csrc/elementwise.cu
and the generic skeleton of csrc/bn_stream.cu), executed as a randomly scheduled model on the host.

The kernels cannot run here, but their synchronisation is a small protocol between a producer thread (bulk
loads into a ring of stages, bulk stores out of two output buffers) and the compute threads, over `full[STAGES]`
barriers with parity waits, block barriers and bulk-group waits.  The model follows the kernel line by line and
checks, under thousands of random interleavings, that it never deadlocks and never lets a load overwrite a
stage that is still being read or a store read a buffer that is still being written.
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
