"""Index arithmetic of the fused wgrad -> reduce-scatter path, emulated on the host.

The CUDA side (conv_wgrad_push_kernel's tile hand-off in csrc/conv_igemm.cu and reduce_staged_sgd_kernel
in csrc/collectives.cu) cannot run here; what can be checked without a GPU is that the two halves agree on
who owns which float4 and where its staged copy lives.  The emulation below uses the same formulas,
line for line, on random bucket geometries and world sizes.
"""
import numpy as np
import pytest


def _emulate(world, start, length, flat_off, rows, ld, block_n, seed):
    rng = np.random.default_rng(seed)
    n = start + length + 64
    grads = [rng.standard_normal(n).astype(np.float32) for _ in range(world)]
    n4 = length >> 2
    per4 = (n4 + world - 1) // world
    stage_base = 8                                   # some other bucket's area comes first
    stage = [np.full(stage_base + world * per4 * 4, np.nan, np.float32) for _ in range(world)]
    pushed_lo, pushed_hi = flat_off, flat_off + rows * ld

    # --- conv_wgrad_push_kernel, tile hand-off (one "tile" = 128 rows x block_n columns)
    for rank in range(world):
        for co_tile in range((rows + 127) // 128):
            for ci_tile in range(ld // block_n):
                for rr in range(128):
                    co_r = co_tile * 128 + rr
                    if co_r >= rows:
                        break
                    row_off = co_r * ld + ci_tile * block_n
                    i4_row = (flat_off + row_off - start) >> 2
                    for c4 in range(block_n >> 2):
                        i4 = i4_row + c4
                        owner = i4 // per4
                        if owner != rank:
                            src = flat_off + row_off + 4 * c4
                            dst = stage_base + ((rank * per4 + (i4 - owner * per4)) << 2)
                            stage[owner][dst:dst + 4] = grads[rank][src:src + 4]

    # --- reduce_staged_sgd_kernel, reduction part
    out = [g.copy() for g in grads]
    for rank in range(world):
        lo = min(per4 * rank, n4)
        hi = min(lo + per4, n4)
        for i in range(lo, hi):
            e = start + (i << 2)
            if pushed_lo <= e < pushed_hi:
                g = grads[rank][e:e + 4].copy()
                for r in range(world):
                    if r != rank:
                        s = stage_base + ((r * per4 + (i - lo)) << 2)
                        g += stage[rank][s:s + 4]
            else:
                g = sum(grads[r][e:e + 4] for r in range(world))
            out[rank][e:e + 4] = g / world
    return grads, out, n4, per4


@pytest.mark.parametrize("world", [2, 3, 4, 8])
@pytest.mark.parametrize("geom", [
    # (start, bucket length, tensor offset inside the flat buffer, rows, row length, column tile)
    (0, 128 * 576 + 512, 256, 128, 576, 64),          # tensor in the middle of its bucket, small tensors around it
    (1024, 256 * 1152 + 1024 + 12, 1024 + 8, 256, 1152, 128),
    (4096, 64 * 576 + 132, 4096 + 132, 64, 576, 64),  # fewer rows than a tile, tensor at the end of the bucket
])
def test_owner_and_staging_indices_agree(world, geom):
    start, length, flat_off, rows, ld, block_n = geom
    assert length % 4 == 0 and flat_off % 4 == 0 and flat_off + rows * ld <= start + length
    grads, out, n4, per4 = _emulate(world, start, length, flat_off, rows, ld, block_n, seed=world)
    mean = sum(grads) / world
    for rank in range(world):
        lo, hi = min(per4 * rank, n4), min(per4 * rank + per4, n4)
        sl = slice(start + 4 * lo, start + 4 * hi)
        assert not np.isnan(out[rank][sl]).any(), "an owned float4 was never staged"
        np.testing.assert_allclose(out[rank][sl], mean[sl], rtol=1e-5, atol=1e-6)
