"""The argument the carve kernel rests on, checked on its own (CPU): the reference's per-view state machine
(kernels/backprojection.c:57-84, run view after view by cl.py:190-227) has an ORDER-INDEPENDENT result that equals a
MINIMUM over per-view outcomes in the device coding 0 carved < 1 seen < 2 unseen.  That is what lets the kernel
(i) visit the views in farthest-point order, (ii) test several views of one voxel at the same time in the sparse
phase and combine them with `min`, (iii) MIN-all-reduce view-sharded replicas, and (iv) expand to the reference's
int32 values only at read-back (`expand_state_kernel` / `tail_kernel`), for every seed value of the grid.
"""
import numpy as np
import pytest

MISS, ZERO, NONZERO = 2, 0, 1  # outcome of one voxel against one view, already in the device coding


def reference_sequential(outcomes, default_value):
    """kernels/backprojection.c:57-84 applied view after view to one voxel."""
    s = default_value
    for o in outcomes:
        if s == -1:       # :65-67 carved voxels return at once
            continue
        if o == MISS:     # :70-72 outside the picture / behind the camera: state untouched
            continue
        if o == ZERO:     # :79-80
            s = -1
        elif s == 0:      # :81-83 `seen` only promotes a grid value of 0
            s = 1
    return s


def device_result(outcomes, default_value, rng):
    """min-combine over an arbitrary order and an arbitrary grouping into simultaneous passes, then the read-back
    expansion of csrc/bp_b200.cu (values_to_device)."""
    order = rng.permutation(len(outcomes))
    st, q = 2, 0
    while q < len(order):
        g = int(rng.choice([1, 2, 4, 8, 16, 32]))
        if st == 0:
            break                                    # a carved voxel is skipped for the remaining views
        st = min([st] + [outcomes[v] for v in order[q:q + g]])
        q += g
    d = default_value
    if d == -1:
        return -1
    return {0: -1, 1: 1 if d == 0 else d, 2: d}[st]


@pytest.mark.parametrize("default_value", [0, 1, 5, -1, -3])
def test_min_combine_in_any_order_equals_the_sequential_state_machine(default_value):
    rng = np.random.default_rng(default_value + 10)
    for trial in range(3000):
        n = int(rng.integers(0, 12))
        p = rng.dirichlet([1, 1, 1])
        outcomes = [int(o) for o in rng.choice([MISS, ZERO, NONZERO], size=n, p=p)]
        assert device_result(outcomes, default_value, rng) == reference_sequential(outcomes, default_value), outcomes


def test_replicas_min_reduce_like_the_sequential_run():
    """View sharding: every rank carves its own views from the seed; the element-wise MIN of the int8
    states is the single-GPU state.  (Plain min over the reference's -1 / 0 / 1 would be wrong: min(0, 1) = 0.)"""
    rng = np.random.default_rng(0)
    for trial in range(2000):
        n, world = int(rng.integers(1, 10)), int(rng.integers(1, 5))
        outcomes = [int(o) for o in rng.choice([MISS, ZERO, NONZERO], size=n)]
        ranks = [min([2] + outcomes[r::world]) for r in range(world)]
        st = min(ranks)
        assert {0: -1, 1: 1, 2: 0}[st] == reference_sequential(outcomes, 0)
    assert min(0, 1) == 0 and reference_sequential([NONZERO], 0) == 1  # the counter-example for the naive coding


def test_speculative_readback_invariant():
    """The argument bp_set_readback_target rests on (csrc/bp_b200.cu, start_/finish_spec_readback), on the CPU: the
    host volume holds the expansion of a SNAPSHOT of the int8 state; later views only ever lower a voxel's state
    (2 unseen -> 1 seen -> 0 carved); rewriting every (i, j) row whose state differs from the snapshot - padding
    included in the comparison, excluded from the write - leaves exactly the expansion of the final state."""
    rng = np.random.default_rng(11)
    ni, ny, nz, nzp = 5, 7, 37, 48
    for default_value in (0, 5, -1):
        d = default_value
        lut = np.array([-1, -1 if d == -1 else (1 if d == 0 else d), d], dtype=np.int32)  # carve_values()
        snap = rng.integers(0, 3, (ni * ny, nzp)).astype(np.int8)
        snap[:, nz:] = rng.integers(-128, 127, (ni * ny, nzp - nz))          # padding holds anything, in both copies
        final = snap.copy()
        rows = rng.random(ni * ny) < 0.3                                    # some rows change after the snapshot
        lower = rng.integers(0, 3, (ni * ny, nz)).astype(np.int8)
        final[rows, :nz] = np.minimum(final[rows, :nz], lower[rows])
        host = lut[snap[:, :nz]]                                            # the early copy
        differs = (final != snap).any(axis=1)                               # patch_rows_kernel's test (whole row)
        assert not differs[~rows].any()
        host[differs] = lut[final[differs, :nz]]
        assert np.array_equal(host, lut[final[:, :nz]])


def test_and_coding_of_the_view_segments_is_the_minimum():
    """carve_segments_kernel merges the segments of one patch with an atomic AND on a transient coding of the states
    (0 carved, 1 seen, 3 unseen instead of 2; bp_kernels.cu: classify_items_kernel / decode_items_kernel).  The claim it
    rests on, byte-parallel as the kernels apply it on 32-bit words: decode(AND of encoded states) == minimum of the
    states, for any number of operands - and the minimum over the views is what kernels/backprojection.c:65-83 computes."""
    rng = np.random.default_rng(3)

    def encode(w):
        return w | ((w >> 1) & np.uint32(0x01010101))

    def decode(w):
        return w - ((w >> 1) & np.uint32(0x01010101))

    for n_ops in (1, 2, 3, 7):
        states = rng.integers(0, 3, size=(n_ops, 4096, 4), dtype=np.uint8)           # 4 voxels per word
        words = states.view(np.uint32)[..., 0]                                        # [n_ops, 4096]
        enc = encode(words)
        assert set(np.unique(enc.view(np.uint8))) <= {0, 1, 3}
        merged = np.bitwise_and.reduce(enc, axis=0)
        got = decode(merged).view(np.uint8).reshape(-1, 4)
        np.testing.assert_array_equal(got, states.min(axis=0))
    # round trip on every byte value the state can hold, padding included (it holds 0, 1 or 2 as well)
    w = np.array([0x00010200, 0x02020202, 0x01000102], dtype=np.uint32)
    np.testing.assert_array_equal(decode(encode(w)), w)
