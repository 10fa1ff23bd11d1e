"""The segmented OR-fill of the stream kernel (lime_b200/csrc/k12_stream.cu: fill_up / fill_down) spreads a flag over
the positions of a segment with ONE addition instead of a Kogge-Stone ladder: with b = c | x the carry into bit i of
x + b is the OR of the x bits below i inside its run.  Restated in numpy and checked against the ladder."""
import numpy as np


def ladder_up(x, c):
    x, c = x.copy(), c.copy()
    for k in (1, 2, 4, 8, 16):
        x |= (x << np.uint32(k)) & c
        c &= c << np.uint32(k)
    return x


def ladder_down(x, c):
    x, c = x.copy(), c.copy()
    for k in (1, 2, 4, 8, 16):
        x |= (x >> np.uint32(k)) & c
        c &= c >> np.uint32(k)
    return x


def brev(v):
    v = v.copy()
    v = ((v >> np.uint32(1)) & np.uint32(0x55555555)) | ((v & np.uint32(0x55555555)) << np.uint32(1))
    v = ((v >> np.uint32(2)) & np.uint32(0x33333333)) | ((v & np.uint32(0x33333333)) << np.uint32(2))
    v = ((v >> np.uint32(4)) & np.uint32(0x0F0F0F0F)) | ((v & np.uint32(0x0F0F0F0F)) << np.uint32(4))
    v = ((v >> np.uint32(8)) & np.uint32(0x00FF00FF)) | ((v & np.uint32(0x00FF00FF)) << np.uint32(8))
    return (v >> np.uint32(16)) | (v << np.uint32(16))


def add_up(x, c):
    b = c | x
    carry_in = (x + b) ^ x ^ b                       # uint32 addition wraps like the kernel's
    return x | (carry_in & c)


def add_down(x, c):
    return brev(add_up(brev(x), brev(c)))


def test_add_carry_fill_equals_the_ladder():
    rng = np.random.default_rng(1)
    n = 2_000_000
    x = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    c = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    x[::3] &= rng.integers(0, 1 << 32, x[::3].shape[0], dtype=np.uint64).astype(np.uint32)       # sparse flags
    c[::5] |= rng.integers(0, 1 << 32, c[::5].shape[0], dtype=np.uint64).astype(np.uint32)       # long runs
    edge = np.array([0, 1, 0x80000000, 0xFFFFFFFF, 0x7FFFFFFF, 0xFFFFFFFE], dtype=np.uint32)
    x = np.concatenate([x, np.repeat(edge, edge.shape[0])])
    c = np.concatenate([c, np.tile(edge, edge.shape[0])])
    with np.errstate(over="ignore"):
        assert np.array_equal(add_up(x, c), ladder_up(x, c))
        assert np.array_equal(add_down(x, c), ladder_down(x, c))
