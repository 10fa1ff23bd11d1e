"""CPU tests of the host half of the narrowed lcp transfer (lime_b200/csrc/host_narrow.cpp):
out[i] = min(lcp[i], 255), plain and through the chunk ring lime_load(LIME_MEM_HOST) uses, and the
claim it rests on: for alpha <= 255, `min(lcp, 255) < alpha` is `lcp < alpha` (ClusterLCP.cpp:150-202
only ever compares lcp with alpha)."""
import ctypes as C

import numpy as np
import pytest

from lime_b200 import _lib


def _lcp(n, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 300, n, dtype=np.uint32)
    if n:
        k = max(1, n // 16)
        a[rng.integers(0, n, k)] = rng.integers(0, 2**32, k, dtype=np.uint64).astype(np.uint32)
        a[rng.integers(0, n, 4)] = np.array([0, 254, 255, 256], dtype=np.uint32)[: min(4, n)] if n >= 4 else a[:1]
        a[-1] = 0xFFFFFFFF
    return a


def _ptr(a):
    return C.c_void_p(a.ctypes.data)


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 63, 64, 65, 4097, (1 << 20) + 3, 3 * (1 << 20) + 77])
@pytest.mark.parametrize("threads", [1, 3, 8])
def test_narrow_plain(n, threads):
    lib = _lib.load()
    lcp = _lcp(n, 7 + n)
    out = np.full(n + 8, 0xAB, dtype=np.uint8)
    assert lib.lime_host_narrow_lcp(_ptr(lcp), _ptr(out), n, threads) == 0
    assert np.array_equal(out[:n], np.minimum(lcp, 255).astype(np.uint8))
    assert (out[n:] == 0xAB).all()           # nothing written past the end


@pytest.mark.parametrize("delay_us", [0, 50])
@pytest.mark.parametrize("n,chunk,threads,slots", [
    (1, 64, 2, 1), (64, 64, 3, 2), (65, 64, 3, 3), (1000, 64, 5, 3), (100003, 4096, 7, 3),
    (100003, 4096, 1, 1), (1 << 20, 1 << 16, 4, 2), ((1 << 21) + 11, 100000, 15, 3), (5, 1 << 20, 16, 3)])
def test_narrow_pipelined(n, chunk, threads, slots, delay_us):
    """narrow_transfer (the loop lime_load runs) over a simulated in-order asynchronous copy queue: a stage
    slot rewritten before its (delayed) copy has run would corrupt the output."""
    lib = _lib.load()
    lcp = _lcp(n, 11 + n)
    out = np.full(n + 8, 0xCD, dtype=np.uint8)
    assert lib.lime_host_narrow_lcp_pipelined(_ptr(lcp), _ptr(out), n, chunk, threads, slots, delay_us) == 0
    assert np.array_equal(out[:n], np.minimum(lcp, 255).astype(np.uint8))
    assert (out[n:] == 0xCD).all()


def test_narrow_pipelined_rejects_bad_arguments():
    lib = _lib.load()
    lcp = _lcp(10, 1)
    out = np.zeros(10, dtype=np.uint8)
    assert lib.lime_host_narrow_lcp_pipelined(_ptr(lcp), _ptr(out), 10, 0, 1, 1, 0) != 0
    assert lib.lime_host_narrow_lcp_pipelined(_ptr(lcp), _ptr(out), 10, 4, 1, 0, 0) != 0
    assert lib.lime_host_narrow_lcp(None, _ptr(out), 10, 1) != 0
    assert lib.lime_host_threads() >= 1


def test_clamped_lcp_decides_every_head_test_for_alpha_up_to_255():
    lcp = _lcp(1 << 16, 3)
    lcp[:600] = np.arange(600, dtype=np.uint32)          # every value around the clamp
    nar = np.minimum(lcp, 255).astype(np.uint8)
    for alpha in (1, 2, 12, 16, 24, 32, 254, 255):
        assert np.array_equal(lcp < alpha, nar < alpha)
    # and the limit is tight: alpha = 256 is not decided by the byte
    assert not np.array_equal(lcp < 256, nar.astype(np.uint32) < 256)


# ---- the bit transfer of lime_set_alpha: bit j of out[i] = lcp[8 i + j] < alpha ---------------------------------
def _bits(lcp, alpha):
    pad = np.ones((-lcp.shape[0]) % 8, dtype=bool)            # positions beyond n read as heads
    return np.packbits(np.concatenate([lcp < alpha, pad]), bitorder="little")


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 31, 32, 33, 511, 512, 513, 4097, (1 << 20) + 3, 3 * (1 << 20) + 77])
@pytest.mark.parametrize("threads,alpha", [(1, 16), (3, 1), (8, 255), (2, 256), (5, 0x80000001)])
def test_pack_plain(n, threads, alpha):
    lib = _lib.load()
    lcp = _lcp(n, 5 + n)
    nb = (n + 7) // 8
    out = np.full(nb + 8, 0xAB, dtype=np.uint8)
    assert lib.lime_host_pack_lcp(_ptr(lcp), _ptr(out), n, alpha, threads) == 0
    assert np.array_equal(out[:nb], _bits(lcp, alpha))
    assert (out[nb:] == 0xAB).all()


@pytest.mark.parametrize("delay_us", [0, 50])
@pytest.mark.parametrize("n,chunk,threads,slots", [
    (1, 512, 2, 1), (512, 512, 3, 2), (513, 512, 3, 3), (5000, 512, 5, 3), (100003, 4096, 7, 3),
    (100003, 4096, 1, 1), (1 << 20, 1 << 16, 4, 2), ((1 << 21) + 11, 1 << 17, 15, 3), (5, 1 << 20, 16, 3)])
def test_pack_pipelined(n, chunk, threads, slots, delay_us):
    lib = _lib.load()
    lcp = _lcp(n, 13 + n)
    nb = (n + 7) // 8
    out = np.full(nb + 8, 0xCD, dtype=np.uint8)
    assert lib.lime_host_pack_lcp_pipelined(_ptr(lcp), _ptr(out), n, chunk, 16, threads, slots, delay_us) == 0
    assert np.array_equal(out[:nb], _bits(lcp, 16))
    assert (out[nb:] == 0xCD).all()


def test_pack_pipelined_rejects_bad_arguments():
    lib = _lib.load()
    lcp = _lcp(10, 1)
    out = np.zeros(10, dtype=np.uint8)
    assert lib.lime_host_pack_lcp_pipelined(_ptr(lcp), _ptr(out), 10, 100, 16, 1, 1, 0) != 0     # chunk not a multiple of 512
    assert lib.lime_host_pack_lcp_pipelined(_ptr(lcp), _ptr(out), 10, 512, 0, 1, 1, 0) != 0       # alpha 0
    assert lib.lime_host_pack_lcp(None, _ptr(out), 10, 16, 1) != 0


# ---- da as 3 bytes per position (document ids below 2^24) ------------------------------------------------------------
def _da3(da):
    return da.astype("<u4").view(np.uint8).reshape(-1, 4)[:, :3].reshape(-1)


@pytest.mark.parametrize("n", [0, 1, 7, 8, 9, 15, 16, 17, 31, 33, 4097, (1 << 20) + 3])
def test_pack_da3_plain(n):
    lib = _lib.load()
    da = np.random.default_rng(n).integers(0, 1 << 24, n, dtype=np.uint32)
    if n:
        da[0], da[-1] = 0xFFFFFF, 0xABCDEF
    for shift in (0, 1, 4):                      # 16-byte aligned output (non-temporal stores) and not
        buf = np.full(3 * n + 64, 0xAB, dtype=np.uint8)
        off = (-buf.ctypes.data) % 16 + shift
        out = buf[off:]
        assert lib.lime_host_pack_da3(_ptr(da), C.c_void_p(buf.ctypes.data + off), n) == 0
        assert np.array_equal(out[:3 * n], _da3(da))
        assert (out[3 * n:] == 0xAB).all() and (buf[:off] == 0xAB).all()    # nothing written outside (the vector loop stores 16 for 12)


@pytest.mark.parametrize("delay_us", [0, 50])
@pytest.mark.parametrize("n,chunk,threads,slots", [
    (1, 512, 2, 1), (513, 512, 3, 3), (5000, 512, 5, 3), (100003, 4096, 7, 3), ((1 << 21) + 11, 1 << 17, 15, 3), (5, 1 << 20, 16, 3)])
def test_pack_lcp_bits_and_da3_pipelined(n, chunk, threads, slots, delay_us):
    """Both halves of a stage slot (lcp bits, da as 3 bytes) through the loop lime_load runs."""
    lib = _lib.load()
    lcp = _lcp(n, 17 + n)
    da = np.random.default_rng(n + 1).integers(0, 1 << 24, n, dtype=np.uint32)
    nb = (n + 7) // 8
    out = np.full(nb + 8, 0xCD, dtype=np.uint8)
    out3 = np.full(3 * n + 8, 0xEF, dtype=np.uint8)
    assert lib.lime_host_pack_lcp_da3_pipelined(_ptr(lcp), _ptr(da), _ptr(out), _ptr(out3), n, chunk, 16, threads, slots, delay_us) == 0
    assert np.array_equal(out[:nb], _bits(lcp, 16)) and (out[nb:] == 0xCD).all()
    assert np.array_equal(out3[:3 * n], _da3(da)) and (out3[3 * n:] == 0xEF).all()
