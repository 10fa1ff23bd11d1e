"""The carry proof behind the fire-and-forget adds (DESIGN.md section 4, K12 scatter): cells are uint8,
four to a 32-bit word, and an update adds t << 8*(g & 3) to the word.  Claim: with the matrix zero
before the pass, (sum of all bytes == sum of all t) <=> no cell's true sum exceeded 255.  Restated
with numpy and checked on random update streams, including streams built to overflow."""
import numpy as np


def _apply(n_words, upd_word, upd_lane, upd_t):
    words = np.zeros(n_words, dtype=np.uint64)
    np.add.at(words, upd_word, upd_t.astype(np.uint64) << (8 * upd_lane.astype(np.uint64)))
    words &= np.uint64(0xFFFFFFFF)                                   # 32-bit wrap of the hardware add
    true = np.zeros((n_words, 4), dtype=np.int64)
    np.add.at(true, (upd_word, upd_lane), upd_t.astype(np.int64))
    byte_sum = int(sum(((words >> np.uint64(8 * k)) & np.uint64(0xFF)).sum() for k in range(4)))
    return byte_sum, int(upd_t.sum()), true


def test_byte_sum_equals_total_iff_no_cell_exceeds_255():
    rng = np.random.default_rng(5)
    seen = {True: 0, False: 0}
    for trial in range(300):
        n_words = int(rng.integers(1, 40))
        n_upd = int(rng.integers(0, 400))
        hot = trial % 3 == 0                                         # concentrate updates so that cells overflow
        w = rng.integers(0, 1 if hot else n_words, size=n_upd)
        lane = rng.integers(0, 4, size=n_upd)
        t = rng.integers(1, 33 if trial % 2 else 256, size=n_upd)
        byte_sum, total, true = _apply(n_words, w, lane, t)
        no_overflow = bool((true <= 255).all())
        assert (byte_sum == total) == no_overflow, (trial, byte_sum, total)
        assert byte_sum <= total
        seen[no_overflow] += 1
    assert seen[True] > 30 and seen[False] > 30


def test_exactly_255_is_not_a_carry_and_256_is():
    for lane in range(4):
        w = np.zeros(2, dtype=np.int64)
        l = np.full(2, lane)
        b, tot, _ = _apply(1, w, l, np.array([200, 55]))
        assert b == tot == 255
        b, tot, _ = _apply(1, w, l, np.array([200, 56]))
        assert tot == 256 and b == (1 if lane < 3 else 0)          # the carry moved into the next byte / out of the word
