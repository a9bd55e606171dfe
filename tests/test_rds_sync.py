"""CPU tests of the RDS block / group synchroniser (rds.cu: RdsSync, host logic only) against Demod_WFM::processNewRdsBit +
checkBlock + getNextRdsGroupData compiled from the reference (demod_wfm.cpp:584-788, oracle/_ref/libpebbleref_app.so)."""
import numpy as np
import pytest

from rds_signal import GROUPS, stream_bits, group_bits, checkword, OFFSET


def _ref():
    from oracle import ref_app
    if not ref_app.available():
        pytest.skip("oracle/_ref not built")
    return ref_app.Demod_WFM(247000, 2048)


def _drain(obj):
    out = []
    while True:                      # a None is "queue empty" OR "same as the last group": ask until it stays None
        g = obj.getNextRdsGroupData()
        if g is None:
            g = obj.getNextRdsGroupData()
            if g is None:
                break
        out.append(g)
    return out


def _both(bits, step=997):
    import pebblesdr_b200 as pb
    r, g = _ref(), pb.RdsSync()
    want, got = [], []
    for i in range(0, len(bits), step):                 # groups are polled between pushes, like the GUI thread does
        r.pushRdsBits(bits[i:i + step]); g.pushBits(bits[i:i + step])
        want.append(_drain(r)); got.append(_drain(g))
    return want, got


def test_code_tables_from_the_generator_polynomial():
    """The checkword generator used by the test signal and the synchroniser's syndrome tables must agree: every block of
    a clean stream is accepted, and the four groups come out in order."""
    import pebblesdr_b200 as pb
    bits = stream_bits(GROUPS, 104 * 12)
    g = pb.RdsSync()
    g.pushBits(bits)
    out = _drain(g)
    assert out[:4] == GROUPS and len(out) == 12


def test_clean_stream_and_arbitrary_bit_offset():
    for lead in (0, 1, 25, 26, 57, 103):
        bits = np.concatenate([np.random.default_rng(lead).integers(0, 2, lead).astype(np.uint8), stream_bits(GROUPS, 104 * 9)])
        want, got = _both(bits)
        assert got == want
        assert sum(len(w) for w in want) >= 7


def test_burst_errors_corrected_and_uncorrectable_blocks_resync():
    rng = np.random.default_rng(7)
    bits = stream_bits(GROUPS, 104 * 60).copy()
    for grp in range(3, 40, 3):                         # bursts of 1..5 bits inside one block: Meggitt decoder corrects them
        p = 104 * grp + int(rng.integers(0, 4)) * 26 + int(rng.integers(0, 16))
        L = 1 + grp % 5
        bits[p:p + L] ^= 1
    for grp in (44, 45):                                # scattered errors in one block: not correctable -> group skipped
        p = 104 * grp + 26
        bits[[p + 1, p + 9, p + 17]] ^= 1
    want, got = _both(bits)
    assert got == want
    flat = [g for w in want for g in w]
    assert len(flat) > 40


def test_loss_of_signal_posts_zero_group_and_resyncs():
    rng = np.random.default_rng(11)
    good = stream_bits(GROUPS, 104 * 10)
    junk = rng.integers(0, 2, 26 * 9).astype(np.uint8)  # more than BLOCK_ERROR_LIMIT bad blocks in a row
    bits = np.concatenate([good, junk, good, rng.integers(0, 2, 5000).astype(np.uint8), good])
    want, got = _both(bits, step=313)
    assert got == want
    flat = [g for w in want for g in w]
    assert (0, 0, 0, 0) in flat


def test_random_bits_never_disagree():
    rng = np.random.default_rng(3)
    bits = rng.integers(0, 2, 400000).astype(np.uint8)
    # sprinkle valid blocks so the state machines leave bit sync now and then
    for k in range(0, len(bits) - 300, 2777):
        bits[k:k + 208] = stream_bits(GROUPS[k % 4:] + GROUPS, 208)
    want, got = _both(bits, step=4001)
    assert got == want
