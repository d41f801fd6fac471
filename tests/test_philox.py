"""CPU: the Philox4x32-10 restatement (oracle/philox.py, twin of gtn_common.cuh:philox4x32_10) against the
Random123 known-answer vectors (Salmon et al., SC'11; kat_vectors of the Random123 distribution)."""
import numpy as np

from oracle import philox as PX

KAT = [  # (counter, key, expected)
    ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def test_philox_known_answers():
    for ctr, key, exp in KAT:
        got = tuple(int(x) for x in PX.philox4x32_10(*ctr, *key))
        assert got == exp, f"ctr {ctr} key {key}: {[hex(g) for g in got]}"


def test_keep_mask_scheme():
    m = PX.keep_mask(1234, PX.SITE_ATT_H, 7, (3, 1024), 0.1)
    assert m.shape == (3, 1024) and set(np.unique(m)) <= {0.0, 1.0}
    assert abs(m.mean() - 0.9) < 0.03
    bits = PX.random_bits(1234, PX.SITE_ATT_H, 7, 3 * 1024)
    assert np.array_equal(m.reshape(-1), (bits >= PX.threshold(0.1)).astype(np.float32))
    # element numbering: element e uses counter (e >> 2, step, site, 0), lane e & 3
    r = PX.philox4x32_10(5, 7, PX.SITE_ATT_H, 0, 1234, 0)
    assert [int(x) for x in r] == [int(b) for b in bits[20:24]]
    assert PX.threshold(0.5) == 1 << 31 and PX.threshold(0.0) == 0
    # different site / step / seed -> different streams
    assert not np.array_equal(bits, PX.random_bits(1234, PX.SITE_DEC_H, 7, 3 * 1024))
    assert not np.array_equal(bits, PX.random_bits(1234, PX.SITE_ATT_H, 8, 3 * 1024))
    assert not np.array_equal(bits, PX.random_bits(1235, PX.SITE_ATT_H, 7, 3 * 1024))
