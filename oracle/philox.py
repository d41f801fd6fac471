"""Counter-based RNG shared by the oracle and the CUDA kernels (TEST INFRASTRUCTURE).

This file is part of ``oracle/``: only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s cpu_baseline / ``--impl reference`` legs may import it.

The reference draws dropout masks from torch's device RNG (``F.dropout`` at
/root/reference/model.py:104,148,149,393,408 and ``nn.Dropout`` at
/root/reference/layers.py:55), which no second implementation can reproduce by seed.
The B200 kernels therefore use a documented counter-based generator,
Philox4x32-10 (Salmon et al., SC'11), keyed as

    key     = (seed_lo, seed_hi)
    counter = (elem >> 2, step, site, 0)       # elem = row * width + col
    r       = philox(counter, key)[elem & 3]   # 32 random bits for that element
    keep    = r >= floor(p * 2**32)            # Bernoulli(1 - p)

and this module is the numpy restatement of the same function, used to build the
identical masks on the host so that reference / oracle / kernel can be compared
element for element.  ``gantron_b200/csrc/gtn_common.cuh:philox4x32_10`` is the
device twin; tests/test_philox.py pins both against the Random123 known-answer vectors.
"""
from __future__ import annotations

import numpy as np

M0 = np.uint64(0xD2511F53)
M1 = np.uint64(0xCD9E8D57)
W0 = 0x9E3779B9
W1 = 0xBB67AE85
MASK32 = np.uint64(0xFFFFFFFF)

# dropout sites (the `site` word of the counter)
SITE_PRENET1 = 1      # model.py:104 first prenet layer
SITE_PRENET2 = 2      # model.py:104 second prenet layer
SITE_ATT_H = 3        # model.py:393 attention_hidden dropout
SITE_DEC_H = 4        # model.py:408 decoder_hidden dropout
SITE_POSTNET0 = 16    # model.py:148/149, + layer index
SITE_DISC0 = 32       # layers.py:55, + layer index
SITE_LIN0 = 40        # layers.py:68 DiscDense dropout of the LinearDiscriminator, + layer index
SITE_ENC0 = 60        # model.py:197 / 217 encoder convolutions, + layer index


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Vectorised Philox4x32-10. All inputs broadcastable uint32 arrays; returns 4 uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint64) & MASK32
    c1 = np.asarray(c1, dtype=np.uint64) & MASK32
    c2 = np.asarray(c2, dtype=np.uint64) & MASK32
    c3 = np.asarray(c3, dtype=np.uint64) & MASK32
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0 = int(k0) & 0xFFFFFFFF
    k1 = int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c0          # 64-bit products of 32-bit values: exact in uint64
        p1 = M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & MASK32
        c0, c1, c2, c3 = (hi1 ^ c1 ^ np.uint64(k0)), lo1, (hi0 ^ c3 ^ np.uint64(k1)), lo0
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return (c0.astype(np.uint32), c1.astype(np.uint32), c2.astype(np.uint32), c3.astype(np.uint32))


def random_bits(seed: int, site: int, step: int, n_elem: int) -> np.ndarray:
    """32 random bits for elements 0..n_elem-1 of (site, step)."""
    n4 = (n_elem + 3) // 4
    blk = np.arange(n4, dtype=np.uint64)
    r = philox4x32_10(blk, np.uint64(step), np.uint64(site), np.uint64(0),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    out = np.stack(r, axis=1).reshape(-1)
    return out[:n_elem]


def threshold(p: float) -> int:
    """Integer drop threshold: element kept iff bits >= threshold(p)."""
    return min(int(np.floor(float(p) * 4294967296.0)), 0xFFFFFFFF)


def keep_mask(seed: int, site: int, step: int, shape, p: float) -> np.ndarray:
    """float32 {0,1} keep mask of `shape` (row-major element numbering)."""
    n = int(np.prod(shape))
    bits = random_bits(seed, site, step, n)
    return (bits >= np.uint32(threshold(p))).astype(np.float32).reshape(shape)
