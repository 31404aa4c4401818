"""Philox4x32-10 in numpy, plus the random-stream layout shared with the kernels.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference draws from CPython's ``random`` (MT19937): ``random()`` for the
Metropolis test (``py_mc/mc.py:2,182``) and for exchange decisions
(``py_mc/mc_mpi.py:126``), ``randrange`` for site selection
(``examples/ising2D.py:32-33``).  A GPU sweep cannot replay a sequential MT
stream, so the product defines a *counter-based* stream (Philox4x32-10, Salmon
et al. SC'11) and parity is checked by feeding the reference the same numbers
through injected ``random``/``randrange`` objects.

Stream layout ("pmc stream v1"), restated in ``py_mc_b200/csrc/pmc_stream.cuh``:

``key      = (seed & 0xffffffff, seed >> 32)``
``counter  = (idx, replica, step_lo, domain<<28 | call<<24 | lane<<18 | step_hi[17:0])``

* domain 0 (Metropolis sweep): ``idx`` = word index ``row*nwords_per_row + wcol``
  of a 64-site word of one colour, ``step = 2*sweep + colour``.  The 32-bit uniform
  ``x32`` of lane ``l`` of the word has two parts.  *Head*: its 8 most significant
  bits are bit ``l`` of the 64-bit *bit planes* 0..7; call ``m`` in 0..3 (lane field 0)
  yields plane ``2m = o0 | o1<<32`` and plane ``2m+1 = o2 | o3<<32``, plane 0 is the
  MSB.  *Tail*: its 24 low bits are ``o0 >> 8`` of the lane's own call (call 15, lane
  field ``l``).  A kernel compares the head bit-serially for all 64 lanes at once
  (which decides 255 of 256 comparisons ``x32 <= thr``) and evaluates the tail only
  for lanes that are still undecided; the value is *defined* by both parts.
* domain 1 (initial configuration): ``idx`` = ``site >> 2`` (natural row-major site
  index), element ``site & 3`` of the output; spin = +1 iff ``x32 >= 2**31``
  (the ``random() >= 0.5`` rule of ``examples/ising2D.py:67-70``).
* domain 2 (replica exchange): ``idx`` = lower temperature index of the pair,
  ``step`` = exchange round, ``x32 = o0``.
* domain 3 (pair-exchange / Kawasaki moves): as domain 0 with ``step = 16*sweep + bond class``.
* domain 4 (hard discs): ``idx`` = replica-local trial index, see ``oracle/harddisc.py``.

The uniform handed to the reference is ``u = x32 * 2**-32`` (exact in float64).
"""
import numpy as np

PHILOX_M0 = np.uint64(0xD2511F53)
PHILOX_M1 = np.uint64(0xCD9E8D57)
PHILOX_W0 = 0x9E3779B9
PHILOX_W1 = 0xBB67AE85

DOMAIN_SWEEP = 0
DOMAIN_INIT = 1
DOMAIN_RX = 2
DOMAIN_PAIR = 3
DOMAIN_DISC = 4

_MASK32 = np.uint64(0xFFFFFFFF)


def philox4x32_10(ctr, key):
    """Philox4x32 with 10 rounds.

    ctr: uint32 array [..., 4]; key: (k0, k1) python ints.  Returns uint32 [..., 4].
    """
    ctr = np.asarray(ctr, dtype=np.uint32)
    c0 = ctr[..., 0].astype(np.uint64)
    c1 = ctr[..., 1].astype(np.uint64)
    c2 = ctr[..., 2].astype(np.uint64)
    c3 = ctr[..., 3].astype(np.uint64)
    k0 = int(key[0]) & 0xFFFFFFFF
    k1 = int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = PHILOX_M0 * c0
        p1 = PHILOX_M1 * c2
        hi0, lo0 = p0 >> np.uint64(32), p0 & _MASK32
        hi1, lo1 = p1 >> np.uint64(32), p1 & _MASK32
        n0 = hi1 ^ c1 ^ np.uint64(k0)
        n2 = hi0 ^ c3 ^ np.uint64(k1)
        c0, c1, c2, c3 = n0, lo1, n2, lo0
        k0 = (k0 + PHILOX_W0) & 0xFFFFFFFF
        k1 = (k1 + PHILOX_W1) & 0xFFFFFFFF
    out = np.stack([c0, c1, c2, c3], axis=-1)
    return out.astype(np.uint32)


def seed_key(seed):
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return (seed & 0xFFFFFFFF, seed >> 32)


HEAD_PLANES = 8
TAIL_CALL = 15


def make_counter(idx, replica, step, domain, call=0, lane=0):
    """Build counters [..., 4] following the stream layout above (idx and lane may be arrays)."""
    idx = np.asarray(idx, dtype=np.uint64)
    step = int(step)
    assert 0 <= step < (1 << 50)
    assert 0 <= call < 16 and 0 <= domain < 16
    ctr = np.empty(idx.shape + (4,), dtype=np.uint32)
    ctr[..., 0] = (idx & _MASK32).astype(np.uint32)
    ctr[..., 1] = np.uint32(int(replica) & 0xFFFFFFFF)
    ctr[..., 2] = np.uint32(step & 0xFFFFFFFF)
    ctr[..., 3] = (np.uint32((domain << 28) | (call << 24) | ((step >> 32) & 0x3FFFF)) |
                   (np.asarray(lane, dtype=np.uint32) << np.uint32(18)))
    return ctr


def word_planes(seed, replica, step, word_idx, call, domain=DOMAIN_SWEEP):
    """The two 64-bit planes (2*call, 2*call+1) for each word index."""
    out = philox4x32_10(make_counter(word_idx, replica, step, domain, call), seed_key(seed))
    o = out.astype(np.uint64)
    even = o[..., 0] | (o[..., 1] << np.uint64(32))
    odd = o[..., 2] | (o[..., 3] << np.uint64(32))
    return even, odd


def word_lane_x32(seed, replica, step, word_idx, domain=DOMAIN_SWEEP):
    """x32 for all 64 lanes of each word: uint32 array [..., 64] (head planes + per-lane tail)."""
    word_idx = np.asarray(word_idx, dtype=np.uint64)
    lanes = np.arange(64, dtype=np.uint64)
    x = np.zeros(word_idx.shape + (64,), dtype=np.uint64)
    for call in range(HEAD_PLANES // 2):
        even, odd = word_planes(seed, replica, step, word_idx, call, domain)
        for j, plane in ((2 * call, even), (2 * call + 1, odd)):
            bits = (plane[..., None] >> lanes) & np.uint64(1)
            x |= bits << np.uint64(31 - j)
    idx = np.broadcast_to(word_idx[..., None], word_idx.shape + (64,))
    lane = np.broadcast_to(lanes.astype(np.uint32), idx.shape)
    tail = philox4x32_10(make_counter(idx, replica, step, domain, TAIL_CALL, lane), seed_key(seed))[..., 0]
    x |= tail.astype(np.uint64) >> np.uint64(HEAD_PLANES)
    return x.astype(np.uint32)


def init_x32(seed, replica, nsites):
    """x32 per natural site index for the initial-configuration stream (domain 1)."""
    ncalls = (nsites + 3) // 4
    out = philox4x32_10(make_counter(np.arange(ncalls), replica, 0, DOMAIN_INIT), seed_key(seed))
    return out.reshape(-1)[:nsites]


def rx_x32(seed, rnd, pair_idx):
    """x32 for exchange round ``rnd`` and temperature pair (t, t+1) with t = pair_idx (domain 2)."""
    out = philox4x32_10(make_counter(np.asarray(pair_idx), 0, rnd, DOMAIN_RX), seed_key(seed))
    return out[..., 0]


def x32_to_uniform(x32):
    """u = x32 * 2**-32, exact in float64; what the reference's ``random()`` is made to return."""
    return np.asarray(x32, dtype=np.float64) * (1.0 / 4294967296.0)
