"""Philox4x32-10 known-answer vectors and the event-stream convention (no GPU needed).

The generator is published (Salmon, Moraes, Dror, Shaw, SC'11); the vectors below are the Random123 distribution's
`kat_vectors` entries for philox4x32 with 10 rounds. oracle/philox_ref.h restates the algorithm; the device copy in
covid-abm_b200/csrc/cabm_device.cuh is written independently and pinned to this one by the GPU parity tests
(every draw of every event must agree for the agent state to be bit-identical).
"""
import ctypes as C

import numpy as np

import oracle_lib

KAT = [
    ((0x00000000, 0x00000000, 0x00000000, 0x00000000), (0x00000000, 0x00000000), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
    ((0xffffffff, 0xffffffff, 0xffffffff, 0xffffffff), (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
    ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
]


def _lib():
    lib = oracle_lib.oracle_backend().lib
    lib.cabm_oracle_philox.argtypes = [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
    lib.cabm_oracle_philox.restype = None
    lib.cabm_oracle_stream_words.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.c_size_t]
    lib.cabm_oracle_stream_words.restype = None
    return lib


def _philox(lib, ctr, key):
    out = (C.c_uint32 * 4)()
    lib.cabm_oracle_philox((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
    return tuple(out)


def _numpy_philox(ctr, key):
    """Independent numpy restatement (uint64 arithmetic), used to cross-check the stream layout."""
    c = [np.uint64(x) for x in ctr]; k = [np.uint64(x) for x in key]
    M0, M1, W0, W1, MASK = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0x9E3779B9), np.uint64(0xBB67AE85), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0 = M0 * c[0]; p1 = M1 * c[2]
        c = [(p1 >> np.uint64(32)) ^ c[1] ^ k[0], p1 & MASK, (p0 >> np.uint64(32)) ^ c[3] ^ k[1], p0 & MASK]
        k = [(k[0] + W0) & MASK, (k[1] + W1) & MASK]
    return tuple(int(x) for x in c)


def test_published_known_answer_vectors():
    lib = _lib()
    for ctr, key, want in KAT:
        assert _philox(lib, ctr, key) == want
        assert _numpy_philox(ctr, key) == want


def test_event_stream_layout():
    """counter = (block, a, b, day << 8 | purpose), key = (seed_lo, seed_hi); draws are consumed word by word."""
    lib = _lib()
    seed, day, purpose, a, b = 0x1234567890ABCDEF, 173, 6, 2_700_000, 991_234
    out = (C.c_uint32 * 11)()
    lib.cabm_oracle_stream_words(seed, day, purpose, a, b, out, 11)
    want = []
    for block in range(3):
        want += list(_numpy_philox((block, a, b, (day << 8) | purpose), (seed & 0xFFFFFFFF, seed >> 32)))
    assert list(out) == want[:11]
    # a different purpose, entity or day gives an unrelated stream
    for alt in [(seed, day, purpose + 1, a, b), (seed, day + 1, purpose, a, b), (seed, day, purpose, a + 1, b), (seed + 1, day, purpose, a, b)]:
        o2 = (C.c_uint32 * 4)()
        lib.cabm_oracle_stream_words(*alt, o2, 4)
        assert list(o2) != want[:4]


def test_uniformity_of_first_words():
    """Cheap sanity on the stream the exposure trials use: first word over consecutive person ids is uniform."""
    lib = _lib()
    n = 20000
    xs = np.empty(n, dtype=np.uint32)
    buf = (C.c_uint32 * 1)()
    for i in range(n):
        lib.cabm_oracle_stream_words(42, 10, 6, i, 7, buf, 1)
        xs[i] = buf[0]
    u = xs.astype(np.float64) / 2.0**32
    hist, _ = np.histogram(u, bins=20, range=(0, 1))
    chi2 = float(((hist - n / 20) ** 2 / (n / 20)).sum())
    assert chi2 < 43.8          # chi-square 19 dof, p = 0.001
    assert abs(np.corrcoef(u[:-1], u[1:])[0, 1]) < 0.03
