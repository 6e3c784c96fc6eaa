"""
Pins of the CPU oracle (oracle/) that need no GPU: known answers, an independent numpy
restatement of the reference's Philox mapping, accuracy of the deterministic math functions,
and the reference's own statistical RNG test.
"""

import numpy as np
import pytest

M0, M1 = 0xD2511F53, 0xCD9E8D57
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = 0xFFFFFFFF


def philoxNumpy(ctr: np.ndarray, key: np.ndarray) -> np.ndarray:
    """Philox4x32-10 (Salmon et al., SC'11) on uint64 lanes; ctr (n,4), key (n,2)."""
    x = [ctr[:, k].astype(np.uint64) for k in range(4)]
    k0, k1 = key[:, 0].astype(np.uint64), key[:, 1].astype(np.uint64)
    for _ in range(10):
        p0 = np.uint64(M0) * x[0]
        p1 = np.uint64(M1) * x[2]
        x = [(p1 >> np.uint64(32)) ^ x[1] ^ k0, p1 & np.uint64(MASK), (p0 >> np.uint64(32)) ^ x[3] ^ k1,
             p0 & np.uint64(MASK)]
        k0 = (k0 + np.uint64(W0)) & np.uint64(MASK)
        k1 = (k1 + np.uint64(W1)) & np.uint64(MASK)
    return np.stack(x, axis=1).astype(np.uint32)


def theiaDraws(key: int, offset: int, stream: int, first: int, n: int) -> np.ndarray:
    """
    sh/random.philox.glsl:15-77 of the reference restated with Python integers: draw i of a
    stream = Philox(ctr = offset + 4 i (128 bit, end-around carry), key + stream (64 bit,
    end-around carry))[i & 3], mapped to [0,1) by sh/random.util.glsl:8-14.
    """
    ctrs, keys = [], []
    for i in range(first, first + n):
        c = offset + ((i << 2) & MASK)
        if c >> 128:
            c = (c & ((1 << 128) - 1)) + 1
        k = key + stream
        if k >> 64:
            k = (k & ((1 << 64) - 1)) + 1
        ctrs.append([(c >> (32 * j)) & MASK for j in range(4)])
        keys.append([k & MASK, (k >> 32) & MASK])
    raw = philoxNumpy(np.array(ctrs, np.uint64), np.array(keys, np.uint64))
    pick = raw[np.arange(n), (np.arange(first, first + n) & 3)]
    f = pick.astype(np.float32) * np.float32(2.0 ** -32)  # uint -> float conversion rounds to nearest
    return np.minimum(f, np.uint32(0x3F7FFFFF).view(np.float32))


def test_philox_random123_known_answers(orc):
    ctr = np.array([[0, 0, 0, 0], [MASK] * 4, [0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344]], np.uint32)
    key = np.array([[0, 0], [MASK] * 2, [0xA4093822, 0x299F31D0]], np.uint32)
    want = np.array([[0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8], [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD],
                     [0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1]], np.uint32)
    assert np.array_equal(philoxNumpy(ctr, key), want)  # the independent restatement itself
    assert np.array_equal(orc.philoxRaw(ctr, key), want)


def test_philox_block_function_matches_numpy(orc):
    rng = np.random.default_rng(1)
    ctr = rng.integers(0, 1 << 32, (4096, 4), dtype=np.uint64).astype(np.uint32)
    key = rng.integers(0, 1 << 32, (4096, 2), dtype=np.uint64).astype(np.uint32)
    assert np.array_equal(orc.philoxRaw(ctr, key), philoxNumpy(ctr, key))


@pytest.mark.parametrize("key,offset,stream,first", [
    (0xC0FFEE, 0, 0, 0),
    (42, (1 << 128) - 5, 7, 0),             # counter wraps: end-around carry
    ((1 << 64) - 1, (1 << 32) - 2, 3, 5),   # key wraps, carry out of counter word 0
    (0xC01DC0FFEE, 1 << 96, 123456, 1000),
])
def test_stream_mapping_matches_independent_restatement(orc, key, offset, stream, first):
    n = 64
    got = orc.philoxFill(key, offset, 1, n, baseStream=stream, baseCount=first)[0]
    want = theiaDraws(key, offset, stream, first, n)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_rng_sink_layout(orc):
    # sh/random.sink.glsl:17-39: u[stream * samples + count]; the 2D variant stores
    # random2D_s(stream, count) = (draw count, draw count + 1), so neighbouring samples overlap
    a = orc.philoxFill(5, 0, 8, 16)
    for s in (0, 3, 7):
        assert np.array_equal(a[s], theiaDraws(5, 0, s, 0, 16))
    b = orc.philoxFill(5, 0, 4, 8, dim=2)
    d = theiaDraws(5, 0, 2, 0, 9)
    assert np.array_equal(b[2][:, 0], d[:8]) and np.array_equal(b[2][:, 1], d[1:])


def test_rng_sink_bases(orc):
    # tests/test_random.py:36-64 of the reference (sizes and bases of test_rngSink / test_rngSink_2D): element [s, c] is
    # draw baseCount + c of stream baseStream + s. The reference checks this with a debug generator that returns
    # 10000 stream + i; generators are kinds here, so the pin is the Philox stream mapping restated with Python integers.
    a = orc.philoxFill(0xC0FFEE, 77, 250, 800, baseStream=12, baseCount=316)
    assert a.shape == (250, 800)
    for s in (0, 17, 249):
        assert np.array_equal(a[s], theiaDraws(0xC0FFEE, 77, 12 + s, 316, 800))
    b = orc.philoxFill(0xC0FFEE, 77, 250, 800, baseStream=12, baseCount=316, dim=2)
    assert b.shape == (250, 800, 2)
    for s in (0, 249):
        d = theiaDraws(0xC0FFEE, 77, 12 + s, 316, 801)
        assert np.array_equal(b[s][:, 0], d[:800]) and np.array_equal(b[s][:, 1], d[1:])


def test_counter_and_key_words_round_trip():
    # tests/test_random.py:125-127 (uvec4 <-> int): the 128 bit offset and the 64 bit key as little endian words
    from theia_b200.random import Counter, Key
    c = Counter()
    c.value = 0xC0110FFC0FFEE
    assert c.value == 0xC0110FFC0FFEE and list(c.word) == [0xFFC0FFEE, 0xC0110, 0, 0]
    c.value = (1 << 128) - 3
    assert c.value == (1 << 128) - 3 and list(c.word) == [0xFFFFFFFD, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF]
    k = Key()
    k.value = 0x0123456789ABCDEF
    assert (k.lo, k.hi, k.value) == (0x89ABCDEF, 0x01234567, 0x0123456789ABCDEF)


def test_uniformity_reference_pin(orc):
    # tests/test_random.py:67-81 of the reference: 20 bins, max deviation < 0.5 %
    u = orc.philoxFill(0xC0FFEE, 0, 4096, 2048).reshape(-1)
    assert u.min() >= 0.0 and u.max() < 1.0
    hist = np.histogram(u, bins=20, range=(0, 1))[0] / u.size
    assert np.abs(hist * 20 - 1.0).max() < 5e-3
    # autoAdvance semantics: the next batch starts where offset + nRNGSamples points
    a = orc.philoxFill(1, 0, 2, 32)
    b = orc.philoxFill(1, 4 * 8, 2, 24)  # offset counts Philox blocks of 4: draw i sits at offset + 4 i
    assert np.array_equal(a[:, 8:], b)


@pytest.mark.parametrize("fn,lo,hi,ulp", [("sin", -20.0, 20.0, 4), ("cos", -20.0, 20.0, 4), ("exp", -80.0, 80.0, 4),
                                          ("log", 1e-30, 1e30, 4)])
def test_deterministic_math_accuracy(orc, fn, lo, hi, ulp):
    rng = np.random.default_rng(3)
    n = 1 << 16
    if fn == "log":
        x = np.exp(rng.uniform(np.log(lo), np.log(hi), n)).astype(np.float32)
    else:
        x = rng.uniform(lo, hi, n).astype(np.float32)
    got = orc.mathEval(fn, x).astype(np.float64)
    want = getattr(np, fn)(x.astype(np.float64))
    want32 = want.astype(np.float32)
    spacing = np.maximum(np.spacing(np.abs(want32)).astype(np.float64), 2.0 ** -149)
    if fn in ("sin", "cos"):
        # absolute accuracy near the zeros, relative elsewhere (the Vulkan bound is 2^-11 absolute)
        err = np.abs(got - want) / np.maximum(spacing, 2.0 ** -24)
    else:
        err = np.abs(got - want) / spacing
    assert err.max() <= ulp, err.max()


def test_deterministic_pow_accuracy(orc):
    rng = np.random.default_rng(4)
    n = 1 << 16
    x = np.exp(rng.uniform(-8.0, 8.0, n)).astype(np.float32)
    y = rng.uniform(-6.0, 6.0, n).astype(np.float32)
    got = orc.mathEval("pow", x, y).astype(np.float64)
    want = np.power(x.astype(np.float64), y.astype(np.float64))
    rel = np.abs(got - want) / want
    assert rel.max() < 1e-6


def test_deterministic_tanh_atan_accuracy(orc):
    # the kernel CDF of sh/response.histogram.kernel.glsl:62-70 is built from these two
    rng = np.random.default_rng(5)
    x = rng.uniform(-30.0, 30.0, 1 << 16).astype(np.float32)
    assert np.abs(orc.mathEval("tanh", x).astype(np.float64) - np.tanh(x.astype(np.float64))).max() < 2.5e-7
    t = np.tan(rng.uniform(-1.57, 1.57, 1 << 16)).astype(np.float32)
    got, want = orc.mathEval("atan", t).astype(np.float64), np.arctan(t.astype(np.float64))
    assert (np.abs(got - want) / np.maximum(np.abs(want), 1e-30)).max() < 4e-7
    assert orc.mathEval("tanh", np.array([0.0, 50.0, -50.0], np.float32)).tolist() == [0.0, 1.0, -1.0]
    assert orc.mathEval("atan", np.array([0.0], np.float32))[0] == 0.0
    assert orc.mathEval("atan", np.array([np.inf], np.float32))[0] == np.float32(np.pi / 2)


def test_math_special_values(orc):
    assert orc.mathEval("exp", np.array([0.0], np.float32))[0] == 1.0
    assert orc.mathEval("log", np.array([1.0], np.float32))[0] == 0.0
    assert orc.mathEval("sin", np.array([0.0], np.float32))[0] == 0.0
    assert orc.mathEval("cos", np.array([0.0], np.float32))[0] == 1.0
    assert orc.mathEval("exp", np.array([-200.0], np.float32))[0] == 0.0
    assert np.isinf(orc.mathEval("exp", np.array([200.0], np.float32))[0])
    assert orc.mathEval("log", np.array([0.0], np.float32))[0] == -np.inf
