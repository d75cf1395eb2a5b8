"""GPU parity of the stand-alone tonemap pass (tonemapAll / tonemap + toColor, RE/RaytracingEngine.cpp:113-214)
through the C ABI.  Bar (north_star): 8-bit pixels within +-1 LSB of the reference; the FP32 pass may differ by one
quantisation step when c*255 lands next to an integer, never more."""
import json
import os

import numpy as np
import pytest

from raytracingengine_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def f4(hdr3):
    h = np.zeros((len(hdr3), 4), dtype=np.float32)
    h[:, :3] = hdr3
    return h


def check(engine, oracle, hdr3, dense=True, **kw):
    h4 = f4(hdr3)
    got = engine.tonemap_host(h4, **kw)
    want = oracle.tonemap_all(h4[:, :3].astype(np.float64))  # the oracle sees the same FP32-representable values
    worst = 0
    for op, a in got.items():
        d = np.abs(a.astype(np.int16) - want[op].astype(np.int16))
        assert d.max() <= 1, f"{_abi.TM_NAMES[op]}: {d.max()} LSB at {np.argwhere(d > 1)[:3].tolist()}"
        worst = max(worst, int(d.max()))
        if dense:  # on ordinary radiance values one-step differences are rare (c*255 next to an integer)
            assert np.count_nonzero(d) <= 0.01 * d.size + 4, f"{_abi.TM_NAMES[op]}: {np.count_nonzero(d)} of {d.size} bytes differ"
    return got, want


def test_kat_vectors(engine, oracle):
    with open(os.path.join(GOLD, "kat.json")) as f:
        kat = json.load(f)["tonemap"]
    hdr = np.array([[np.nan if v is None else v for v in row] for row in kat["hdr"]], dtype=np.float64)
    got = engine.tonemap_host(f4(hdr))
    for op in range(_abi.TM_COUNT):
        d = np.abs(got[op].astype(np.int16) - np.array(kat["rgb8"][op], dtype=np.int16))
        assert d.max() <= 1, _abi.TM_NAMES[op]


@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 40000])
def test_random_hdr_all_operators(engine, oracle, n):
    rng = np.random.default_rng(n)
    hdr = np.exp(rng.normal(-1.0, 2.0, (n, 3)))  # log-normal radiance, 1e-4 .. 1e3
    check(engine, oracle, hdr)


def test_edge_values(engine, oracle):
    """0 (0/0 in the luminance operators), negatives, huge values, inf and NaN: ClampVec3 maps NaN -> 0."""
    vals = [0.0, -0.0, 1e-30, -0.5, -1.0, -2.0, 1.0, 5.0, 25.0, 1e10, 1e30, np.inf, -np.inf, np.nan]
    hdr = np.array([[a, b, c] for a in vals for b in vals[::3] for c in vals[1::4]], dtype=np.float64)
    check(engine, oracle, hdr, dense=False)  # 1e10 / (1e10 + 1) truncates to 254 in FP64 and is 1.0 in FP32: still 1 LSB


def test_rendered_frame(engine, oracle):
    arrays, cam = scenes.config3(scale=0.1)
    engine.upload(arrays, cam)
    hdr = engine.render_host(outputs=("hdr",))["hdr"]
    got = engine.tonemap_host(hdr)
    want = oracle.tonemap_all(hdr[:, :3].astype(np.float64))
    for op in range(_abi.TM_COUNT):
        assert np.abs(got[op].astype(np.int16) - want[op].astype(np.int16)).max() <= 1


def test_operator_subsets_and_parameters(engine, port_oracle):
    rng = np.random.default_rng(11)
    hdr = np.exp(rng.normal(0.0, 1.5, (4099, 3)))
    h4 = f4(hdr)
    full = engine.tonemap_host(h4)
    for ops in ([6], [0, 3], [1, 2, 4, 5]):
        part = engine.tonemap_host(h4, ops=ops)
        assert sorted(part) == sorted(ops)
        for op in ops:
            assert np.array_equal(part[op], full[op])
    # reinhardExtended*(.., max_white) and reinhardJodie(.., a) with non-default parameters
    got = engine.tonemap_host(h4, ops=[2, 3, 4], max_white=2.5, jodie_a=0.4)
    for op in (2, 3, 4):
        want = port_oracle.tonemap_one(op, h4[:, :3].astype(np.float64), max_white=2.5, jodie_a=0.4)
        assert np.abs(got[op].astype(np.int16) - want.astype(np.int16)).max() <= 1


def test_max_luminance(engine):
    rng = np.random.default_rng(5)
    hdr = f4(np.exp(rng.normal(0.0, 2.0, (100003, 3))))
    hdr[17] = [np.nan, 1.0, 1.0, 0.0]  # NaN luminance is ignored
    dev = engine.malloc(hdr.nbytes)
    engine.h2d(dev, hdr)
    got = engine.max_luminance(dev, len(hdr))
    lum = hdr[:, 0].astype(np.float64) * 0.2126 + hdr[:, 1] * 0.7152 + hdr[:, 2] * 0.0722
    assert abs(got - np.nanmax(lum)) <= 1e-5 * np.nanmax(lum)
    assert engine.max_luminance(dev, 0) == 0.0


def test_auto_white_device_path(engine, port_oracle):
    """rte_max_luminance_device + rte_tonemap_auto_white (the stream-ordered auto max-white pair of the multi-GPU
    pipeline) = rte_tonemap with the same max_white passed from the host."""
    rng = np.random.default_rng(11)
    hdr = f4(np.exp(rng.normal(0.0, 1.5, (40001, 3))))
    n = len(hdr)
    dev = engine.malloc(hdr.nbytes)
    white = engine.malloc(4)
    outs = [engine.malloc(3 * n + 16) for _ in range(2)]
    engine.h2d(dev, hdr)
    engine.max_luminance_device(dev, n, white)
    w = np.zeros(1, np.float32)
    engine.d2h(w, white)
    assert w[0] == np.float32(engine.max_luminance(dev, n))
    for op in (2, 3):
        ptrs = [None] * _abi.TM_COUNT
        ptrs[op] = outs[0]
        engine.tonemap_auto_white(dev, n, ptrs, white, ops=[op])
        ptrs[op] = outs[1]
        engine.tonemap_device(dev, n, ptrs, ops=[op], max_white=float(w[0]))
        a, b = np.zeros((n, 3), np.uint8), np.zeros((n, 3), np.uint8)
        engine.d2h(a, outs[0]); engine.d2h(b, outs[1])
        assert np.array_equal(a, b)
        want = port_oracle.tonemap_one(op, hdr[:, :3].astype(np.float64), max_white=float(w[0]), jodie_a=0.18)
        assert np.abs(a.astype(np.int16) - want.astype(np.int16)).max() <= 1


def test_exact_host_tonemap_is_byte_exact(engine, oracle):
    """rte_tonemap_exact_host (what tonemap() / tonemapAll() of the host mirrors call): the FP64 operators over the FP64 HDR —
    the reference's own bytes, on ordinary radiance, on the KAT vectors and on the edge values (0/0, negatives, inf, NaN)."""
    with open(os.path.join(GOLD, "kat.json")) as f:
        kat = json.load(f)["tonemap"]
    rng = np.random.default_rng(3)
    vals = [0.0, -0.0, 1e-30, -0.5, -1.0, 1.0, 5.0, 25.0, 1e10, 1e30, np.inf, -np.inf, np.nan]
    hdr = np.concatenate([np.exp(rng.normal(-1.0, 2.0, (50001, 3))),
                          np.array([[np.nan if v is None else v for v in row] for row in kat["hdr"]], dtype=np.float64),
                          np.array([[a, b, c] for a in vals for b in vals[::3] for c in vals[1::4]], dtype=np.float64)])
    got = engine.tonemap_exact_host(hdr)
    want = oracle.tonemap_all(hdr)
    for op in range(_abi.TM_COUNT):
        assert np.array_equal(got[op], want[op]), f"{_abi.TM_NAMES[op]}: {np.count_nonzero(got[op] != want[op])} bytes differ"
    part = engine.tonemap_exact_host(hdr, ops=[_abi.TM_ACES])
    assert sorted(part) == [_abi.TM_ACES] and np.array_equal(part[_abi.TM_ACES], want[_abi.TM_ACES])
    # the operator functions themselves: Vec3 -> Vec3 in FP64
    _, fl = oracle.tonemap_all(hdr[:50001], want_float=True)
    for op in range(_abi.TM_COUNT):
        v = engine.tonemap_values_host(op, hdr[:50001], max_white=5.0, jodie_a=0.18)
        np.testing.assert_allclose(v, fl[op], rtol=1e-13, atol=1e-300, equal_nan=True)
