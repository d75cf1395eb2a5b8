"""rte_max_luminance_noalloc (include/rte.h): the allocation-free variant of rte_max_luminance_device
(no counterpart in the reference).  Its reduction scratch is a file-scope static __device__ array, so the
call itself allocates nothing.  Checked through the C ABI against the existing entry point on two calls in a row (the second
call sees the first call's leftovers in the shared array), at the size limit, and for the rejection beyond it.

This file is named to be collected LAST."""
import numpy as np
import pytest

from raytracingengine_b200 import _abi
from raytracingengine_b200.engine import RteError

pytestmark = pytest.mark.gpu


def _hdr(n, seed, scale=1.0):
    rng = np.random.default_rng(seed)
    h = np.zeros((n, 4), np.float32)
    h[:, :3] = np.exp(rng.normal(0.0, 2.0, (n, 3))) * scale
    return h


def _both(engine, dev_hdr, n, out_a, out_b):
    """(existing entry point, new entry point) on the same device buffer."""
    engine.max_luminance_device(dev_hdr, n, out_a)
    engine.max_luminance_noalloc(dev_hdr, n, out_b)
    engine.synchronize()
    a, b = np.zeros(1, np.float32), np.zeros(1, np.float32)
    engine.d2h(a, out_a)
    engine.d2h(b, out_b)
    return float(a[0]), float(b[0])


def test_noalloc_equals_existing_entry_point_two_calls_in_a_row(engine):
    n1, n2 = 1_000_003, 40_001  # the second call is smaller: stale partials of the first stay behind in the shared array
    h1, h2 = _hdr(n1, 1, scale=50.0), _hdr(n2, 2, scale=0.01)
    h1[17] = [np.nan, 1.0, 1.0, 0.0]  # NaN luminance is ignored by both
    dev = engine.malloc(h1.nbytes)
    out_a, out_b = engine.malloc(4), engine.malloc(4)
    for rep, (h, n) in enumerate(((h1, n1), (h2, n2), (h1, n1))):
        engine.h2d(dev, h)
        a, b = _both(engine, dev, n, out_a, out_b)
        lum = h[:, 0].astype(np.float64) * 0.2126 + h[:, 1] * 0.7152 + h[:, 2] * 0.0722
        want = float(np.nanmax(lum))
        assert abs(b - a) <= 1e-6 * a, (rep, a, b)          # same value as rte_max_luminance_device
        assert abs(b - want) <= 1e-5 * want, (rep, b, want)  # and the tolerance of tests/test_gpu_tonemap.py::test_max_luminance
    assert h2[:, :3].max() < 0.5 * a  # (the smaller frame really had a smaller maximum than the leftovers)


def test_noalloc_sizes_and_limit(engine):
    out_a, out_b = engine.malloc(4), engine.malloc(4)
    for n in (1, 255, 4096, 4097, 12289):  # one block, block boundary, several blocks with a ragged tail
        h = _hdr(n, 100 + n)
        dev = engine.malloc(h.nbytes)
        engine.h2d(dev, h)
        a, b = _both(engine, dev, n, out_a, out_b)
        assert a == b, (n, a, b)
    # n_pixels == 0 writes 0
    engine.h2d(out_b, np.full(1, 7.0, np.float32))
    engine.max_luminance_noalloc(0, 0, out_b)
    engine.synchronize()
    z = np.ones(1, np.float32)
    engine.d2h(z, out_b)
    assert z[0] == 0.0
    # beyond the stated maximum: a clear error, nothing launched, the context stays usable
    dev = engine.malloc(64)
    with pytest.raises(RteError, match="RTE_NOALLOC_MAX_PIXELS") as ei:
        engine.max_luminance_noalloc(dev, _abi.NOALLOC_MAX_PIXELS + 1, out_b)
    assert ei.value.status == _abi.RTE_ERR_LIMIT
    h = _hdr(5000, 9)
    dev = engine.malloc(h.nbytes)
    engine.h2d(dev, h)
    a, b = _both(engine, dev, 5000, out_a, out_b)
    assert a == b


def test_noalloc_at_a_frame_size(engine):
    """A 3840x2160 frame (the bench's size): 2025 partials, value equal to the existing entry point's."""
    n = 3840 * 2160
    h = _hdr(n, 3)
    h[n - 1, :3] = 1.0e7  # the maximum sits in the last pixel (far above the tail of the log-normal samples)
    dev = engine.malloc(h.nbytes)
    out_a, out_b = engine.malloc(4), engine.malloc(4)
    engine.h2d(dev, h)
    a, b = _both(engine, dev, n, out_a, out_b)
    assert a == b and abs(a - 1.0e7) <= 1e-5 * 1.0e7
