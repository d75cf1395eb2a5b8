"""Row sharding of one frame across the GPUs of a box — host-side index logic only.

Pixels are independent in the reference (RE/Scene.h:321-325: each iteration reads the immutable scene and
writes finalImage[idx] only), so the frame shards with no data-path collective.  Two layouts:

  * contiguous bands   rank g renders rows [g*H/G, (g+1)*H/G)                 (rte_render_opts.row0/row1)
  * interleaved stripes rank g renders rows g*S + k*G*S + [0,S), k = 0,1,...  (stripe_rows=S, stripe_period=G*S)

Stripes balance the load (sky / ceiling rows are cheap, sphere-dense rows are not) and are what bench.py
uses; the RGB8 stripes of all ranks are all-gathered (NCCL) and de-interleaved when the PPM is written.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np


def band_of(rank: int, world: int, height: int) -> Tuple[int, int]:
    """Contiguous row band of `rank` (the last rank absorbs nothing: bands differ by at most one row)."""
    return (rank * height) // world, ((rank + 1) * height) // world


def stripe_opts(rank: int, world: int, height: int, stripe: int) -> dict:
    """rte_render_opts fields of `rank` for interleaved stripes of `stripe` rows."""
    if stripe <= 0 or stripe % 4:
        raise ValueError("stripe must be a positive multiple of 4 (the kernel's tile height)")
    return dict(row0=min(rank * stripe, height), row1=height, stripe_rows=stripe, stripe_period=world * stripe)


def stripe_row_list(rank: int, world: int, height: int, stripe: int) -> List[int]:
    """Global row index of every band-local row of `rank`, in buffer order."""
    rows = []
    start = rank * stripe
    while start < height:
        rows.extend(range(start, min(start + stripe, height)))
        start += world * stripe
    return rows


def stripe_row_count(rank: int, world: int, height: int, stripe: int) -> int:
    return len(stripe_row_list(rank, world, height, stripe))


def deinterleave(gathered: np.ndarray, world: int, height: int, width: int, stripe: int, rows_max: int) -> np.ndarray:
    """(world * rows_max * width, C) all-gathered stripes -> (height * width, C) frame in reference order
    (row-major, row 0 = top, RE/Scene.h:214-216)."""
    g = np.asarray(gathered)
    c = g.shape[-1] if g.ndim > 1 else 1
    g = g.reshape(world, rows_max * width, c)
    out = np.empty((height, width, c), dtype=g.dtype)
    for r in range(world):
        rows = stripe_row_list(r, world, height, stripe)
        out[rows] = g[r, :len(rows) * width].reshape(len(rows), width, c)
    return out.reshape(height * width, c)
