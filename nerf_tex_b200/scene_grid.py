"""Uniform grids over the instancer's scene (host side of CommitScene, instancer/src/instancer.cpp:420-424).

The reference hands its patch instances and meshes to Embree, which builds BVHs.  Here the device kernels walk a
uniform grid (csrc/ntx_instancer.cu::grid_walk): patch instances sit on a surface at roughly one size, which is the
case a grid serves as well as a hierarchy, and a grid is two flat arrays that travel through the C ABI as pointers.
Built once per scene with vectorised NumPy; `cell_start[ncell + 1]` / `items[]` list the items whose box, inflated by
`pad`, overlaps a cell (cell index = (z * ny + y) * nx + x).  The inflation is what makes the walk conservative: its
floating-point stepping may clip a cell corner by a few ulps, and an item the ray really hits always reaches at least
`pad` into one of the cells the ray crosses."""

from __future__ import annotations

from typing import Dict

import numpy as np

MAX_RES = 256
MAX_CELLS = 1 << 22


def build_grid(box_min: np.ndarray, box_max: np.ndarray, items_per_cell: float = 2.0) -> Dict[str, np.ndarray]:
    """box_min / box_max [n,3]: world boxes of the items.  Returns lo[3], cell[3], res[3], cell_start, items."""
    box_min, box_max = np.asarray(box_min, np.float64), np.asarray(box_max, np.float64)
    n = box_min.shape[0]
    assert n > 0 and box_min.shape == box_max.shape == (n, 3)
    lo, hi = box_min.min(0), box_max.max(0)
    extent = np.maximum(hi - lo, 1e-6)
    pad = 1e-4 * float(extent.max())
    lo, hi = lo - 2 * pad, hi + 2 * pad
    extent = hi - lo
    # cell size: about the mean item size per axis (an item then overlaps ~8 cells), coarsened until the grid fits
    # (rounded down: an axis along which the scene is about one item thick — patches on a surface — gets a single layer
    # instead of two layers that both list everything)
    size = np.maximum((box_max - box_min).mean(0), extent / MAX_RES)
    res = np.clip(np.floor(extent / size), 1, MAX_RES).astype(np.int64)
    while res.prod() > MAX_CELLS or res.prod() > max(64, int(8 * n / items_per_cell)):
        a = int(np.argmax(res))
        if res[a] == 1:
            break
        res[a] = (res[a] + 1) // 2
    cell = extent / res
    c0 = np.clip(np.floor((box_min - pad - lo) / cell), 0, res - 1).astype(np.int64)
    c1 = np.clip(np.floor((box_max + pad - lo) / cell), 0, res - 1).astype(np.int64)
    span = c1 - c0 + 1                                                   # [n,3] cells per axis
    per_item = span.prod(1)
    total = int(per_item.sum())
    owner = np.repeat(np.arange(n), per_item)
    k = np.arange(total) - np.repeat(np.cumsum(per_item) - per_item, per_item)      # index within the item's block
    sx, sy = span[owner, 0], span[owner, 1]
    cx = c0[owner, 0] + k % sx
    cy = c0[owner, 1] + (k // sx) % sy
    cz = c0[owner, 2] + k // (sx * sy)
    cell_id = (cz * res[1] + cy) * res[0] + cx
    order = np.lexsort((owner, cell_id))                                 # items of a cell in ascending id
    ncell = int(res.prod())
    counts = np.bincount(cell_id, minlength=ncell)
    cell_start = np.zeros(ncell + 1, np.int64)
    cell_start[1:] = np.cumsum(counts)
    assert cell_start[-1] < 2 ** 31
    return {"lo": lo.astype(np.float32), "cell": cell.astype(np.float32), "res": res.astype(np.int32),
            "cell_start": cell_start.astype(np.int32), "items": owner[order].astype(np.int32), "pad": pad}


def instance_boxes(transformations: np.ndarray, b_0, b_1):
    """World-space bounding boxes of the instanced patch boxes (corners of [b_0, b_1] through every 4x4 transform)."""
    T = np.asarray(transformations, np.float64).reshape(-1, 4, 4)
    b0, b1 = np.asarray(b_0, np.float64), np.asarray(b_1, np.float64)
    corners = np.array([[(b0, b1)[(c >> a) & 1][a] for a in range(3)] + [1.0] for c in range(8)])     # [8,4]
    world = np.einsum("iab,cb->ica", T[:, :3, :], corners)                                             # [I,8,3]
    return world.min(1), world.max(1)


def triangle_boxes(V: np.ndarray, F: np.ndarray):
    P = np.asarray(V, np.float64)[np.asarray(F, np.int64)]                                             # [m,3,3]
    return P.min(1), P.max(1)
