"""Synthetic structured P1 meshes (numpy, host side) — SURVEY.md §8(d).

cube_kuhn(N): (N+1)^3 lattice nodes, lexicographic ids i + (N+1)(j + (N+1)k); every hexahedron is split into the 6
Kuhn tetrahedra sharing the main diagonal (corner 000 -> 111), all positively oriented.  square(N): 2 triangles per quad,
counter-clockwise.  Optional interior jitter makes the element matrices generic; optional permutations expose
unstructured access.  These generators stand in for Arcane's mesh reader + cellNode()/VariableNodeReal3 views.
"""
import itertools

import numpy as np


def kuhn_table():
    """6 tets of the unit cube as corner indices (corner id = dx + 2*dy + 4*dz), positively oriented."""
    tets = []
    for perm in itertools.permutations(range(3)):
        c = [0, 0, 0]
        path = [0]
        for ax in perm:
            c[ax] = 1
            path.append(c[0] + 2 * c[1] + 4 * c[2])
        # orientation = sign of the permutation; swap two vertices of odd ones
        inv = sum(1 for a in range(3) for b in range(a + 1, 3) if perm[a] > perm[b])
        if inv % 2 == 1:
            path[1], path[2] = path[2], path[1]
        tets.append(path)
    return np.array(tets, dtype=np.int64)


def cube_kuhn(N, L=1.0, perturb=0.25, seed=1234, shuffle_cells=None, shuffle_nodes=None):
    """Return (cell_node int32 [6N^3,4], coords float64 [(N+1)^3,3])."""
    n1 = N + 1
    ax = np.arange(n1, dtype=np.float64) * (L / N)
    z, y, x = np.meshgrid(ax, ax, ax, indexing="ij")
    coords = np.stack([x.ravel(), y.ravel(), z.ravel()], axis=1)
    if perturb and N > 1:
        rng = np.random.default_rng(seed)
        h = L / N
        idx = np.arange(n1)
        interior = ((idx[:, None, None] > 0) & (idx[:, None, None] < N) & (idx[None, :, None] > 0) & (idx[None, :, None] < N)
                    & (idx[None, None, :] > 0) & (idx[None, None, :] < N)).ravel()
        jit = rng.uniform(-perturb * h, perturb * h, size=coords.shape)
        coords[interior] += jit[interior]
    hi = np.arange(N, dtype=np.int64)
    kk, jj, ii = np.meshgrid(hi, hi, hi, indexing="ij")
    base = (ii + n1 * (jj + n1 * kk)).ravel()
    corner = np.array([dx + n1 * (dy + n1 * dz) for dz in (0, 1) for dy in (0, 1) for dx in (0, 1)], dtype=np.int64)
    # corner id = dx + 2*dy + 4*dz
    off = corner[kuhn_table()]  # [6,4]
    cells = (base[:, None, None] + off[None, :, :]).reshape(-1, 4).astype(np.int32)
    return _shuffle(cells, coords, shuffle_cells, shuffle_nodes)


def square(N, L=1.0, perturb=0.25, seed=1234, shuffle_cells=None, shuffle_nodes=None):
    """Return (cell_node int32 [2N^2,3], coords float64 [(N+1)^2,3]) with z = 0."""
    n1 = N + 1
    ax = np.arange(n1, dtype=np.float64) * (L / N)
    y, x = np.meshgrid(ax, ax, indexing="ij")
    coords = np.stack([x.ravel(), y.ravel(), np.zeros(n1 * n1)], axis=1)
    if perturb and N > 1:
        rng = np.random.default_rng(seed)
        h = L / N
        idx = np.arange(n1)
        interior = ((idx[:, None] > 0) & (idx[:, None] < N) & (idx[None, :] > 0) & (idx[None, :] < N)).ravel()
        jit = rng.uniform(-perturb * h, perturb * h, size=(n1 * n1, 2))
        coords[interior, :2] += jit[interior]
    hi = np.arange(N, dtype=np.int64)
    jj, ii = np.meshgrid(hi, hi, indexing="ij")
    base = (ii + n1 * jj).ravel()
    off = np.array([[0, 1, n1 + 1], [0, n1 + 1, n1]], dtype=np.int64)
    cells = (base[:, None, None] + off[None, :, :]).reshape(-1, 3).astype(np.int32)
    return _shuffle(cells, coords, shuffle_cells, shuffle_nodes)


def _shuffle(cells, coords, shuffle_cells, shuffle_nodes):
    if shuffle_nodes is not None:
        rng = np.random.default_rng(shuffle_nodes)
        perm = rng.permutation(coords.shape[0])  # new id of old node i is perm[i]
        new_coords = np.empty_like(coords)
        new_coords[perm] = coords
        coords = new_coords
        cells = perm[cells].astype(np.int32)
    if shuffle_cells is not None:
        rng = np.random.default_rng(shuffle_cells)
        cells = cells[rng.permutation(cells.shape[0])]
    return np.ascontiguousarray(cells), np.ascontiguousarray(coords)


def boundary_nodes(coords, axis, value, tol=1e-12):
    """Node ids lying on the plane coords[:, axis] == value (e.g. the "horizontal" group z = 0 / y = 0)."""
    return np.nonzero(np.abs(coords[:, axis] - value) <= tol)[0].astype(np.int32)


def boundary_faces(cells, coords, axis, value, tol=1e-12):
    """Boundary faces (edges in 2D, triangles in 3D) whose nodes all lie on the plane; node order as in the cell."""
    nv = cells.shape[1]
    on = np.abs(coords[:, axis] - value) <= tol
    faces = []
    if nv == 3:
        local = [(0, 1), (1, 2), (2, 0)]
    else:
        local = [(0, 2, 1), (0, 1, 3), (1, 2, 3), (0, 3, 2)]  # outward for positively oriented tets
    for lf in local:
        f = cells[:, lf]
        m = on[f].all(axis=1)
        faces.append(f[m])
    return np.ascontiguousarray(np.concatenate(faces, axis=0).astype(np.int32))


def kuhn_counts(N):
    """(nb_cell, nb_node, nb_edge, nnzb) of cube_kuhn(N) — SURVEY.md §8 table."""
    nn = (N + 1) ** 3
    nc = 6 * N ** 3
    e = 3 * N * (N + 1) ** 2 + 3 * N * N * (N + 1) + N ** 3
    return nc, nn, e, 2 * e + nn


def kuhn_pattern(N):
    """Closed-form BSR pattern of cube_kuhn(N) in the canonical order (diagonal first, then ascending neighbour ids):
    (row_index int32 [Nn+1], columns int32 [nnzb]).  Two lattice nodes share an edge of the Kuhn triangulation iff their
    index difference is a non-zero monotone 0/1 vector or its negative (7 + 7 offsets; interior rows have 15 entries,
    SURVEY.md §8).  Independent of the edge-key algorithm (femutils/BSRFormat.h:602-744), so it pins the pattern at sizes
    where a sequential sort-based restatement takes minutes."""
    n1 = N + 1
    nn = n1 ** 3
    pos = [(dx, dy, dz) for dz in (0, 1) for dy in (0, 1) for dx in (0, 1) if (dx, dy, dz) != (0, 0, 0)]
    offs = sorted(pos + [(-a, -b, -c) for a, b, c in pos], key=lambda d: d[0] + n1 * (d[1] + n1 * d[2]))
    idx = np.arange(n1, dtype=np.int32)
    k, j, i = np.meshgrid(idx, idx, idx, indexing="ij")
    i, j, k = i.ravel(), j.ravel(), k.ravel()
    node = np.arange(nn, dtype=np.int32)
    cand = np.empty((nn, 15), dtype=np.int32)
    keep = np.empty((nn, 15), dtype=bool)
    cand[:, 0] = node
    keep[:, 0] = True
    for q, (dx, dy, dz) in enumerate(offs, start=1):
        cand[:, q] = node + np.int32(dx + n1 * (dy + n1 * dz))
        keep[:, q] = (i + dx >= 0) & (i + dx <= N) & (j + dy >= 0) & (j + dy <= N) & (k + dz >= 0) & (k + dz <= N)
    row_index = np.zeros(nn + 1, dtype=np.int64)
    np.cumsum(keep.sum(axis=1), out=row_index[1:])
    return row_index.astype(np.int32), np.ascontiguousarray(cand[keep])


def square_pattern(N):
    """Closed-form BSR pattern of square(N), canonical order: neighbours at +-1, +-(N+1), +-(N+2) (the quad diagonal),
    7 entries per interior row."""
    n1 = N + 1
    nn = n1 * n1
    pos = [(1, 0), (0, 1), (1, 1)]
    offs = sorted(pos + [(-a, -b) for a, b in pos], key=lambda d: d[0] + n1 * d[1])
    idx = np.arange(n1, dtype=np.int32)
    j, i = np.meshgrid(idx, idx, indexing="ij")
    i, j = i.ravel(), j.ravel()
    node = np.arange(nn, dtype=np.int32)
    cand = np.empty((nn, 7), dtype=np.int32)
    keep = np.empty((nn, 7), dtype=bool)
    cand[:, 0] = node
    keep[:, 0] = True
    for q, (dx, dy) in enumerate(offs, start=1):
        cand[:, q] = node + np.int32(dx + n1 * dy)
        keep[:, q] = (i + dx >= 0) & (i + dx <= N) & (j + dy >= 0) & (j + dy <= N)
    row_index = np.zeros(nn + 1, dtype=np.int64)
    np.cumsum(keep.sum(axis=1), out=row_index[1:])
    return row_index.astype(np.int32), np.ascontiguousarray(cand[keep])
