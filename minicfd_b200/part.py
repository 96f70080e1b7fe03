"""Host-side mirror of `mini::mesh::part::Part` (include/mini/mesh/part.hpp) for in-memory meshes.

Only the bookkeeping of the hot path is mirrored: which cells are local / inner / inter / ghost, their
ids, the faces, and the halo lists.  Numbering follows the reference (citations inline); what the
reference leaves implementation-defined (ghost ids iterate a std::unordered_map, part.hpp:768-775) is
canonicalised to ascending global id.  CGNS / partition-file I/O is out of scope.
"""
from __future__ import annotations

import numpy as np

from .mesh import HEX_FACE_NODES, HexMesh


def dual_graph(mesh: HexMesh):
    """Face-adjacent cell pairs (i < j): the dual graph METIS_MeshToDual builds with n_common >= 3
    (include/mini/mesh/shuffler.hpp:526-527), incl. periodic identifications."""
    cells = mesh.cells
    n = len(cells)
    quads = cells[:, HEX_FACE_NODES].reshape(n * 6, 4)
    if mesh.periodic is not None and len(mesh.periodic):
        # alias the nodes of every second quad to those of its partner (only on that face)
        a = mesh.periodic[:, 0, :]
        b = mesh.periodic[:, 1, :]
        keyb = np.sort(b, axis=1)
        lut = {tuple(k): (bb, aa) for k, bb, aa in zip(map(tuple, keyb), b, a)}
        sq = np.sort(quads, axis=1)
        for row in range(len(quads)):
            hit = lut.get(tuple(sq[row]))
            if hit is not None:
                bb, aa = hit
                m = {int(x): int(y) for x, y in zip(bb, aa)}
                quads[row] = [m[int(v)] for v in quads[row]]
    keys = np.sort(quads, axis=1)
    order = np.lexsort(keys.T[::-1])
    ks = keys[order]
    same = (ks[1:] == ks[:-1]).all(axis=1)
    first = order[:-1][same] // 6
    second = order[1:][same] // 6
    i = np.minimum(first, second)
    j = np.maximum(first, second)
    pairs = np.stack([i, j], axis=1)
    pairs = pairs[np.lexsort((pairs[:, 1], pairs[:, 0]))]
    return pairs


class Part:
    """One part of a partitioned mesh, as rank `rank` of `n_parts` sees it."""

    def __init__(self, mesh: HexMesh, cell_parts=None, rank: int = 0, n_parts: int = 1, pairs=None):
        self.mesh = mesh
        self.rank, self.size = rank, n_parts
        n = mesh.n_cells
        parts = np.zeros(n, dtype=np.int32) if cell_parts is None else np.asarray(cell_parts, dtype=np.int32)
        self.cell_parts = parts
        pairs = dual_graph(mesh) if pairs is None else pairs
        pi, pj = parts[pairs[:, 0]], parts[pairs[:, 1]]
        mine_i, mine_j = pi == rank, pj == rank
        cross = pairs[(mine_i ^ mine_j)]
        # orient cross pairs as (local, remote)
        swap = parts[cross[:, 0]] != rank
        cross = np.where(swap[:, None], cross[:, ::-1], cross)
        local = np.flatnonzero(parts == rank)
        is_inter = np.zeros(n, dtype=bool)
        is_inter[cross[:, 0]] = True
        # part.hpp:746-767: inner cells first, then inter cells, each ascending
        inner = local[~is_inter[local]]
        inter = local[is_inter[local]]
        self.local_global = np.concatenate([inner, inter])
        self.n_inner, self.n_local = len(inner), len(local)
        self.ghost_global = np.unique(cross[:, 1])
        self.n_ghost = len(self.ghost_global)
        self.global_ids = np.concatenate([self.local_global, self.ghost_global])
        self.global_to_id = -np.ones(n, dtype=np.int64)
        self.global_to_id[self.global_ids] = np.arange(len(self.global_ids))
        # part.hpp:800-805, 904-925: per neighbour part (ascending), cells in ascending global id
        remote_part = parts[cross[:, 1]]
        self.neighbours = np.unique(remote_part)
        self.send_cells, self.recv_cells = [], []
        for q in self.neighbours:
            sel = remote_part == q
            self.send_cells.append(self.global_to_id[np.unique(cross[sel, 0])])
            self.recv_cells.append(self.global_to_id[np.unique(cross[sel, 1])])
        # boundary faces held by local cells, keyed by the sorted quad
        self.cell_nodes = mesh.cells[self.global_ids]
        if len(mesh.bfaces):
            quads = mesh.cells[:, HEX_FACE_NODES].reshape(n * 6, 4)
            keys = np.sort(quads, axis=1)
            bkeys = np.sort(mesh.bfaces, axis=1)
            allk = np.concatenate([keys, bkeys])
            _, inv = np.unique(allk, axis=0, return_inverse=True)
            inv = inv.reshape(-1)
            owner = -np.ones(inv.max() + 1, dtype=np.int64)
            owner[inv[:n * 6]] = np.arange(n * 6) // 6   # boundary quads have a single owner
            holder = owner[inv[n * 6:]]
            keep = parts[holder] == rank
            self.bfaces = mesh.bfaces[keep]
            self.btags = mesh.btags[keep]
        else:
            self.bfaces = np.zeros((0, 4), dtype=np.int64)
            self.btags = np.zeros((0,), dtype=np.int32)

    # names follow part.hpp:784-797, 1022-1069
    def CountLocalCells(self):
        return self.n_local

    def CountGhostCells(self):
        return self.n_ghost

    def GetCellDataSize(self, kfields):
        return self.n_local * kfields

    def scatter_to_global(self, column, kfields, out):
        """Place this part's column (local id order) into a global-cell-order array."""
        out.reshape(-1, kfields)[self.local_global] = column.reshape(-1, kfields)
        return out

    def gather_from_global(self, global_column, kfields):
        return global_column.reshape(-1, kfields)[self.local_global].reshape(-1).copy()


def lobatto_points(degree: int):
    """include/mini/integrator/lobatto.hpp:38-57"""
    if degree == 2:
        return np.array([-1.0, 0.0, 1.0])
    if degree == 3:
        return np.array([-1.0, -np.sqrt(1.0 / 5), np.sqrt(1.0 / 5), 1.0])
    raise ValueError("degree must be 2 or 3")


def cell_points(nodes, cell_nodes, degree: int):
    """Global coordinates of the solution points, [n_cells, N, 3], node order ijk = (i*Q+j)*Q+k
    (include/mini/integrator/hexahedron.hpp:66-80; trilinear map coordinate/hexahedron.hpp:413-428)."""
    x = lobatto_points(degree)
    ref = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                    [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]], dtype=np.float64)
    I, J, K = np.meshgrid(x, x, x, indexing="ij")
    loc = np.stack([I.ravel(), J.ravel(), K.ravel()], axis=1)            # [N, 3]
    shape = np.prod(1 + ref[None, :, :] * loc[:, None, :], axis=2) / 8   # [N, 8]
    return np.einsum("nv,cvd->cnd", shape, nodes[cell_nodes])
