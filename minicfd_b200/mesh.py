"""Synthetic hexahedral meshes (host side, numpy only).

The reference reads gmsh/CGNS meshes (`input/*.geo`, `mini/mesh/cgns.hpp`); neither gmsh nor CGNS is
available here, so the configurations of BASELINE.json are reproduced as structured boxes of
Hexahedron8 cells in CGNS node order (`mini/coordinate/hexahedron.hpp:378-389`), optionally warped and
with per-cell node rotations so that every holder/sharer face orientation occurs.
"""
from __future__ import annotations

import dataclasses

import numpy as np

# CGNS HEXA_8 local faces -> vertex slots (mini/coordinate/hexahedron.hpp:378-389)
HEX_FACE_NODES = np.array(
    [[0, 3, 2, 1], [0, 1, 5, 4], [1, 2, 6, 5], [2, 3, 7, 6], [0, 4, 7, 3], [4, 5, 6, 7]], dtype=np.int64)

# boundary tags of a box
XMIN, XMAX, YMIN, YMAX, ZMIN, ZMAX = range(6)


@dataclasses.dataclass
class HexMesh:
    nodes: np.ndarray        # [n_nodes, 3] float64
    cells: np.ndarray        # [n_cells, 8] int64, CGNS HEXA_8 order
    bfaces: np.ndarray       # [n_bfaces, 4] int64
    btags: np.ndarray        # [n_bfaces] int32
    dims: tuple | None = None  # (nx, ny, nz) when structured
    periodic: np.ndarray | None = None  # [n_pairs, 2, 4] int64: matched quads (minus side, plus side)

    @property
    def n_cells(self) -> int:
        return int(self.cells.shape[0])


def box(nx: int, ny: int, nz: int, lo=(0.0, 0.0, 0.0), hi=(1.0, 1.0, 1.0), *, warp: float = 0.0,
        seed: int = 0, periodic: bool = False) -> HexMesh:
    """nx*ny*nz Hexahedron8 cells, cell id = ix + nx*(iy + ny*iz).

    warp > 0 displaces interior nodes by warp*h uniformly at random (non-affine cells).
    periodic=True emits no boundary faces but `periodic` face pairs instead (GPU-only variant, the
    reference has no periodic BC).
    """
    xs = np.linspace(lo[0], hi[0], nx + 1)
    ys = np.linspace(lo[1], hi[1], ny + 1)
    zs = np.linspace(lo[2], hi[2], nz + 1)
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    nodes = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)

    def nid(i, j, k):
        return i + (nx + 1) * (j + (ny + 1) * k)

    if warp > 0.0:
        rng = np.random.default_rng(seed)
        h = np.array([(hi[d] - lo[d]) / n for d, n in enumerate((nx, ny, nz))])
        K, J, I = np.meshgrid(np.arange(nz + 1), np.arange(ny + 1), np.arange(nx + 1), indexing="ij")
        interior = ((I > 0) & (I < nx) & (J > 0) & (J < ny) & (K > 0) & (K < nz)).ravel()
        nodes[interior] += warp * h * rng.uniform(-1, 1, size=(int(interior.sum()), 3))

    k, j, i = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    i, j, k = i.ravel(), j.ravel(), k.ravel()
    cells = np.stack([nid(i, j, k), nid(i + 1, j, k), nid(i + 1, j + 1, k), nid(i, j + 1, k),
                      nid(i, j, k + 1), nid(i + 1, j, k + 1), nid(i + 1, j + 1, k + 1), nid(i, j + 1, k + 1)],
                     axis=1).astype(np.int64)

    def cid(ii, jj, kk):
        return ii + nx * (jj + ny * kk)

    bf, bt, pairs = [], [], []
    sides = [
        (XMIN, 4, lambda a, b: cid(0, a, b), (ny, nz)), (XMAX, 2, lambda a, b: cid(nx - 1, a, b), (ny, nz)),
        (YMIN, 1, lambda a, b: cid(a, 0, b), (nx, nz)), (YMAX, 3, lambda a, b: cid(a, ny - 1, b), (nx, nz)),
        (ZMIN, 0, lambda a, b: cid(a, b, 0), (nx, ny)), (ZMAX, 5, lambda a, b: cid(a, b, nz - 1), (nx, ny)),
    ]
    quads = {}
    for tag, lf, cfun, (na, nb) in sides:
        b, a = np.meshgrid(np.arange(nb), np.arange(na), indexing="ij")
        c = cfun(a.ravel(), b.ravel())
        q = cells[c][:, HEX_FACE_NODES[lf]]
        quads[tag] = q
        bf.append(q)
        bt.append(np.full(len(c), tag, dtype=np.int32))
    mesh = HexMesh(nodes=nodes, cells=cells, bfaces=np.concatenate(bf), btags=np.concatenate(bt),
                   dims=(nx, ny, nz))
    if periodic:
        # vertex m of the low-side quad coincides (up to the period) with vertex m of the high-side one
        matched = {XMAX: [1, 5, 6, 2], YMAX: [3, 2, 6, 7], ZMAX: [4, 7, 6, 5]}
        for (lo_tag, hi_tag), (tag, lf, cfun, (na, nb)) in zip(((XMIN, XMAX), (YMIN, YMAX), (ZMIN, ZMAX)),
                                                              (sides[1], sides[3], sides[5])):
            b, a = np.meshgrid(np.arange(nb), np.arange(na), indexing="ij")
            hi_quads = cells[cfun(a.ravel(), b.ravel())][:, matched[hi_tag]]
            pairs.append(np.stack([quads[lo_tag], hi_quads], axis=1))
        mesh.periodic = np.concatenate(pairs).astype(np.int64)
        mesh.bfaces = np.zeros((0, 4), dtype=np.int64)
        mesh.btags = np.zeros((0,), dtype=np.int32)
    return mesh


# the 24 orientation-preserving relabelings of a hexahedron's 8 vertices
def _hex_rotations() -> np.ndarray:
    ref = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1],
                    [-1, -1, 1], [1, -1, 1], [1, 1, 1], [-1, 1, 1]])
    out = []
    import itertools
    for perm in itertools.permutations(range(3)):
        for signs in itertools.product((1, -1), repeat=3):
            R = np.zeros((3, 3), dtype=int)
            for r in range(3):
                R[r, perm[r]] = signs[r]
            if round(np.linalg.det(R)) != 1:
                continue
            rot = ref @ R.T
            out.append([int(np.where((ref == rot[v]).all(axis=1))[0][0]) for v in range(8)])
    return np.array(out, dtype=np.int64)


HEX_ROTATIONS = _hex_rotations()


def scramble(mesh: HexMesh, seed: int = 0, *, rotate: bool = True, permute: bool = True) -> HexMesh:
    """Randomly rotate each cell's local frame and permute cell order: an 'unstructured' numbering of the
    same geometry, exercising every holder/sharer face pairing and orientation."""
    rng = np.random.default_rng(seed)
    cells = mesh.cells.copy()
    if rotate:
        r = rng.integers(0, len(HEX_ROTATIONS), size=len(cells))
        cells = np.take_along_axis(cells, HEX_ROTATIONS[r], axis=1)
    if permute:
        cells = cells[rng.permutation(len(cells))]
    return HexMesh(nodes=mesh.nodes, cells=cells, bfaces=mesh.bfaces, btags=mesh.btags, dims=None,
                   periodic=mesh.periodic)


def block_partition(mesh: HexMesh, px: int, py: int, pz: int) -> np.ndarray:
    """Block decomposition of a structured box: the METIS stand-in of SURVEY.md section 8d (METIS is
    not available; `cell_parts` is an input shared by the oracle and the GPU build)."""
    nx, ny, nz = mesh.dims
    c = np.arange(mesh.n_cells)
    i, j, k = c % nx, (c // nx) % ny, c // (nx * ny)
    bi, bj, bk = i * px // nx, j * py // ny, k * pz // nz
    return (bi + px * (bj + py * bk)).astype(np.int32)


def factor3(n: int) -> tuple[int, int, int]:
    """n = px*py*pz with the factors as equal as possible (px >= py >= pz)."""
    best = (n, 1, 1)
    for a in range(1, n + 1):
        if n % a:
            continue
        for b in range(1, n // a + 1):
            if (n // a) % b:
                continue
            c = n // a // b
            t = tuple(sorted((a, b, c), reverse=True))
            if max(t) - min(t) < max(best) - min(best):
                best = t
    return best
