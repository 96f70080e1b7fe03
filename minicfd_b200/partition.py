"""`partition/<rank>.txt`: the per-part index file between `mini::mesh::Shuffler` and `mini::mesh::part::Part`.

The reference partitions a CGNS mesh with METIS, reorders nodes, cells and boundary faces so that every part owns
contiguous ranges (`Shuffler::Shuffle`, include/mini/mesh/shuffler.hpp:241-304), writes the reordered mesh to
`<case>/shuffled.cgns` and one text file per part (`Shuffler::WritePartitionInfo`, shuffler.hpp:307-506) from which
`Part` rebuilds its local cells, ghost cells, adjacency and boundary faces (part.hpp:535-631, 689-745, 806-931,
1456-1549).  CGNS and HDF5 are not on this image, so the mesh itself stays in memory (`ShuffledMesh`, the stand-in for
shuffled.cgns: one zone, section 1 = the HEXA_8 cells, sections 2.. = one QUAD_4 section per boundary tag); the text
files are written and parsed in the reference's format, line for line:

    # i_zone i_node_head i_node_tail
    # i_part i_node_metis                       nodes this part owns that other parts' cells use
    # i_part i_node_metis i_zone i_node         nodes of other parts that this part's cells use
    # i_zone i_sect i_cell_head i_cell_tail
    # i_cell_metis j_cell_metis                 adjacent cells, both in this part (i < j)
    # i_part i_cell_metis j_cell_metis i_node_npe j_node_npe     adjacent cells, j in part i_part
    # i_zone i_sect i_face_head i_face_tail
    #

"metis" ids are positions in the mesh as it was BEFORE the reordering (0-based); zone, section, node, cell and face
ids are CGNS ids of the reordered mesh (1-based, ranges [head, tail)).
"""
from __future__ import annotations

import dataclasses
import os

import numpy as np

from .mesh import HEX_FACE_NODES, HexMesh
from .part import Part, dual_graph


def get_node_parts(cells, cell_parts, n_nodes):
    """metis::GetNodeParts (metis.hpp:314-335): a node belongs to the smallest part among the cells that use it."""
    node_parts = np.full(n_nodes, np.iinfo(np.int32).max, dtype=np.int32)
    np.minimum.at(node_parts, cells.reshape(-1), np.repeat(cell_parts, cells.shape[1]))
    return node_parts


def get_new_order(parts):
    """GetNewOrder (shuffler.hpp:35-63): stable order by part; returns (new_to_old, old_to_new)."""
    new_to_old = np.argsort(parts, kind="stable")
    old_to_new = np.empty_like(new_to_old)
    old_to_new[new_to_old] = np.arange(len(parts))
    return new_to_old, old_to_new


@dataclasses.dataclass
class ShuffledMesh:
    """What shuffled.cgns holds (one zone): reordered nodes and cells, per boundary tag a section of quads, and the
    MetisIndex / PartIndex fields (shuffler.hpp:241-304, 318-353)."""
    nodes: np.ndarray            # [n_nodes, 3]
    node_metis: np.ndarray       # MetisIndex on nodes
    node_parts: np.ndarray       # PartIndex on nodes
    cells: np.ndarray            # [n_cells, 8] 0-based new node ids (CGNS stores them 1-based)
    cell_metis: np.ndarray
    cell_parts: np.ndarray
    face_sections: list          # per boundary tag (ascending): dict(tag, quads [n, 4], parts [n], first CGNS id)
    n_parts: int

    def hex_mesh(self) -> HexMesh:
        quads = [s["quads"] for s in self.face_sections]
        tags = [np.full(len(s["quads"]), s["tag"], dtype=np.int32) for s in self.face_sections]
        return HexMesh(nodes=self.nodes, cells=self.cells,
                       bfaces=np.concatenate(quads) if quads else np.zeros((0, 4), dtype=np.int64),
                       btags=np.concatenate(tags) if tags else np.zeros((0,), dtype=np.int32))


def shuffle(mesh: HexMesh, cell_parts, n_parts: int) -> ShuffledMesh:
    """Shuffler::Shuffle for one zone with one cell section (shuffler.hpp:241-304) + FillFaceParts (a boundary face goes
    with the cell that holds it, shuffler.hpp:189-239)."""
    cell_parts = np.asarray(cell_parts, dtype=np.int32)
    node_parts = get_node_parts(mesh.cells, cell_parts, len(mesh.nodes))
    n2o_nodes, o2n_nodes = get_new_order(node_parts)
    n2o_cells, _ = get_new_order(cell_parts)
    cells = o2n_nodes[mesh.cells[n2o_cells]]
    # the holder of every boundary quad: the cell that has its four nodes
    n = mesh.n_cells
    sections = []
    if len(mesh.bfaces):
        keys = np.sort(mesh.cells[:, HEX_FACE_NODES].reshape(n * 6, 4), axis=1)
        lut = {tuple(k): i // 6 for i, k in enumerate(map(tuple, keys))}
        holder = np.array([lut[tuple(k)] for k in map(tuple, np.sort(mesh.bfaces, axis=1))])
        first = n + 1                                   # CGNS element ids go on after the cells
        for tag in np.unique(mesh.btags):
            sel = np.flatnonzero(mesh.btags == tag)
            parts = cell_parts[holder[sel]]
            n2o_faces, _ = get_new_order(parts)
            sections.append({"tag": int(tag), "quads": o2n_nodes[mesh.bfaces[sel][n2o_faces]], "parts": parts[n2o_faces],
                             "first": first})
            first += len(sel)
    return ShuffledMesh(nodes=mesh.nodes[n2o_nodes], node_metis=n2o_nodes.astype(np.int64), node_parts=node_parts[n2o_nodes],
                        cells=cells, cell_metis=n2o_cells.astype(np.int64), cell_parts=cell_parts[n2o_cells],
                        face_sections=sections, n_parts=n_parts)


def _range_of(parts, p, first):
    """[head, tail) of the contiguous run of part p in a part-sorted field whose first CGNS id is `first`."""
    idx = np.flatnonzero(parts == p)
    if len(idx) == 0:
        return None
    return first + int(idx[0]), first + int(idx[-1]) + 1


def write_partition_info(case_dir: str, mesh: HexMesh, cell_parts, sm: ShuffledMesh, pairs=None):
    """Shuffler::WritePartitionInfo (shuffler.hpp:307-506).  `mesh`, `cell_parts`: the mesh before the reordering (its
    cell and node order is the metis numbering); `sm` = shuffle(mesh, cell_parts, n_parts)."""
    cell_parts = np.asarray(cell_parts, dtype=np.int32)
    n_parts = sm.n_parts
    node_parts = get_node_parts(mesh.cells, cell_parts, len(mesh.nodes))
    pairs = dual_graph(mesh) if pairs is None else pairs
    # the dual graph as METIS stores it: for cell i its neighbours j in ascending order (CSR); shuffler.hpp:405-434
    both = np.concatenate([pairs, pairs[:, ::-1]])
    both = both[np.lexsort((both[:, 1], both[:, 0]))]
    new_node_of_metis = np.empty(len(mesh.nodes), dtype=np.int64)
    new_node_of_metis[sm.node_metis] = np.arange(len(mesh.nodes))
    inner = [[] for _ in range(n_parts)]
    inter = [dict() for _ in range(n_parts)]          # part -> other part -> [(i, j)]
    adj_nodes = [dict() for _ in range(n_parts)]      # part -> other part -> set of metis nodes
    send_nodes = [dict() for _ in range(n_parts)]
    for i in range(mesh.n_cells):                     # cells use nodes of other parts (shuffler.hpp:407-417)
        pi = int(cell_parts[i])
        for node in mesh.cells[i]:
            pn = int(node_parts[node])
            if pn != pi:
                adj_nodes[pi].setdefault(pn, set()).add(int(node))
                send_nodes[pn].setdefault(pi, set()).add(int(node))
    for i, j in both:
        pi, pj = int(cell_parts[i]), int(cell_parts[j])
        if pi == pj:
            if i < j:
                inner[pi].append((int(i), int(j)))
        else:
            inter[pi].setdefault(pj, []).append((int(i), int(j)))
            for node in mesh.cells[j]:                # the ghost cell's nodes (shuffler.hpp:425-433)
                pn = int(node_parts[node])
                if pn != pi:
                    adj_nodes[pi].setdefault(pn, set()).add(int(node))
                    send_nodes[pn].setdefault(pi, set()).add(int(node))
    os.makedirs(os.path.join(case_dir, "partition"), exist_ok=True)
    for p in range(n_parts):
        with open(os.path.join(case_dir, "partition", f"{p}.txt"), "w") as f:
            f.write("# i_zone i_node_head i_node_tail\n")
            r = _range_of(sm.node_parts, p, 1)
            if r:
                f.write(f"1 {r[0]} {r[1]}\n")
            f.write("# i_part i_node_metis\n")
            for q in sorted(send_nodes[p]):
                for node in sorted(send_nodes[p][q]):
                    f.write(f"{q} {node}\n")
            f.write("# i_part i_node_metis i_zone i_node\n")
            for q in sorted(adj_nodes[p]):
                for node in sorted(adj_nodes[p][q]):
                    f.write(f"{q} {node} 1 {int(new_node_of_metis[node]) + 1}\n")
            f.write("# i_zone i_sect i_cell_head i_cell_tail\n")
            r = _range_of(sm.cell_parts, p, 1)
            if r:
                f.write(f"1 1 {r[0]} {r[1]}\n")
            f.write("# i_cell_metis j_cell_metis\n")
            for i, j in inner[p]:
                f.write(f"{i} {j}\n")
            f.write("# i_part i_cell_metis j_cell_metis i_node_npe j_node_npe\n")
            for q in sorted(inter[p]):
                for i, j in inter[p][q]:
                    f.write(f"{q} {i} {j} 8 8\n")
            f.write("# i_zone i_sect i_face_head i_face_tail\n")
            for s, sec in enumerate(sm.face_sections):
                r = _range_of(sec["parts"], p, sec["first"])
                if r:
                    f.write(f"1 {s + 2} {r[0]} {r[1]}\n")
            f.write("#\n")


@dataclasses.dataclass
class PartitionInfo:
    """One partition/<rank>.txt, parsed block by block the way part.hpp:535-560, 577-613, 689-694, 806-827, 1456-1462 read it."""
    node_ranges: list        # (zone, head, tail)
    send_nodes: list         # (part, metis node)
    adj_nodes: list          # (part, metis node, zone, node)
    cell_ranges: list        # (zone, sect, head, tail)
    inner_adjs: list         # (i, j) metis cells
    inter_adjs: list         # (part, i, j, npe_i, npe_j)
    face_ranges: list        # (zone, sect, head, tail)


def read_partition_info(path: str) -> PartitionInfo:
    blocks, cur = [], None
    with open(path) as f:
        for line in f:
            if line.startswith("#"):
                cur = []
                blocks.append(cur)
            else:
                cur.append(tuple(int(x) for x in line.split()))
    if len(blocks) != 8 or blocks[7]:
        raise ValueError(f"{path}: expected seven blocks closed by '#'")
    return PartitionInfo(*blocks[:7])


def part_from_partition(info: PartitionInfo, sm: ShuffledMesh, rank: int) -> Part:
    """`Part::Part(case, rank, size)` up to the geometry (part.hpp:442-476): local cells from the cell ranges
    (part.hpp:689-745), inner / inter from the inter-part adjacency (part.hpp:904-916), ghost cells = its far ends
    (part.hpp:828-903), send / receive lists per neighbouring part in ascending metis id (std::map, part.hpp:904-925),
    boundary faces from the face ranges (part.hpp:1456-1549).  Returns the same object as Part(mesh, cell_parts, rank,
    n_parts) builds from the whole mesh; ids refer to `sm.hex_mesh()`."""
    mesh = sm.hex_mesh()
    n = mesh.n_cells
    new_of_metis = np.empty(n, dtype=np.int64)
    new_of_metis[sm.cell_metis] = np.arange(n)
    p = Part.__new__(Part)
    p.mesh, p.rank, p.size = mesh, rank, sm.n_parts
    local = np.concatenate([np.arange(h - 1, t - 1) for (_, _, h, t) in info.cell_ranges]) if info.cell_ranges \
        else np.zeros(0, dtype=np.int64)
    parts = -np.ones(n, dtype=np.int32)
    parts[local] = rank
    send, recv = {}, {}
    for q, i, j, _, _ in info.inter_adjs:
        send.setdefault(q, set()).add(i)
        recv.setdefault(q, set()).add(j)
        parts[new_of_metis[j]] = q
    p.cell_parts = parts
    is_inter = np.zeros(n, dtype=bool)
    for q in send:
        is_inter[new_of_metis[sorted(send[q])]] = True
    inner, inter = local[~is_inter[local]], local[is_inter[local]]
    p.local_global = np.concatenate([inner, inter])
    p.n_inner, p.n_local = len(inner), len(local)
    # ghost cells: canonical order = ascending id in the mesh at hand (part.py, module docstring)
    ghosts = sorted({int(new_of_metis[j]) for q in recv for j in recv[q]})
    p.ghost_global = np.array(ghosts, dtype=np.int64)
    p.n_ghost = len(ghosts)
    p.global_ids = np.concatenate([p.local_global, p.ghost_global])
    p.global_to_id = -np.ones(n, dtype=np.int64)
    p.global_to_id[p.global_ids] = np.arange(len(p.global_ids))
    p.neighbours = np.array(sorted(send), dtype=np.int32)
    p.send_cells = [p.global_to_id[np.sort(new_of_metis[sorted(send[q])])] for q in sorted(send)]
    p.recv_cells = [p.global_to_id[np.sort(new_of_metis[sorted(recv[q])])] for q in sorted(send)]
    p.cell_nodes = mesh.cells[p.global_ids]
    quads, tags = [], []
    for (_, sect, head, tail) in info.face_ranges:
        sec = sm.face_sections[sect - 2]
        quads.append(sec["quads"][head - sec["first"]:tail - sec["first"]])
        tags.append(np.full(tail - head, sec["tag"], dtype=np.int32))
    p.bfaces = np.concatenate(quads) if quads else np.zeros((0, 4), dtype=np.int64)
    p.btags = np.concatenate(tags) if tags else np.zeros((0,), dtype=np.int32)
    return p


def partition_and_shuffle(case_dir: str, mesh: HexMesh, n_parts: int, cell_parts=None) -> ShuffledMesh:
    """`Shuffler::PartitionAndShuffle` (shuffler.hpp:510-546) without its file I/O: dual graph of the cells
    (METIS_MeshToDual, n_common = 3), the cells' parts, the nodes' parts (metis::GetNodeParts), the reordering, and
    the per-part index files under `<case_dir>/partition/`.  METIS_PartGraphKway itself is not on this image
    (DESIGN.md section 6): `cell_parts` comes from the caller, or, failing that, consecutive runs of cells in mesh
    order — a valid input for everything downstream, not METIS's answer."""
    if cell_parts is None:
        cell_parts = (np.arange(mesh.n_cells) * n_parts // max(mesh.n_cells, 1)).astype(np.int32)
    pairs = dual_graph(mesh)
    sm = shuffle(mesh, cell_parts, n_parts)
    write_partition_info(case_dir, mesh, cell_parts, sm, pairs=pairs)
    return sm
