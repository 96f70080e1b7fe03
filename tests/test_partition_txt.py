"""partition/<rank>.txt (minicfd_b200/partition.py): the index file between mini::mesh::Shuffler and
mini::mesh::part::Part, written and parsed in the reference's format (include/mini/mesh/shuffler.hpp:307-506,
include/mini/mesh/part.hpp:535-631, 689-745, 806-931, 1456-1549)."""
import os

import numpy as np
import pytest

import cases
import oracle_check as oc
from minicfd_b200 import mesh as meshmod, partition
from minicfd_b200.part import Part, cell_points

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "partition_two_cells")


def test_written_files_match_the_hand_checked_ones(tmp_path):
    m = meshmod.box(2, 1, 1)
    partition.write_partition_info(str(tmp_path), m, [0, 1], partition.shuffle(m, [0, 1], 2))
    for r in range(2):
        assert open(tmp_path / "partition" / f"{r}.txt").read() == open(os.path.join(GOLDEN, f"{r}.txt")).read()


def test_reader_takes_the_blocks_in_the_reference_order():
    info = partition.read_partition_info(os.path.join(GOLDEN, "0.txt"))
    assert info.node_ranges == [(1, 1, 9)] and info.cell_ranges == [(1, 1, 1, 2)]
    assert info.inner_adjs == [] and info.inter_adjs == [(1, 0, 1, 8, 8)]
    assert (1, 2, 1, 9) in info.adj_nodes and len(info.send_nodes) == 8 and len(info.face_ranges) == 5


def _shuffled(n_parts, seed):
    m = meshmod.scramble(meshmod.box(6, 5, 4, warp=0.1, seed=1), seed=2)
    parts = np.random.default_rng(seed).integers(0, n_parts, m.n_cells).astype(np.int32)   # as ragged as METIS may make it
    return m, parts, partition.shuffle(m, parts, n_parts)


@pytest.mark.parametrize("n_parts", [2, 5])
def test_part_from_the_file_equals_part_from_the_whole_mesh(tmp_path, n_parts):
    """shuffle -> write -> read -> Part: every table (local ids with inner cells first, ghost ids, neighbours, send and
    receive lists, boundary faces) equals what Part builds when it sees the whole mesh and the partition vector."""
    m, parts, sm = _shuffled(n_parts, seed=n_parts)
    assert np.all(np.diff(sm.cell_parts) >= 0) and np.all(np.diff(sm.node_parts) >= 0)     # contiguous ranges per part
    assert np.array_equal(sm.nodes, m.nodes[sm.node_metis]) and np.array_equal(sm.node_metis[sm.cells], m.cells[sm.cell_metis])
    partition.write_partition_info(str(tmp_path), m, parts, sm)
    for r in range(n_parts):
        info = partition.read_partition_info(str(tmp_path / "partition" / f"{r}.txt"))
        a = partition.part_from_partition(info, sm, r)
        b = Part(sm.hex_mesh(), sm.cell_parts, r, n_parts)
        for k in ("local_global", "ghost_global", "global_ids", "neighbours", "bfaces", "btags", "cell_nodes"):
            assert np.array_equal(getattr(a, k), getattr(b, k)), (r, k)
        assert (a.n_inner, a.n_local, a.n_ghost) == (b.n_inner, b.n_local, b.n_ghost)
        for x, y in zip(a.send_cells + a.recv_cells, b.send_cells + b.recv_cells):
            assert np.array_equal(x, y)
        # the nodes the file says this part needs from others are exactly those of its cells it does not own
        used = np.unique(sm.cells[a.global_ids])
        foreign = {int(sm.node_metis[x]) for x in used if sm.node_parts[x] != r}
        assert foreign == {node for (_, node, _, _) in info.adj_nodes}
        assert all(sm.node_metis[i_node - 1] == node for (_, node, _, i_node) in info.adj_nodes)


def test_partition_and_shuffle_writes_one_file_per_part(tmp_path):
    m = meshmod.scramble(meshmod.box(4, 4, 2), seed=3)
    sm = partition.partition_and_shuffle(str(tmp_path), m, 4)
    assert sorted(os.listdir(tmp_path / "partition")) == ["0.txt", "1.txt", "2.txt", "3.txt"]
    n_local = 0
    for r in range(4):
        p = partition.part_from_partition(partition.read_partition_info(str(tmp_path / "partition" / f"{r}.txt")), sm, r)
        n_local += p.n_local
    assert n_local == m.n_cells


def test_cpp_reader_builds_the_same_part(tmp_path):
    """include/mini/b200/partition.hpp: ReadPartitionInfo + BuildPartFromPartition (the C++ side of part.hpp:442-476
    without its CGNS reads) on the files this module wrote: the shim's Part gets the same local order, ghost cells, halo
    lists and boundary faces as the Python part.  No GPU involved."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    m, parts, sm = _shuffled(3, seed=11)
    partition.write_partition_info(str(tmp_path), m, parts, sm)
    names = {0: "XMin", 1: "XMax", 2: "YMin", 3: "YMax", 4: "ZMin", 5: "ZMax"}
    with open(tmp_path / "mesh.bin", "wb") as f:
        i64 = lambda *v: f.write(np.asarray(v, dtype=np.int64).tobytes())
        i64(len(sm.nodes)); f.write(np.ascontiguousarray(sm.nodes, dtype=np.float64).tobytes())
        i64(len(sm.cells)); f.write(np.ascontiguousarray(sm.cells, dtype=np.int64).tobytes())
        f.write(np.ascontiguousarray(sm.cell_metis, dtype=np.int64).tobytes())
        i64(len(sm.face_sections))
        for sec in sm.face_sections:
            name = names[sec["tag"]].encode()
            i64(len(name)); f.write(name); i64(sec["first"]); i64(len(sec["quads"]))
            f.write(np.ascontiguousarray(sec["quads"], dtype=np.int64).tobytes())
    exe = tmp_path / "partition_reader"
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-I", os.path.join(root, "include"),
                           os.path.join(root, "tests", "drivers", "partition_reader.cpp"), "-o", str(exe),
                           "-L", os.path.join(root, "minicfd_b200"), "-lminicfd_b200", "-Wl,-rpath," + os.path.join(root, "minicfd_b200")])
    tag_order = [sec["tag"] for sec in sm.face_sections]      # the shim numbers its tags in the order of the sections
    for r in range(3):
        out = subprocess.run([str(exe), str(tmp_path), str(r), "3", str(tmp_path / "mesh.bin")], capture_output=True, text=True, timeout=60)
        assert out.returncode == 0, out.stderr
        rows = {l.split()[0]: [int(x) for x in l.split()[1:]] for l in out.stdout.splitlines() if not l.startswith("local")}
        p = partition.part_from_partition(partition.read_partition_info(str(tmp_path / "partition" / f"{r}.txt")), sm, r)
        assert out.stdout.splitlines()[0] == f"local {p.n_local} ghost {p.n_ghost} inner {p.n_inner}"
        assert rows["ids"] == sm.cell_metis[p.global_ids].tolist()
        assert rows["neighbours"] == p.neighbours.tolist()
        assert rows["send_counts"] == [len(x) for x in p.send_cells]
        assert rows["send"] == np.concatenate(p.send_cells).tolist() and rows["recv"] == np.concatenate(p.recv_cells).tolist()
        assert rows["bfaces"] == p.bfaces.reshape(-1).tolist()
        assert [tag_order[t] for t in rows["btags"]] == p.btags.tolist()
        assert rows["cells"] == p.cell_nodes.reshape(-1).tolist()


@pytest.mark.gpu
def test_parts_from_the_files_advance_like_the_whole_mesh(tmp_path):
    """Three parts built from their partition files, advanced on one GPU with the device-side ghost exchange, against
    the oracle on the mesh before the reordering (the files' metis ids lead back to it)."""
    import torch
    from minicfd_b200 import capi, peer
    from minicfd_b200.spatial import Spatial
    m, parts, sm = _shuffled(3, seed=7)
    partition.write_partition_info(str(tmp_path), m, parts, sm)
    P = [partition.part_from_partition(partition.read_partition_info(str(tmp_path / "partition" / f"{r}.txt")), sm, r) for r in range(3)]
    u = cases.noisy_state(cell_points(m.nodes, m.cells, 2), seed=9)            # [metis cell][135]
    o = oc.Oracle(m, 2, 0, capi.HLLC)
    for tag in range(6):
        o.set_bc(tag, 0)
    S = [Spatial(p, degree=2, riemann=capi.HLLC, device=0) for p in P]
    streams = [torch.cuda.Stream() for _ in S]
    u_new = u.reshape(-1, 135)[sm.cell_metis].reshape(-1)                      # the same field on the reordered cells
    for sp, p, st in zip(S, P, streams):
        for tag in range(6):
            sp.SetInviscidWall(tag)
        sp.ctx.set_stream(st.cuda_stream)
        sp.SetSolutionColumn(p.gather_from_global(u_new, 135))
    peer.connect_local(S, P)
    ref, t = u.copy(), 0.0
    for _ in range(3):
        capi.peer_step_group([sp.ctx for sp in S], 3, t, 2e-3)
        dt = o.time_step(ref, 2e-3, 3)
        ref = o.rk_update(ref, 3, t, dt)
        t += dt
    got = np.zeros(m.n_cells * 135)
    for sp, p in zip(S, P):
        sp.ctx.peer_check()
        p.scatter_to_global(sp.GetSolutionColumn(), 135, got)
    back = np.empty_like(got).reshape(-1, 135)
    back[sm.cell_metis] = got.reshape(-1, 135)
    assert np.abs(back.reshape(-1) - ref).max() / np.abs(ref).max() < 1e-12
