"""The shuffler's index work on the GPU against the reference's own golden vectors
(test/mesh/shuffler.cpp:29-65, test/mesh/metis.cpp:169-181) and, at size, against a plain restatement of
include/mini/mesh/shuffler.hpp:35-105 and metis.hpp:314-335 in numpy.  Integer results: bit-exact."""
import numpy as np
import pytest

from minicfd_b200 import mesh as meshmod, shuffler

pytestmark = pytest.mark.gpu


def test_get_new_order_and_shuffle_data_golden():            # test/mesh/shuffler.cpp:29-49
    parts = [1, 0, 0, 2, 0, 1, 2, 1, 0, 2]
    n2o, o2n = shuffler.GetNewOrder(parts)
    assert n2o.tolist() == [1, 2, 4, 8, 0, 5, 7, 3, 6, 9]
    assert o2n[n2o].tolist() == list(range(10))
    data = np.array([1.0, 0.0, 0.1, 2.0, 0.2, 1.1, 2.1, 1.2, 0.3, 2.2])
    shuffler.ShuffleData(n2o, data)
    assert data.tolist() == [0.0, 0.1, 0.2, 0.3, 1.0, 1.1, 1.2, 2.0, 2.1, 2.2]


def test_shuffle_connectivity_golden():                       # test/mesh/shuffler.cpp:50-65
    cells = np.array([1, 2, 6, 2, 3, 6, 6, 3, 5, 3, 4, 5], dtype=np.int64)
    shuffler.ShuffleConnectivity([5, 4, 3, 2, 1, 0], [1, 2, 3, 0], 3, cells)
    assert cells.tolist() == [5, 4, 1, 1, 4, 2, 4, 3, 2, 6, 5, 1]


def test_get_node_parts_golden():                             # test/mesh/metis.cpp:169-181
    got = shuffler.GetNodeParts([0, 4, 8, 11, 14], [0, 2, 3, 1, 2, 4, 5, 3, 6, 8, 7, 8, 9, 7], 10, [2, 0, 0, 1], 3)
    assert got.tolist() == [2, 2, 0, 0, 0, 0, 0, 0, 0, 1]


@pytest.mark.parametrize("n,n_parts", [(1, 1), (31, 3), (1024, 8), (1025, 8), (300000, 64), (74088, 1000)])
def test_get_new_order_at_size(n, n_parts):
    rng = np.random.default_rng(n)
    parts = rng.integers(0, n_parts, n)
    n2o, o2n = shuffler.GetNewOrder(parts)
    expect = np.argsort(parts, kind="stable")                  # (part, index) lexicographic, shuffler.hpp:42-45
    assert np.array_equal(n2o, expect)
    inverse = np.empty(n, dtype=np.int64); inverse[expect] = np.arange(n)
    assert np.array_equal(o2n, inverse)


def test_partition_pipeline_on_a_box():
    """The chain of Shuffler::PartitionAndShuffle after METIS (shuffler.hpp:526-546) on BASELINE config 2's mesh with a
    block partition standing in for METIS: node parts, new orders of nodes and cells, coordinates and connectivity
    shuffled -- every table equal to the numpy restatement, and the mesh still the same mesh."""
    m = meshmod.box(42, 42, 42, hi=(10.0, 10.0, 10.0))
    cell_parts = meshmod.block_partition(m, 2, 2, 2).astype(np.int64)
    n_cells, n_nodes = len(m.cells), len(m.nodes)
    range_ = np.arange(n_cells + 1, dtype=np.int64) * 8
    node_parts = shuffler.GetNodeParts(range_, m.cells.reshape(-1), n_nodes, cell_parts, 8)
    expect = np.full(n_nodes, 8, dtype=np.int64)
    np.minimum.at(expect, m.cells.reshape(-1), np.repeat(cell_parts, 8))
    assert np.array_equal(node_parts, expect)
    cn2o, co2n = shuffler.GetNewOrder(cell_parts)
    nn2o, no2n = shuffler.GetNewOrder(node_parts)
    assert np.array_equal(cn2o, np.argsort(cell_parts, kind="stable"))
    assert np.array_equal(nn2o, np.argsort(node_parts, kind="stable"))
    xyz = m.nodes.copy()
    shuffler.ShuffleData(nn2o, xyz)
    assert np.array_equal(xyz, m.nodes[nn2o])
    conn = (m.cells + 1).astype(np.int64).copy()                # CGNS ids are 1-based
    shuffler.ShuffleConnectivity(no2n, cn2o, 8, conn)
    assert np.array_equal(conn, no2n[m.cells[cn2o]] + 1)
    assert np.array_equal(xyz[conn - 1], m.nodes[m.cells[cn2o]])    # same corner coordinates, cell by cell
    assert (np.diff(cell_parts[cn2o]) >= 0).all()                    # part-contiguous


def test_bad_input_is_rejected():
    from minicfd_b200 import capi
    with pytest.raises(capi.McfdError):
        shuffler.GetNewOrder([0, -1, 2])
    with pytest.raises(capi.McfdError):
        shuffler.ShuffleConnectivity([0, 1], [0], 3, np.array([1, 2, 3], dtype=np.int64))
    with pytest.raises(capi.McfdError):
        shuffler.GetNodeParts([0, 2], [0, 5], 3, [0], 1)
