"""The step before the path: the deterministic index work of `mini::mesh::Shuffler` on the GPU.

METIS itself (the `cell_parts` array) stays with the host; what the reference then does with it -- node parts by the
min-part rule, the part-contiguous order, reordering of data and connectivity -- is integer work these functions hand
to `libminicfd_b200.so`.  Results are bit-exact (include/mini/mesh/metis.hpp:314-335, mesh/shuffler.hpp:35-105).
"""
from __future__ import annotations

import numpy as np

from . import capi


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _run(rc):
    if rc != 0:
        raise capi.McfdError(f"libminicfd_b200 error {rc}: {capi.load().mcfd_last_error(None).decode()}")


def GetNodeParts(range_, nodes, n_nodes, cell_parts, n_parts, device=0):
    """metis::GetNodeParts(mesh, cell_parts, n_parts): a node shared by several parts belongs to the smallest."""
    range_, nodes, cell_parts = _i64(range_), _i64(nodes), _i64(cell_parts)
    out = np.empty(int(n_nodes), dtype=np.int64)
    _run(capi.load().mcfd_get_node_parts(device, int(n_nodes), len(cell_parts), capi._ip(range_), capi._ip(nodes),
                                         capi._ip(cell_parts), int(n_parts), capi._ip(out)))
    return out


def GetNewOrder(parts, device=0):
    """mini::mesh::GetNewOrder(parts, n) -> (new_to_old, old_to_new): indices sorted by (part, index)."""
    parts = _i64(parts)
    n2o, o2n = np.empty(len(parts), dtype=np.int64), np.empty(len(parts), dtype=np.int64)
    _run(capi.load().mcfd_get_new_order(device, len(parts), capi._ip(parts), capi._ip(n2o), capi._ip(o2n)))
    return n2o, o2n


def ShuffleData(new_to_old, data, device=0):
    """mini::mesh::ShuffleData(new_to_old, old_data), in place on the rows of a float64 array."""
    new_to_old = _i64(new_to_old)
    assert data.dtype == np.float64 and data.flags.c_contiguous and len(data) == len(new_to_old)
    width = data.size // max(len(data), 1)
    _run(capi.load().mcfd_shuffle_data(device, len(new_to_old), max(width, 1), capi._ip(new_to_old), capi._dp(data)))
    return data


def ShuffleConnectivity(old_to_new_for_nodes, new_to_old_for_cells, npe, cell_nodes, device=0):
    """mini::mesh::ShuffleConnectivity(...), in place: 1-based node ids renamed, cells reordered."""
    o2n, n2o = _i64(old_to_new_for_nodes), _i64(new_to_old_for_cells)
    assert cell_nodes.dtype == np.int64 and cell_nodes.flags.c_contiguous and cell_nodes.size == len(n2o) * npe
    _run(capi.load().mcfd_shuffle_connectivity(device, len(n2o), int(npe), len(o2n), capi._ip(o2n), capi._ip(n2o),
                                               capi._ip(cell_nodes)))
    return cell_nodes
