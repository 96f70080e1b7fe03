"""Frame output: the host side of `mini::mesh::vtk::Writer<Part>::WriteSolutions` (include/mini/mesh/vtk.hpp:526-668).

The values at the 27 points of every cell's VTK tri-quadratic hexahedron come from the device
(`mcfd_sample_vtk`, one kernel over the resident state); this module only lays them out as the reference's
`.vtu` piece per part plus the `.pvtu` index written by rank 0: same elements, same arrays in the same order, same
number formatting (`operator<<(double)` = "%g", points base64-encoded behind a UInt64 byte count).
"""
from __future__ import annotations

import base64
import os

import numpy as np

FIELD_NAMES = ("Density", "MomentumX", "MomentumY", "MomentumZ", "EnergyStagnationDensity")  # demo/euler/sourceless.cpp:109-110
HEXAHEDRON27 = 29   # vtk.hpp:139, chosen for 8-corner cells at vtk.hpp:407-408


def _ascii(values) -> str:
    """`for (v : values) vtu << v << ' ';` with the stream's default format (vtk.hpp:600-602)."""
    return "".join("%g " % v for v in values)


def _base64_block(array: np.ndarray) -> str:
    """vtk.hpp:638-646: the byte count as one UInt64, encoded on its own, then the encoded bytes."""
    raw = np.ascontiguousarray(array).tobytes()
    head = base64.b64encode(np.uint64(len(raw)).tobytes()).decode()
    return head + base64.b64encode(raw).decode()


def write_pvtu(directory, soln_name, n_parts, field_names=FIELD_NAMES, point_data=(), cell_data=()):
    """vtk.hpp:541-573 (rank 0 only)."""
    lines = ['<VTKFile type="PUnstructuredGrid" version="1.0" byte_order="LittleEndian" header_type="UInt64">',
             '  <PUnstructuredGrid GhostLevel="1">', "    <PPointData>"]
    for name in list(field_names) + [n for n, _ in point_data]:
        lines.append(f'      <PDataArray type="Float64" Name="{name}"/>')
    lines += ["    </PPointData>", "    <PCellData>"]
    for name, _ in cell_data:
        lines.append(f'      <PDataArray type="Float64" Name="{name}"/>')
    lines += ["    </PCellData>", "    <PPoints>",
              '      <PDataArray type="Float64" Name="Points" NumberOfComponents="3"/>', "    </PPoints>"]
    for r in range(n_parts):
        lines.append(f'    <Piece Source="./{soln_name}/{r}.vtu"/>')
    lines += ["  </PUnstructuredGrid>", "</VTKFile>", ""]
    with open(os.path.join(directory, soln_name + ".pvtu"), "w") as f:
        f.write("\n".join(lines))


def _assemble(path, n_cells, fields, point_data, cell_data, points_block, connectivity, offsets, types):
    """vtk.hpp:574-667 from ready-made array bodies (bytes): fields / point_data / cell_data = [(name, body)]."""
    n_points = n_cells * 27
    out = [b'<VTKFile type="UnstructuredGrid" version="1.0" byte_order="LittleEndian" header_type="UInt64">',
           b"  <UnstructuredGrid>", b'    <Piece NumberOfPoints="%d" NumberOfCells="%d">' % (n_points, n_cells),
           b"      <PointData>"]
    for name, body in list(fields) + list(point_data):
        out += [b'        <DataArray type="Float64" Name="%s" format="ascii">' % name.encode(), body, b"        </DataArray>"]
    out += [b"      </PointData>", b"      <CellData>"]
    for name, body in cell_data:
        out += [b'        <DataArray type="Float64" Name="%s" format="ascii">' % name.encode(), body, b"        </DataArray>"]
    out += [b"      </CellData>", b"      <Points>",
            b'        <DataArray type="Float64" Name="Points" NumberOfComponents="3" format="binary">',
            points_block, b"        </DataArray>", b"      </Points>", b"      <Cells>",
            b'        <DataArray type="Int32" Name="connectivity" format="ascii">', connectivity, b"        </DataArray>",
            b'        <DataArray type="Int32" Name="offsets" format="ascii">', offsets, b"        </DataArray>",
            b'        <DataArray type="UInt8" Name="types" format="ascii">', types,
            b"        </DataArray>", b"      </Cells>", b"    </Piece>", b"  </UnstructuredGrid>", b"</VTKFile>", b""]
    with open(path, "wb") as f:
        f.write(b"\n".join(out))


def write_vtu(path, points, values, field_names=FIELD_NAMES, point_data=(), cell_data=()):
    """One part's piece with every number formatted on the host.  points [n_cells][27][3], values [n_cells][27][K];
    point_data = [(name, array[n_cells*27])], cell_data = [(name, array[n_cells])]."""
    n_cells = points.shape[0]
    n_points = n_cells * 27
    flat = values.reshape(n_points, -1)
    _assemble(path, n_cells, [(name, _ascii(flat[:, k]).encode()) for k, name in enumerate(field_names)],
              [(name, _ascii(np.asarray(a).reshape(-1)).encode()) for name, a in point_data],
              [(name, _ascii(np.asarray(a).reshape(-1)).encode()) for name, a in cell_data],
              _base64_block(points.astype(np.float64)).encode(),
              "".join("%d " % i for i in range(n_points)).encode(),
              "".join("%d " % (27 * (i + 1)) for i in range(n_cells)).encode(), (b"%d " % HEXAHEDRON27) * n_cells)


class Writer:
    """`mini::mesh::vtk::Writer<Part>` for one Spatial: AddPointData / AddCellData / WriteSolutions."""

    def __init__(self, spatial, modal=False, field_names=FIELD_NAMES):
        self.spatial, self.modal, self.field_names = spatial, bool(modal), tuple(field_names)
        self.point_data, self.cell_data = [], []
        self._points = None          # the geometry does not change between frames

    def AddPointData(self, name, func):
        """func(points[n][3], values[n][K]) -> array[n] (vtk.hpp:505-508; vectorised over the points)."""
        self.point_data.append((name, func))

    def AddCellData(self, name, func):
        """func() -> array[n_local] (vtk.hpp:517-520)."""
        self.cell_data.append((name, func))

    def WriteSolutions(self, directory, soln_name):
        """vtk.hpp:526-668.  The values are sampled AND turned into text on the device (mcfd_sample_vtk,
        mcfd_vtk_component_text, mcfd_format_*); the host encodes the (static) points once and writes bytes."""
        ctx = self.spatial.ctx
        part = self.spatial.part
        n_cells = ctx.n_local
        need_values = bool(self.point_data)          # extra point data is computed by host callbacks
        if self._points is None:
            self._points, _ = ctx.sample_vtk(self.modal, points=True)
            self._points_block = _base64_block(self._points).encode()
            self._cells = (ctx.format_integers(n_cells * 27, 0, 1), ctx.format_integers(n_cells, 27, 27),
                           ctx.format_integers(n_cells, HEXAHEDRON27, 0))
        values = None
        if need_values:
            _, values = ctx.sample_vtk(self.modal, points=False)
        else:
            ctx.sample_vtk_on_device(self.modal)
        fields = [(name, ctx.vtk_component_text(k)) for k, name in enumerate(self.field_names)]
        rank, size = getattr(part, "rank", 0), getattr(part, "size", 1)
        if rank == 0:
            os.makedirs(os.path.join(directory, soln_name), exist_ok=True)
            write_pvtu(directory, soln_name, size, self.field_names, self.point_data, self.cell_data)
        if self.spatial.comm is not None:
            import torch.distributed as dist
            dist.barrier(group=self.spatial.comm.group)      # part.hpp:1574
        pts = self._points.reshape(-1, 3)
        pd = [(n, ctx.format_doubles(f(pts, values.reshape(len(pts), -1)))) for n, f in self.point_data]
        cd = [(n, ctx.format_doubles(f())) for n, f in self.cell_data]
        _assemble(os.path.join(directory, soln_name, f"{rank}.vtu"), n_cells, fields, pd, cd, self._points_block,
                  *self._cells)
