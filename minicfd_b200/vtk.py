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


def write_vtu(path, points, values, field_names=FIELD_NAMES, point_data=(), cell_data=()):
    """vtk.hpp:574-667.  points [n_cells][27][3], values [n_cells][27][K]; point_data = [(name, array[n_cells*27])],
    cell_data = [(name, array[n_cells])]."""
    n_cells = points.shape[0]
    n_points = n_cells * 27
    flat = values.reshape(n_points, -1)
    out = ['<VTKFile type="UnstructuredGrid" version="1.0" byte_order="LittleEndian" header_type="UInt64">',
           "  <UnstructuredGrid>", f'    <Piece NumberOfPoints="{n_points}" NumberOfCells="{n_cells}">',
           "      <PointData>"]
    for k, name in enumerate(field_names):
        out += [f'        <DataArray type="Float64" Name="{name}" format="ascii">', _ascii(flat[:, k]),
                "        </DataArray>"]
    for name, array in point_data:
        out += [f'        <DataArray type="Float64" Name="{name}" format="ascii">', _ascii(np.asarray(array).reshape(-1)),
                "        </DataArray>"]
    out += ["      </PointData>", "      <CellData>"]
    for name, array in cell_data:
        out += [f'        <DataArray type="Float64" Name="{name}" format="ascii">', _ascii(np.asarray(array).reshape(-1)),
                "        </DataArray>"]
    out += ["      </CellData>", "      <Points>",
            '        <DataArray type="Float64" Name="Points" NumberOfComponents="3" format="binary">',
            _base64_block(points.astype(np.float64)), "        </DataArray>", "      </Points>", "      <Cells>",
            '        <DataArray type="Int32" Name="connectivity" format="ascii">',
            "".join("%d " % i for i in range(n_points)), "        </DataArray>",
            '        <DataArray type="Int32" Name="offsets" format="ascii">',
            "".join("%d " % (27 * (i + 1)) for i in range(n_cells)), "        </DataArray>",
            '        <DataArray type="UInt8" Name="types" format="ascii">', ("%d " % HEXAHEDRON27) * n_cells,
            "        </DataArray>", "      </Cells>", "    </Piece>", "  </UnstructuredGrid>", "</VTKFile>", ""]
    with open(path, "w") as f:
        f.write("\n".join(out))


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
        ctx = self.spatial.ctx
        part = self.spatial.part
        if self._points is None:
            self._points, values = ctx.sample_vtk(self.modal, points=True)
        else:
            _, values = ctx.sample_vtk(self.modal, points=False)
        rank, size = getattr(part, "rank", 0), getattr(part, "size", 1)
        names = self.point_data, self.cell_data
        if rank == 0:
            os.makedirs(os.path.join(directory, soln_name), exist_ok=True)
            write_pvtu(directory, soln_name, size, self.field_names, *names)
        if self.spatial.comm is not None:
            import torch.distributed as dist
            dist.barrier(group=self.spatial.comm.group)      # part.hpp:1574
        pts = self._points.reshape(-1, 3)
        pd = [(n, f(pts, values.reshape(len(pts), -1))) for n, f in self.point_data]
        cd = [(n, f()) for n, f in self.cell_data]
        write_vtu(os.path.join(directory, soln_name, f"{rank}.vtu"), self._points, values, self.field_names, pd, cd)
