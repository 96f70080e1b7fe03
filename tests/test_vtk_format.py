"""The host side of the frame output (minicfd_b200/vtk.py) on the CPU: fed with the numbers the unmodified reference
collected (tests/golden/vtk_*.npz), the writers must reproduce the reference's .vtu and .pvtu files byte for byte --
arrays, order, number formatting, base64 framing of the points."""
import os

import numpy as np
import pytest

import cases
from minicfd_b200 import vtk

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", [c[0] for c in cases.VTK_GOLDEN_CASES])
def test_writers_reproduce_the_reference_files(name, tmp_path):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    vtk.write_vtu(str(tmp_path / "0.vtu"), g["points"], g["values"])
    assert open(tmp_path / "0.vtu").read() == str(g["vtu"])
    vtk.write_pvtu(str(tmp_path), "Frame0", 1)
    assert open(tmp_path / "Frame0.pvtu").read() == str(g["pvtu"])


def test_extra_point_and_cell_data_and_parts(tmp_path):
    pts = np.zeros((2, 27, 3)); val = np.ones((2, 27, 5))
    vtk.write_vtu(str(tmp_path / "1.vtu"), pts, val, point_data=[("Pressure", np.full(54, 0.4))],
                  cell_data=[("CellViscosity1", np.array([0.0, 1e-5]))])
    text = open(tmp_path / "1.vtu").read()
    assert '<DataArray type="Float64" Name="Pressure" format="ascii">\n' + "0.4 " * 54 + "\n" in text
    assert '<DataArray type="Float64" Name="CellViscosity1" format="ascii">\n0 1e-05 \n' in text
    assert 'NumberOfPoints="54" NumberOfCells="2"' in text and "27 54 \n" in text and "29 29 \n" in text
    vtk.write_pvtu(str(tmp_path), "Frame3", 4, point_data=[("Pressure", None)], cell_data=[("CellViscosity1", None)])
    text = open(tmp_path / "Frame3.pvtu").read()
    assert text.count("<Piece Source=") == 4 and '<Piece Source="./Frame3/3.vtu"/>' in text
    assert '<PDataArray type="Float64" Name="Pressure"/>' in text and '<PDataArray type="Float64" Name="CellViscosity1"/>' in text
