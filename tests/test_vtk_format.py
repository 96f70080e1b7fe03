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


def _special_values():
    ties = [12345.25, 1000005.0, 999999.5, 0.5, 2.5, 1234565.0, 123456.5, 0.000123456500000000004, 8.5e-5, 99999.95]
    edges = [0.0, -0.0, 1.0, -1.0, 10.0, 1e5, 1e6, 123456.0, 1234567.0, 999999.0, 999999.4, 1e-4, 1e-5, 9.9999949e-5,
             9.999995e-5, 0.0001234567, 3.14159265358979, 2.718281828e100, 1e100, 1e-100, 1.5e300, 2.5e-300, 1e-310,
             4.9e-324, 1.7976931348623157e308, float("inf"), float("-inf"), float("nan"), 1e22, 1e23, 5e-324, 0.1, 0.3,
             1 / 3, 2 / 3, 100000.5, 99999.5, 0.99999949999, 0.9999995, 1 - 2 ** -53]
    return np.array(ties + edges + [-x for x in ties + edges if x == x])


def test_device_number_formatting_source_on_the_host():
    """format.cuh's FormatG6 compiled for the host: every token it is sure of equals printf's "%g "; what it is not
    sure of (ties, values beyond its power table) is flagged for the host; flags are rare on ordinary data."""
    import emuapi
    rng = np.random.default_rng(5)
    mant = rng.uniform(-10, 10, 400000)
    values = np.concatenate([_special_values(), mant, mant * 10.0 ** rng.integers(-20, 20, mant.size),
                             rng.uniform(0.1, 10, 100000) * 10.0 ** rng.integers(-290, 300, 100000),
                             np.round(rng.uniform(-1e7, 1e7, 100000)), np.round(rng.uniform(-100, 100, 100000), 3)])
    tokens, unsure = emuapi.format_g6(values)
    wrong = [(v, t) for v, t, u in zip(values, tokens, unsure) if not u and t != (b"%g " % v)]
    assert not wrong, wrong[:10]
    for v in (12345.25, 1000005.0, 999999.5, 1e-310, 1.5e305):     # a true tie / outside the table: must be flagged
        assert emuapi.format_g6([v])[1][0]
    assert unsure[len(_special_values()):len(_special_values()) + 800000].mean() < 1e-4
    ints = np.concatenate([[0, 1, 9, 10, 27, 2 ** 31 - 1, -1, -120], rng.integers(0, 2 ** 31, 1000)])
    assert emuapi.format_integer(ints) == [b"%d " % i for i in ints]
