"""Frame output: `mcfd_sample_vtk` and the .vtu/.pvtu writers against what the unmodified
reference's `mini::mesh::vtk::Writer<Part>` collects and writes (tests/golden/vtk_*.npz, made by
tests/golden/make_vtk_golden.py).  Sampled values: 1e-12 relative.  Files: every structural line identical; the ASCII
numbers carry six significant digits, so they are compared as numbers at one unit of the sixth digit and, counted as
text, must agree for all but a handful of values that sit on a rounding boundary."""
import base64
import os
import subprocess

import numpy as np
import pytest

import cases
from minicfd_b200 import capi, mesh as meshmod, vtk
from minicfd_b200.part import Part
from minicfd_b200.spatial import Spatial

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
TOL = 1e-12


def relerr(a, b):
    return np.abs(a - b).max() / np.abs(b).max()


def spatial_for(g, degree, kind):
    m = meshmod.HexMesh(nodes=g["nodes"], cells=g["cells"], bfaces=g["bfaces"], btags=g["btags"])
    scheme = capi.DG_GENERAL if kind == "projection" else capi.FR_LOBATTO
    sp = Spatial(Part(m), scheme=scheme, degree=degree, riemann=capi.HLLC, device=0, k_local=(kind == "local"))
    cases.apply_bcs_spatial(sp, cases.MIXED_BCS)
    if kind == "extrapolation":
        sp.SetLimiter(*cases.WENO)
    sp.SetSolutionColumn(g["u"])
    return sp


def compare_vtu(got: str, ref: str):
    """Structural lines identical; data lines equal as numbers.  Returns (values compared, values whose text differs)."""
    a, b = got.split("\n"), ref.split("\n")
    assert len(a) == len(b)
    n_values = n_text_diff = 0
    binary = False
    for la, lb in zip(a, b):
        if lb.lstrip().startswith("<") or not lb:
            assert la == lb
            binary = 'format="binary"' in lb
            continue
        if binary:     # UInt64 byte count (12 characters of base64), then the Float64 bytes
            assert la[:12] == lb[:12]
            xa = np.frombuffer(base64.b64decode(la[12:]), dtype=np.float64)
            xb = np.frombuffer(base64.b64decode(lb[12:]), dtype=np.float64)
            assert xa.shape == xb.shape and np.abs(xa - xb).max() <= 1e-14 * np.abs(xb).max()
            continue
        ta, tb = la.split(), lb.split()
        assert len(ta) == len(tb)
        xa, xb = np.array(ta, dtype=np.float64), np.array(tb, dtype=np.float64)
        assert np.all(np.abs(xa - xb) <= 1.01e-5 * np.abs(xb) + 1e-13)
        n_values += len(tb)
        # rounding noise around zero (the transverse momenta of a 1-D flow) prints as different noise on both sides
        n_text_diff += sum(x != y for x, y, v in zip(ta, tb, xb) if abs(v) > 1e-10)
    return n_values, n_text_diff


@pytest.mark.parametrize("name,degree,scheme,kind", cases.VTK_GOLDEN_CASES)
def test_sample_vtk_golden(name, degree, scheme, kind, tmp_path):
    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    sp = spatial_for(g, degree, kind)
    assert relerr(sp.GetSolutionColumn(), g["state"]) < TOL
    modal = kind in ("extrapolation", "projection")
    points, values = sp.ctx.sample_vtk(modal)
    assert relerr(points, g["points"]) < 1e-14
    assert relerr(values, g["values"]) < TOL
    _, again = sp.ctx.sample_vtk(modal, points=False)       # frames after the first skip the static geometry
    assert np.array_equal(again, values)
    w = vtk.Writer(sp, modal=modal)
    w.WriteSolutions(str(tmp_path), "Frame0")
    assert open(tmp_path / "Frame0.pvtu").read() == str(g["pvtu"])
    n, diff = compare_vtu(open(tmp_path / "Frame0" / "0.vtu").read(), str(g["vtu"]))
    nc = len(g["cells"])
    assert n == (5 * 27 + 27 + 2) * nc and diff <= max(2, n // 2000)    # fields, connectivity, offsets, types


def test_nodal_p2_samples_are_the_solution_nodes():
    """At p = 2 the Lobatto nodes are the VTK points: sampling is a permutation of the stored values, bit for bit."""
    g = np.load(os.path.join(GOLDEN, "vtk_nodal_p2.npz"))
    sp = spatial_for(g, 2, "nodal")
    _, values = sp.ctx.sample_vtk(False)
    u = g["u"].reshape(-1, 3, 3, 3, 5)
    for p, (a, b, c) in enumerate([(-1, -1, -1), (1, -1, -1), (1, 1, -1), (-1, 1, -1), (0, 0, 0)]):
        at = p if p < 4 else 26
        assert np.array_equal(values[:, at], u[:, a + 1, b + 1, c + 1])


def test_sample_vtk_at_size_reproduces_a_trilinear_field():
    """Property at BASELINE config 2's size: a field that is linear in x, y, z is reproduced exactly (to rounding) at
    every sample point by all evaluators, on 42^3 cells."""
    m = meshmod.box(42, 42, 42, hi=(10.0, 10.0, 10.0))
    f = lambda x: np.stack([1 + 0.01 * x[..., 0], 0.02 * x[..., 1], -0.03 * x[..., 2], 0.01 * (x[..., 0] - x[..., 1]),
                            3 + 0.02 * x[..., 2]], axis=-1)
    for degree, modal in ((2, False), (2, True), (3, False)):
        sp = Spatial(Part(m), degree=degree, riemann=capi.HLLC, device=0)
        sp.Approximate(f)
        points, values = sp.ctx.sample_vtk(modal)
        assert points.shape == (42 ** 3, 27, 3)
        assert np.abs(values - f(points)).max() < 1e-12


def test_driver_writes_frames_like_the_reference(tmp_path):
    """demo/euler/shock_tube.cpp -DVISCOSITY with a frame directory: VtkWriter::AddCellData("CellViscosity1..5") and
    VtkWriter::WriteSolutions(part, "Frame0" / "Frame10") as demo/euler/sourceless.cpp:122-126, 161-162, 219-221 spell
    them, against the reference run on the same problem."""
    import refapi
    if not refapi.available(2):
        pytest.skip("oracle/_ref not shipped")
    import test_gpu_cpp_shim as shim
    exe = tmp_path / "shock_tube_viscosity"
    subprocess.check_call(["g++", "-DUSE_HLLC", "-DVISCOSITY", "-std=c++20", "-O2", "-I", os.path.join(ROOT, "include"),
                           "-I", os.path.join(ROOT, "include", "mini", "b200", "nompi"),
                           os.path.join(ROOT, "demo", "euler", "shock_tube.cpp"), "-o", str(exe),
                           "-L", os.path.join(ROOT, "minicfd_b200"), "-lminicfd_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "minicfd_b200")])
    ours = tmp_path / "ours"
    ours.mkdir()
    res = subprocess.run([str(exe), "10", str(tmp_path / "column.bin"), str(ours)], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    theirs = tmp_path / "theirs"
    theirs.mkdir()
    shim._sod_viscosity_reference(10, frames=str(theirs))
    for frame in ("Frame0", "Frame10"):
        assert open(ours / (frame + ".pvtu")).read() == open(theirs / (frame + ".pvtu")).read()
        n, diff = compare_vtu(open(ours / frame / "0.vtu").read(), open(theirs / frame / "0.vtu").read())
        assert n == 300 * (27 * 5 + 5 + 27 + 2) and diff <= n // 1000


def test_numbers_formatted_on_the_device_equal_printf():
    """mcfd_format_doubles / mcfd_format_integers: the packed text equals `"%g " % v` for every value, including the
    ones the device hands to the host (ties, denormals, huge values), NaNs, infinities and signed zeros."""
    import test_vtk_format as fmt
    g = np.load(os.path.join(GOLDEN, "vtk_nodal_p2.npz"))
    sp = spatial_for(g, 2, "nodal")
    rng = np.random.default_rng(11)
    mant = rng.uniform(-10, 10, 300000)
    values = np.concatenate([fmt._special_values(), mant, mant * 10.0 ** rng.integers(-30, 30, mant.size),
                             np.round(rng.uniform(-1e7, 1e7, 50000)), np.round(rng.uniform(-100, 100, 50000), 2)])
    expect = b"".join(b"%g " % v for v in values)
    assert bytes(sp.ctx.format_doubles(values)) == expect
    assert bytes(sp.ctx.format_doubles(values[:5])) == b"".join(b"%g " % v for v in values[:5])     # scratch is reused
    ties = np.full(70000, 12345.25)                       # more undecidable values than the device's list holds
    assert bytes(sp.ctx.format_doubles(ties)) == b"12345.2 " * 70000
    assert bytes(sp.ctx.format_doubles(np.zeros(0))) == b""
    assert bytes(sp.ctx.format_integers(5000, 0, 1)) == b"".join(b"%d " % i for i in range(5000))
    assert bytes(sp.ctx.format_integers(3000, 27, 27)) == b"".join(b"%d " % (27 * (i + 1)) for i in range(3000))
    assert bytes(sp.ctx.format_integers(7, 29, 0)) == b"29 " * 7


def test_component_text_equals_host_formatting_of_the_samples():
    g = np.load(os.path.join(GOLDEN, "vtk_extrapolation_p3.npz"))
    sp = spatial_for(g, 3, "extrapolation")
    _, values = sp.ctx.sample_vtk(True)
    sp.ctx.sample_vtk_on_device(True)
    for k in range(5):
        assert bytes(sp.ctx.vtk_component_text(k)) == b"".join(b"%g " % v for v in values[:, :, k].reshape(-1))


def test_frame_at_size_device_text_equals_host_text(tmp_path):
    """BASELINE config 2's mesh (42^3 cells, 2 M points): the frame written with device-side text is the file the
    host-side formatting writes, byte for byte."""
    m = meshmod.box(42, 42, 42, hi=(10.0, 10.0, 10.0))
    from minicfd_b200 import fields
    sp = Spatial(Part(m), degree=2, riemann=capi.HLLC, device=0)
    sp.Approximate(lambda x: fields.vortex()(x, 0.3))
    w = vtk.Writer(sp)
    w.WriteSolutions(str(tmp_path), "Frame0")
    points, values = sp.ctx.sample_vtk(False)
    vtk.write_vtu(str(tmp_path / "host.vtu"), points, values)
    assert open(tmp_path / "Frame0" / "0.vtu", "rb").read() == open(tmp_path / "host.vtu", "rb").read()
