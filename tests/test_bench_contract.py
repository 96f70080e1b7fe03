"""bench.py's contract as far as it can be checked without a GPU: the algorithmic bytes per configuration, the labels,
and the JSON line of the reference arm (oracle/_ref on the host cores)."""
import copy
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _args(**over):
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        a = bench.parse()
    finally:
        sys.argv = argv
    for k, v in over.items():
        setattr(a, k, v)
    return a


def test_algorithmic_bytes_follow_the_survey_figure():
    b, parts, dof = bench.algorithmic_bytes(_args())
    assert dof == 135 and b == pytest.approx(24 + 192 / 135)
    assert sum(parts.values()) == pytest.approx(b)
    b3, _, dof3 = bench.algorithmic_bytes(_args(degree=3))
    assert dof3 == 320 and b3 == pytest.approx(24 + 192 / 320)
    bv, pv, _ = bench.algorithmic_bytes(_args(viscous=0.01))
    assert bv == pytest.approx((3 * 1080 + 2 * 8 * (15 * 27 + 15 * 54) + 192) / 135)   # state + traces written and read once
    assert bench.algorithmic_bytes(_args(limiter=True))[0] > b and bench.algorithmic_bytes(_args(energy_viscosity=True))[0] > bv


def test_labels_name_the_configuration_that_runs():
    a = _args()
    assert "p=2 Euler" in bench.metric_name(a) and "p=3" in bench.metric_name(_args(degree=3))
    assert "Navier-Stokes" in bench.metric_name(_args(viscous=0.01))
    assert bench.stage_kernel_name(a).startswith("mcfd::NodeStageKernel") and "affine" in bench.stage_kernel_name(a)
    assert bench.stage_kernel_name(_args(warp=0.1)).startswith("mcfd::LineStageKernel")
    assert bench.stage_kernel_name(_args(stage_kernel="lines")).startswith("mcfd::LineStageKernel")
    assert bench.stage_kernel_name(_args(degree=3)).startswith("mcfd::StageKernel<4")
    assert bench.stage_kernel_name(_args(scheme="dgm")).startswith("mcfd::DgmStageKernel")
    # DRAM traffic from the committed capture is reported for the profiled configuration only
    shape = json.load(open(os.path.join(ROOT, "profiles", "stage_kernel_traffic.json")))["shape"]
    assert bench.traffic_shape(a, 1) == shape
    for other in (_args(degree=3), _args(viscous=0.01), _args(warp=0.1), _args(cells=32), _args(stage_kernel="lines")):
        assert bench.traffic_shape(other, 1) != shape
    assert bench.traffic_shape(a, 2) != shape
    assert "config 5" in bench.config_dict(_args(degree=3, cells=32), 1, (32, 32, 32), "1x1x1")["workload"]
    assert [n for n, _ in bench.FEATURES] == ["line_kernel", "viscous", "limiter", "energy_viscosity", "dgm", "config5_p3"]


def test_reference_arm_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--cpu-cells", "12"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["unit"] == "DOF-updates/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["gpu_launches"] == 0 and line["dtype"] == "f64"
    assert line["e2e"] == {"value": line["value"], "unit": "DOF-updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert line["cpu_baseline"]["value"] == line["value"] and line["cpu_baseline"]["kind"] in ("reference", "port")
    assert "sample_of" in line["config"] and "12^3" in line["config"]["workload"]
