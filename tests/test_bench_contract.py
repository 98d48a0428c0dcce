"""The bench line's contract (keys the driver and the judge read), checked on the lines committed under profiles/ — the
same code path printed them on the B200 box; no GPU needed here."""
import json
from pathlib import Path

import pytest

PROFILES = Path(__file__).resolve().parent.parent / "profiles"
LINES = ["r1_bench_full_1gpu.json", "r1_scale_spmd_8.json", "r1_bench_spmd_2gpu_peer_stores.json"]


@pytest.mark.parametrize("name", LINES)
def test_committed_bench_lines_carry_the_contract_keys(name):
    text = (PROFILES / name).read_text().strip()
    assert len(text.splitlines()) == 1                       # ONE JSON line
    d = json.loads(text)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["metric"] == "mandelbrot_mpix_per_s" and d["unit"] == "Mpix/s" and d["vs_baseline"] is None
    assert d["warmup"] >= 3 and d["gpu_launches"] > 0 and "workload" in d["config"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"])
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(d["roofline"])
    assert abs(d["roofline"]["frac"] - d["roofline"]["achieved"] / d["roofline"]["peak"]) < 1e-9
    assert {"value", "unit", "cores", "kind", "sample"} <= set(d["cpu_baseline"]) and d["cpu_baseline"]["kind"] == "port"
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert d["parity"]["mismatches"] == 0


def test_reference_arm_line():
    d = json.loads((PROFILES / "r1_bench_reference_arm.json").read_text().strip())
    assert d["impl"] == "reference" and d["metric"] == "mandelbrot_mpix_per_s"
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["cpu_baseline"]["value"] == d["value"]
