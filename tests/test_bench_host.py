"""Host-side pieces of bench.py that must not take the bench line down on a box without NVML or
nvidia-smi (this container): the clock sampler degrades to "unavailable", in every mode."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


@pytest.mark.parametrize("mode", ["nvml", "smi", "off"])
def test_clock_sampler_degrades_without_a_gpu(monkeypatch, mode):
    import bench

    monkeypatch.setenv("NFB_BENCH_CLOCKS", mode)
    cs = bench.ClockSampler(0)
    cs.start()
    assert cs.ready() or cs.how is not None
    out = cs.stop()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    if out["sm_mhz"] is None:  # no device here: said so, not raised
        assert out["reasons"] == ["clock sampler unavailable"] or out.get("samples", 0) == 0


def test_reason_bits_match_nvml_header():
    """nvml.h nvmlClocksEventReasons: the bits the sampler turns into the recipe's reason names."""
    import bench

    pynvml = pytest.importorskip("pynvml")
    B = bench.ClockSampler.BITS
    assert B["sw_power_cap"] == pynvml.nvmlClocksThrottleReasonSwPowerCap
    assert B["hw_slowdown"] == pynvml.nvmlClocksThrottleReasonHwSlowdown
    assert B["sw_thermal_slowdown"] == pynvml.nvmlClocksThrottleReasonSwThermalSlowdown
    assert B["hw_thermal_slowdown"] == pynvml.nvmlClocksThrottleReasonHwThermalSlowdown
    assert B["hw_power_brake_slowdown"] == pynvml.nvmlClocksThrottleReasonHwPowerBrakeSlowdown
