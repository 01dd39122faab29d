"""Host-side pieces of bench.py that need no GPU: workload naming for the two BASELINE.json configs
it can time, the bounded CPU sample sizes, and the reference arm's JSON contract on a tiny sample."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_workload_names():
    old = bench.WORKLOAD["spp"]
    try:
        bench.WORKLOAD["spp"] = 500
        assert "configs[1]" in bench.WORKLOAD_NAME() and "500spp" in bench.WORKLOAD_NAME()
        bench.WORKLOAD["spp"] = 4096
        assert "configs[4]" in bench.WORKLOAD_NAME() and "4096spp" in bench.WORKLOAD_NAME()
        bench.WORKLOAD["spp"] = 7
        assert "non-standard" in bench.WORKLOAD_NAME()
    finally:
        bench.WORKLOAD["spp"] = old


def test_cpu_sample_is_bounded():
    for cores in (1, 8, 16, 64, 256):
        spp = bench.cpu_sample_spp(cores, 1200, 675)
        assert 1 <= spp <= 32


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "cpu_mt_tracer")),
                    reason="oracle/_ref/cpu_mt_tracer not built")
def test_mt_program_timing_record():
    rec = bench.run_reference_mt(96, 2, target_s=0.2)
    assert rec and rec["unit"] == "Msamples/s" and rec["threads"] == 2 and rec["value"] > 0
    assert "Time Cost" in rec["sample"]


@pytest.mark.skipif(not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "ref_harness")),
                    reason="oracle/_ref/ref_harness not built")
def test_reference_cpu_runner_counts_every_sample():
    samples, seconds, kind = bench.run_reference_cpu(1, 2, 48, 27)
    assert samples == 48 * 27 and seconds > 0 and kind == "reference"
