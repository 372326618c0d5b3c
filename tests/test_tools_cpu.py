"""CPU: the profile tooling reproduces the committed capture summary from the committed ncu launch list, so the
`roofline.traffic` that bench.py reports can be traced back to the raw per-launch numbers."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dram_capture_summary_matches_its_launch_list(tmp_path):
    src = os.path.join(ROOT, "profiles", "r02_dram_config3_n1.csv")
    out = str(tmp_path / "capture")
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dram_capture.py"), src, out, "config3, N=1"], check=True,
                   capture_output=True)
    got = json.load(open(out + ".json"))
    want = json.load(open(os.path.join(ROOT, "profiles", "r02_dram_config3_n1.json")))
    for key in ("phase_bytes", "phase_ms_serialised", "kernels", "total_ms_serialised", "total_dram_bytes"):
        assert got[key] == want[key], key
    # every launch of a step belongs to a phase, and the phases are the ones bench.py reports
    sys.path.insert(0, ROOT)
    import bench
    assert set(got["phase_bytes"]) == set(bench.PHASES)


def test_bench_reads_the_capture_of_its_workload():
    """bench.py takes `roofline.traffic` from profiles/r02_dram_<workload>_n<N>.json: the file it would read for the
    default workload exists and names the dominant phase."""
    cap = json.load(open(os.path.join(ROOT, "profiles", "r02_dram_config3_n1.json")))
    assert cap["phase_bytes"]["sort"] > 0 and "ncu" in cap["source"]
