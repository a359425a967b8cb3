"""The parts of ``bench.py``'s contract that do not need a GPU: both arms describe the same
configuration, clock records of several ranks merge conservatively, the final metrics come out of the
logger's store the way the JSON line reports them, and the command line offers the documented arms."""

import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def test_both_arms_print_the_same_config_block():
    from baseline import reference_runner

    ours = bench.bench_config(8, 6750)
    src = open(reference_runner.__file__).read()
    assert "bench_config" in src  # the reference arm takes its block from the same function
    assert ours["nodes"] == 8 and ours["global_batch"] == 8 * bench.BATCH and ours["topology"] == "full"
    assert "flushed between rounds" in ours["l2"]
    assert json.loads(json.dumps(ours)) == ours


def test_clock_records_merge_to_the_worst_case():
    a = {"sm_mhz": 1965.0, "sm_max_mhz": 1965.0, "power_w_max": 200.0, "samples": 3, "reasons": [], "how": "nvml"}
    b = {"sm_mhz": 1800.0, "sm_max_mhz": 1965.0, "power_w_max": 300.0, "samples": 2, "reasons": ["sw_power_cap"], "how": "nvml"}
    m = bench.merge_clocks([a, None, b])
    assert m["sm_mhz"] == 1800.0 and m["sm_max_mhz"] == 1965.0 and m["power_w_max"] == 300.0
    assert m["reasons"] == ["sw_power_cap"] and m["ranks"] == 2 and m["samples"] == 5
    assert bench.merge_clocks([None]) is None and bench.merge_clocks([]) is None


def test_final_metrics_take_each_nodes_last_value():
    logs = {"exp": {"n0": {"test_loss": [(0, 2.0), (3, 0.5)], "test_metric": [(3, 0.9)]},
                    "n1": {"test_loss": [(3, 0.7), (1, 1.5)], "test_metric": [(3, 0.8)]}}}  # fmt: skip
    fm = bench.final_metrics(logs)
    assert abs(fm["test_loss"] - 0.6) < 1e-12 and abs(fm["test_metric"] - 0.85) < 1e-12
    assert fm["round"] == 3 and fm["nodes_reporting"] == 2
    assert bench.final_metrics({}) == {}


def test_command_line_offers_both_arms_and_the_precision_arm():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl", "--dtype", "--eager", "--ref-fast", "--control"):
        assert flag in out.stdout, flag


def test_reference_arm_reports_when_the_reference_is_not_installed(tmp_path, monkeypatch):
    monkeypatch.setattr(bench, "ROOT", str(tmp_path))
    monkeypatch.delenv("RANK", raising=False)
    monkeypatch.delenv("WORLD_SIZE", raising=False)
    res = bench.run_reference(bench.argparse.Namespace(steps=1, warmup=3, ref_fast=False, timeout=10.0, ref_budget=5.0))
    assert res["impl"] == "reference" and "unavailable" in res


def test_home_device_defaults_to_local_rank(monkeypatch):
    """One process per GPU: framework threads select the process's GPU before their first CUDA call;
    the default comes from the launcher's LOCAL_RANK, an explicit setting wins, and without either
    nothing is selected (several GPUs driven from one process, as in the in-process tests)."""
    from p2pfl_b200 import ops

    monkeypatch.delenv("LOCAL_RANK", raising=False)
    ops.set_home_device(None)
    assert ops.home_device() is None
    monkeypatch.setenv("LOCAL_RANK", "3")
    assert ops.home_device() == 3
    ops.set_home_device(5)
    try:
        assert ops.home_device() == 5
    finally:
        ops.set_home_device(None)
    monkeypatch.setenv("LOCAL_RANK", "not-a-number")
    assert ops.home_device() is None
    ops.thread_uses_cuda()  # no device here: must be a harmless no-op


def test_control_plane_harness_runs_a_mixed_line():
    """bench/control_plane_bench.py end to end, small: 6 nodes on a line, two per process (links inside a
    process are direct calls, links between processes sockets), 3 voted trainers, the emulated fused
    exchange and the pipelined round."""
    import math
    import socket

    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    port = 20000 + port % 10000  # below the ephemeral range: the harness uses port .. port + nodes
    out = subprocess.run(
        [sys.executable, os.path.join(ROOT, "bench", "control_plane_bench.py"), "--nodes", "6", "--per-proc", "2", "--topology", "line",
         "--train-set", "3", "--rounds", "6", "--warmup", "2", "--device-ms", "1.0", "--port", str(port), "--stuck-after", "90"],
        capture_output=True, text=True, timeout=400,
    )  # fmt: skip
    assert out.returncode == 0, out.stderr[-2000:]
    assert "STUCK" not in out.stderr and "Traceback" not in out.stderr, out.stderr[-2000:]
    line = json.loads([l for l in out.stdout.splitlines() if l.startswith("{")][-1])
    assert line["nodes"] == 6 and line["per_proc"] == 2 and line["local_shortcut"] is True
    assert math.isfinite(line["ms_per_round"]) and line["ms_per_round"] > 1.0
    assert line["msgs_per_round_node0"].get("in:vote_train_set", 0) > 0
