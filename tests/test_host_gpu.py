"""The reference's start-up path with the probe wired in, on a real GPU: device map -> grade ->
ListAndWatch -> GET /probe -> GET /metrics (host/plugin_sim.cpp), including the per-slice child
process (exercised with the GPU's own UUID, since the box has no MIG partitions)."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "k8s-gpu-device-plugin_b200", "host")


def run_sim(*args):
    r = subprocess.run([os.path.join(HOST, "plugin_sim"), *args], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    return json.loads(r.stdout)


def test_plugin_sim_healthy_path():
    d = run_sim("--slice", os.path.join(HOST, "gpuprobe-slice"))
    assert d["devices"] >= 1 and d["unhealthy"] == 0 and d["log"] == []
    # ListAndWatch: first list all Healthy; then exactly one resend per transition (down, up)
    assert len(d["lists"]) == 3
    assert all(x["health"] == "Healthy" for x in d["lists"][0])
    assert d["lists"][1][0]["health"] == "Unhealthy" and d["lists"][2][0]["health"] == "Healthy"
    # GET /probe: reference envelope, one result per device incl. the child-process slice
    assert d["probe"]["code"] == 0 and d["probe"]["msg"] == "success"
    res = d["probe"]["data"]
    assert len(res) == d["devices"] + 1
    for r in res:
        assert r["healthy"] is True and r["error"] == 0 and r["pattern_mismatches"] == 0
        assert r["copy_gbps"] > 100 and r["verify_gbps"] > 100 and r["gemm_tflops"] > 1
    assert any(r["index"] == "0:0" for r in res)  # the slice went through gpuprobe-slice
    assert d["unknown"] == {"code": 404, "data": None, "msg": "unknown device GPU-00000000-0000-0000-0000-000000000000"}
    assert d["version"]["data"] == "version : 0.0.1" and d["health"]["data"] == "ok"
    m = d["metrics"]
    for fam in ("gpu_probe_healthy", "gpu_probe_hbm_copy_gbps", "gpu_probe_hbm_triad_gbps", "gpu_probe_hbm_verify_gbps",
                "gpu_probe_pattern_mismatches", "gpu_probe_gemm_tflops", "gpu_probe_duration_seconds"):
        assert f"# TYPE {fam} gauge" in m


def test_plugin_sim_floor_makes_device_unhealthy():
    d = run_sim("--min-copy", "1000000", "--mask", "1")
    assert d["unhealthy"] == d["devices"]
    assert all(x["health"] == "Unhealthy" for x in d["lists"][0])  # kubelet never sees it Healthy
    assert "failed_mask=1" in d["log"][0]


def test_manager_scenario_fabric_and_isolation():
    """PluginManager on real GPUs: grade + fabric inside loadPlugins, then the isolated re-probe path: a child
    whose device reports a sticky fault, a child that never answers (killed at the deadline), and the way back
    to Healthy each time, with the plugin's own process and context untouched."""
    d = run_sim("--scenario", "manager", "--slice", os.path.join(HOST, "gpuprobe-slice"), "--fabric-mib", "64")
    assert d["devices"] >= 1 and d["unhealthy"] == 0
    steps = {s["step"]: s for s in d["steps"]}
    assert steps["isolated"]["verdict"] == "Healthy" and steps["isolated"]["error"] == 0
    assert steps["child_fault"]["verdict"] == "Unhealthy" and steps["child_fault"]["error"] == -9   # GP_ERR_DEVICE_LOST
    assert steps["after_fault"]["verdict"] == "Healthy"
    assert steps["child_hang"]["verdict"] == "Unhealthy" and steps["child_hang"]["error"] == -9
    assert steps["after_hang"]["verdict"] == "Healthy" and steps["in_process"]["verdict"] == "Healthy"
    assert d["isolated_probes"] == 5
    # kubelet saw: all Healthy, then down/up for the fault, down/up for the hang
    first_id = d["lists"][0][0]["ID"]
    seq = [next(x["health"] for x in l if x["ID"] == first_id) for l in d["lists"]]
    assert seq == ["Healthy", "Unhealthy", "Healthy", "Unhealthy", "Healthy"]
    assert any("did not answer within 1500 ms and was killed" in l for l in d["log"])
    if d["devices"] >= 2:
        f = d["fabric"]
        assert f["ngpus"] == d["devices"] and f["healthy"] == 1 and f["data_ok"] == 1
        assert 100 < f["read_min"] <= f["read_max"] < 1000 and 100 < f["write_min"] <= f["write_max"] < 1000
        assert f["allreduce_error"] in (0, -4)   # NCCL may be absent outside the python process
        assert len(d["preferred"]) == 2
        assert 'gpu_probe_p2p_gbps{dir=\"read\",dst=\"1\",src=\"0\"}' in d["metrics"] or "gpu_probe_p2p_gbps{dir=\"read\"" in d["metrics"]
        assert "gpu_probe_fabric_healthy" in d["metrics"]
