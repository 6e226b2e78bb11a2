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
    # NCCL prints its version banner on stdout when NCCL_DEBUG is set in the environment: take the JSON line
    doc = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(doc) == 1, r.stdout[:2000]
    return json.loads(doc[0])


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
    # --assume-idle: this pytest process holds a CUDA context on GPU 0, which NVML rightly reports as a tenant
    d = run_sim("--scenario", "manager", "--slice", os.path.join(HOST, "gpuprobe-slice"), "--fabric-mib", "64",
                "--assume-idle", "1")
    assert d["devices"] >= 1 and d["unhealthy"] == 0 and d["busy_skipped"] == 0
    assert "gpu_probe_hbm_copy_gbps" in d["metrics"]   # the devices really were graded at load (not skipped as busy)
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


def test_slice_binary_at_1g_slice_scale():
    """No MIG on this pool, so the child is run exactly as ProbeSlice would run it for a 1g.23gb slice of a
    B200 — arrays of SliceBytes(16 GiB, 23 GiB) = a quarter of the slice, grids sized for the slice's 18 SMs —
    on the whole GPU addressed by its own UUID (CUDA_VISIBLE_DEVICES takes it exactly as it takes a MIG-
    UUID).  What this covers: the sizing logic, the launchers at slice scale, the verdict and the POD
    result over the pipe.  What it cannot: a real compute instance (DESIGN.md says so)."""
    import ctypes
    import sys
    sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
    import gpuprobe
    with gpuprobe.Probe() as p:
        uuid = p.device_uuid(0)
    slice_mem = 23 << 30
    hbm = min(16 << 30, slice_mem // 4) & ~((1 << 20) - 1)          # GpuProbeBackend::SliceBytes
    assert hbm == 6174015488 & ~((1 << 20) - 1)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=uuid)
    r = subprocess.run([os.path.join(HOST, "gpuprobe-slice"), "15", str(hbm), str(hbm), "4", "2048", "3", "1",
                        "--sm-limit", "18"], capture_output=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr
    assert len(r.stdout) == ctypes.sizeof(gpuprobe.DeviceResult)
    res = gpuprobe.DeviceResult.from_buffer_copy(r.stdout)
    assert res.struct_size == ctypes.sizeof(gpuprobe.DeviceResult) and res.healthy == 1 and res.error == 0
    assert res.probed_mask == gpuprobe.GP_PROBE_ALL_SINGLE and res.sm_count == 18
    assert res.hbm_bytes == hbm and res.pattern_bytes == hbm and res.pattern_passes == 6
    assert res.arena_bytes <= 3 * hbm + (1 << 30) < slice_mem                    # fits the slice with room to spare
    assert res.gemm.samples_bad == 0 and 1 < res.gemm.tflops_best < 400          # 9 CTA pairs, not 74
    assert res.copy_gbps_median > 1000 and res.verify_gbps_median > 1000


def test_manager_leaves_a_gpu_with_a_tenant_alone():
    """The same start-up path WITHOUT --assume-idle: this pytest process holds a CUDA context on GPU 0, so NVML
    reports a tenant there.  loadPlugins must not give that GPU the full probe (it keeps Healthy, ungraded) and
    must not run the fabric probe, which would drive every GPU at once."""
    import sys
    import torch
    torch.zeros(1, device="cuda:0")   # make sure the tenant exists
    sys.path.insert(0, os.path.join(ROOT, "k8s-gpu-device-plugin_b200"))
    import gpuprobe
    with gpuprobe.Probe() as p:
        if not p.nvml_snapshot(p.device_uuid(0)).available:
            pytest.skip("NVML not loadable on this box: busy-ness cannot be told")
    d = run_sim("--scenario", "manager", "--fabric-mib", "64")
    assert d["devices"] >= 1 and d["unhealthy"] == 0 and d["busy_skipped"] >= 1
    assert d["fabric"]["ngpus"] == 0
    assert all(x["health"] == "Healthy" for x in d["lists"][0])
