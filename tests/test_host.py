"""Host-side logic without a GPU: runs the C++ self-test of the reference-interface mirror
(Device/Devices/BuildDevice/DeviceMap, grading dispatch, ListAndWatch feed, metrics text,
HTTP envelopes) with a scripted probe backend."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "k8s-gpu-device-plugin_b200", "host")


def test_host_selftest():
    exe = os.path.join(HOST, "host_selftest")
    if not os.path.exists(exe):
        subprocess.check_call(["make", "-C", ROOT, "host"])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    assert "host_selftest ok" in r.stdout


def test_go_glue_is_present_and_marked_uncompiled():
    """The Go files cannot be compiled here (no toolchain); they must say so and must bind
    only symbols that include/gpuprobe.h declares."""
    import re
    hdr = open(os.path.join(ROOT, "include", "gpuprobe.h")).read()
    declared = set(re.findall(r"\b(gp_[a-z0-9_]+)\s*\(", re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)))
    godir = os.path.join(ROOT, "k8s-gpu-device-plugin_b200", "go")
    files = [f for f in os.listdir(godir) if f.endswith(".go")]
    assert files
    for f in files:
        src = open(os.path.join(godir, f)).read()
        assert "NOT COMPILED" in src, f
        for sym in set(re.findall(r"C\.(gp_[a-z0-9_]+)\(", src)):
            assert sym in declared, (f, sym)
