import os, sys
sys.path.insert(0, "k8s-gpu-device-plugin_b200")
import gpuprobe
p = gpuprobe.Probe(keep_arena=1)
for s in (64 << 20, 1 << 30):
    for d, name in ((0, "read"), (1, "write"), (2, "bidir")):
        m = p.probe_p2p(2, s, d)
        print(os.environ.get("GPUPROBE_P2P_SHAPE", "0"), name, s >> 20, "MiB", round(m[0][1], 1), round(m[1][0], 1), "local", round(m[0][0], 1))
