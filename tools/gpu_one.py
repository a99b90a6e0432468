"""One workload under one schedule, a few device-resident decode calls (the command profiled by ncu on the GPU box).
    python tools/gpu_one.py <workload> <split|fused> [steps]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ["SGB_SCHED"] = sys.argv[2] if len(sys.argv) > 2 else "split"
import tools.gpu_sched_compare as gc
print(json.dumps(gc.run(sys.argv[1] if len(sys.argv) > 1 else "target", os.environ["SGB_SCHED"], int(sys.argv[3]) if len(sys.argv) > 3 else 2)))
