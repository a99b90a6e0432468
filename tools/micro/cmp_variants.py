import os, sys, json, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for v in sys.argv[2:]:
    env = dict(os.environ)
    if v != "main": env["SGB_LIB"] = os.path.join(ROOT, "signals_b200", "libsignals_b200_%s.so" % v)
    for wl in sys.argv[1].split(","):
        for rep in range(2):
            out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gpu_one.py"), wl, "split", "10"], env=env, capture_output=True, text=True).stdout
            r = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
            print(v, wl, "step %.4f prep %.4f emit %.4f recur %.4f" % (r["ms_per_step"], r["prep_ms"], r["emit_ms"], r["decode_ms"]), r["checksum"])
