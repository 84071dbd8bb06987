"""Time the per-stage profile of each library variant given on the command line (run on the GPU box)."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for lib in sys.argv[1:]:
    env = dict(os.environ, GSR_B200_LIB=os.path.join(ROOT, lib))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "100", "--warmup", "5", "--no-cpu-baseline", "--no-loops"],
                       env=env, capture_output=True, text=True)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        print(os.path.basename(lib), "ms/step %.3f" % d["ms_per_step"], {k: round(v["ms"], 3) for k, v in d["stages"].items()}, flush=True)
    except Exception as ex:
        print(os.path.basename(lib), "FAILED", ex, r.stderr[-500:], flush=True)
