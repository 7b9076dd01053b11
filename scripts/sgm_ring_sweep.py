"""Sweep of the SGM ring shapes (stage size, ring budget per kernel): `python scripts/sgm_ring_sweep.py [D]`.
Every line is scripts/post_chain_ms.py with both sides on one stream and STEREO_SGM_RING set."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
D = sys.argv[1] if len(sys.argv) > 1 else "228"
SHAPES = ["32,96,24,72,16,72",      # default
          "32,96,12,72,8,72",       # one step per stage, deeper rings
          "32,96,24,110,16,110",    # two blocks per SM, deeper rings
          "32,96,36,72,24,72",      # three steps per stage
          "32,96,24,54,16,54",      # four blocks per SM
          "16,96,24,72,16,72",      # horizontal: smaller stages, deeper ring
          "32,64,24,72,16,72",      # horizontal: three blocks per SM
          "48,110,24,72,16,72",     # horizontal: larger stages
          "60,200,24,72,16,72"]     # horizontal: one block per SM, long stages
for shape in SHAPES:
    e = dict(os.environ)
    e.update({"STEREO_POST_SERIAL": "1", "STEREO_SGM_RING": shape})
    r = subprocess.run([sys.executable, os.path.join(HERE, "post_chain_ms.py"), D], env=e, capture_output=True, text=True)
    try:
        out = json.loads(r.stdout.strip().splitlines()[-1])
        print(json.dumps({"shape": shape, "sgm_ms": out["kernel_class_ms"]["sgm"], "post_chain_ms": out["post_chain_ms"]}), flush=True)
    except Exception:
        print(json.dumps({"shape": shape, "error": r.stderr[-300:]}), flush=True)
