"""A/B timing of the post chain's fused steps on one full-size pair: `python scripts/post_ab.py [D] > out.jsonl`.

Every configuration runs scripts/post_chain_ms.py in its own process (the library reads its measurement knobs once):
STEREO_POST_SERIAL=1 puts both sides on one stream so that the kernel-class times are of kernels running alone,
STEREO_CBCA_KERNEL selects the form of the fused CBCA kernel (csrc/cbca_fused.cu).  (The fusion knobs this script
was written for - STEREO_POST_FUSE, the bulk-copy ring CBCA - were measured and removed with the code they
switched: profiles/r02_post_chain_ab.md.)"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
D = sys.argv[1] if len(sys.argv) > 1 else "228"
CONFIGS = [
    ("default (two streams)", {}),
    ("serial", {"STEREO_POST_SERIAL": "1"}),
    ("serial, CBCA round-1 form", {"STEREO_POST_SERIAL": "1", "STEREO_CBCA_KERNEL": "1"}),
    ("two streams, CBCA round-1 form", {"STEREO_CBCA_KERNEL": "1"}),
]
for name, env in CONFIGS:
    e = dict(os.environ)
    e.update(env)
    r = subprocess.run([sys.executable, os.path.join(HERE, "post_chain_ms.py"), D], env=e, capture_output=True, text=True)
    try:
        out = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        out = {"error": r.stderr[-500:]}
    out["config"] = name
    out["env"] = env
    print(json.dumps(out), flush=True)
