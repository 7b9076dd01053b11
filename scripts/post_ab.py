"""A/B timing of the post chain's fused steps on one full-size pair: `python scripts/post_ab.py [D] > out.jsonl`.

Every configuration runs scripts/post_chain_ms.py in its own process (the library reads its measurement knobs once):
STEREO_POST_SERIAL=1 puts both sides on one stream so that the kernel-class times are of kernels running alone,
STEREO_POST_FUSE clears single fusions (see csrc/post_api.cu), STEREO_CBCA_KERNEL=1 takes the CBCA kernel with
per-thread global loads instead of the bulk-copy ring."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
D = sys.argv[1] if len(sys.argv) > 1 else "228"
CONFIGS = [
    ("default (two streams, all fused)", {}),
    ("serial, all fused", {"STEREO_POST_SERIAL": "1"}),
    ("serial, nothing fused (round-1 structure)", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "0"}),
    ("serial, only scale-on-load", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "1"}),
    ("serial, only moments in CBCA", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "2"}),
    ("serial, only WTA in CBCA", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "4"}),
    ("serial, only WTA in SGM", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "8"}),
    ("serial, all fused, CBCA with global loads", {"STEREO_POST_SERIAL": "1", "STEREO_CBCA_KERNEL": "1"}),
    ("serial, nothing fused, CBCA with global loads", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "0", "STEREO_CBCA_KERNEL": "1"}),
    ("serial, all but scale-on-load", {"STEREO_POST_SERIAL": "1", "STEREO_POST_FUSE": "14"}),
    ("two streams, all but scale-on-load", {"STEREO_POST_FUSE": "14"}),
    ("two streams, nothing fused", {"STEREO_POST_FUSE": "0"}),
    ("two streams, all fused, CBCA with global loads", {"STEREO_CBCA_KERNEL": "1"}),
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
