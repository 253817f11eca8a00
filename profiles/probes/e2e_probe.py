"""e2e leg of bench.py against the number of shards and the zero-copy variant (host buffers in / out)."""
import json
import subprocess
import sys

for g in (2, 4, 8, 16):
    for extra in ([], ["--e2e-zero-copy"]):
        r = subprocess.run([sys.executable, "bench.py", "--no-cpu", "--no-agents", "--steps", "600", "--e2e-groups", str(g)] + extra,
                           capture_output=True, text=True)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            print(g, extra, round(d["e2e"]["value"] / 1e6, 2), "M env-steps/s", d["finite"], flush=True)
        except Exception:
            print(g, extra, "fail", r.stderr[-300:], flush=True)
