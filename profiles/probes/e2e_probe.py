import subprocess, json, sys
for g in (2, 4, 8):
    for extra in ([], ["--e2e-zero-copy"]):
        r = subprocess.run([sys.executable, "bench.py", "--no-cpu", "--steps", "600", "--e2e-groups", str(g)] + extra, capture_output=True, text=True)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            print(g, extra, d["e2e"]["value"], d["finite"], flush=True)
        except Exception as e:
            print(g, extra, "fail", r.stderr[-400:], flush=True)
