"""A/B timing of library variants on the same box: runs bench.py once per variant (WGBV_LIBRARY) and prints one line each.

usage: python profiles/ab_bench.py [--nodes N] [--steps K] name=path/to/libwgbv_variant.so[@ENV:VALUE...] ... (path 'default': the in-tree build)
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
while args and args[0].startswith("--"):
    extra += args[:2]
    args = args[2:]
if not any(a == "--steps" for a in extra):
    extra += ["--steps", "5"]
for spec in args:
    name, _, path = spec.partition("=")
    path, *sets = path.split("@")
    env = dict(os.environ)
    for kv in sets:
        env[kv.split(":")[0]] = kv.split(":")[1]
    if path and path != "default":
        env["WGBV_LIBRARY"] = os.path.join(ROOT, path) if not os.path.isabs(path) else path
    else:
        env.pop("WGBV_LIBRARY", None)
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--warmup", "3", "--no-e2e", "--no-cpu-baseline", "--no-labels", "--no-transpose", "--no-configs"] + extra
    res = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    try:
        j = json.loads(res.stdout.strip().splitlines()[-1])
        st = {k: round(v, 2) for k, v in j["roofline"].get("stage_ms_rank0", {}).items()}
        print(f"{name:12s} {j['ms_per_step']:8.3f} ms/step verified={j.get('verified')} launches={j.get('gpu_launches')} {st}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"{name:12s} FAILED rc={res.returncode} {e}\n{res.stderr[-2000:]}", flush=True)
