"""Runs bench.py once per value of one wm_set_int tuning knob and prints views/s, ms per step and one breakdown entry.
    python tools/sweep_knob.py attn_general_variant 4 5 6 7 [--entry attn_local]"""
import json
import os
import subprocess
import sys

root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
entry = "attn_local"
if "--entry" in args:
    i = args.index("--entry")
    entry = args[i + 1]
    del args[i:i + 2]
knob, values = args[0], args[1:]
for v in values:
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--steps", "5", "--warmup", "3", "--no-cpu-baseline",
                        "--set", f"{knob}={v}"], capture_output=True, text=True, cwd=root)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not line:
        print(knob, v, "FAILED", r.stderr[-500:])
        continue
    d = json.loads(line[-1])
    print(f"{knob}={v}: {d['value']:.1f} views/s, {d['ms_per_step']:.2f} ms/step, {entry} {d['breakdown_ms'][entry]}, "
          f"sm {d['clocks']['sm_mhz']} MHz", flush=True)
