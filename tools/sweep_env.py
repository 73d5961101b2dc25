#!/usr/bin/env python3
"""Development aid (GPU box): run bench.py under several environment / flag variations and print one summary line each.

    python tools/sweep_env.py [--config 2] [--math parity] "NAME=VAL,NAME=VAL" "NAME=VAL" ...      ("" = defaults)
"""
import argparse, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("--config", default="2")
ap.add_argument("--math", default="parity")
ap.add_argument("--steps", default="10")
ap.add_argument("variants", nargs="*")
a = ap.parse_args()
for v in (a.variants or [""]):
    env = dict(os.environ)
    for kv in filter(None, v.split(",")):
        k, _, val = kv.partition("=")
        env[k] = val
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", a.config, "--math", a.math, "--steps", a.steps, "--no-strong", "--no-cpu"],
                       env=env, capture_output=True, text=True)
    try:
        d = json.loads([l for l in p.stdout.splitlines() if l.startswith("{")][-1])
        km = d["kernel_ms"]
        print(f"{v or 'default':40s} value {d['value']:8.1f} ms/step {d['ms_per_step']:7.3f} single {d['single_frame_ms']:7.3f} e2e {d['e2e']['value']:8.1f} "
              f"frac {d['roofline']['frac']:.4f} step-frac {d['roofline']['frac_of_timed_step']:.4f} | " + " ".join(f"{k}={x:.3f}" for k, x in km.items()), flush=True)
    except Exception as e:
        print(f"{v or 'default':40s} FAILED rc={p.returncode} {e}\n{p.stderr[-1500:]}", flush=True)
