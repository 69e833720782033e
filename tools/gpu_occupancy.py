"""Occupancy sweep (experiment): times one kernel variant with RT_EXTRA_SMEM throttling the resident CTAs/SM."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
name = sys.argv[1]
for extra in [int(x) for x in sys.argv[2:]]:
    env = dict(os.environ, RT_EXTRA_SMEM=str(extra))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gpu_variants.py"), name], env=env, capture_output=True, text=True)
    print("extra_smem", extra, r.stdout.strip()[:160] or r.stderr[-300:], flush=True)
