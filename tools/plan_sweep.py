"""Experiment driver: frame time at a stream count under B2TTS_GEMM_PLAN overrides, one projection shape at a time.
    python tools/plan_sweep.py <streams> [base plan]"""
import json, os, subprocess, sys
streams = int(sys.argv[1])
base = sys.argv[2] if len(sys.argv) > 2 else ""
SHAPES = {"p_qkv": (4096, 1024), "p_o": (1024, 2048), "p_gu": (6144, 1024), "p_down": (1024, 3072),
          "t_qkv": (4096, 2048), "t_o": (2048, 2048), "t_gu": (12288, 2048), "t_down": (2048, 6144)}
OPTS = [(32, 1), (48, 1), (64, 1), (64, 2), (64, 4), (96, 1), (128, 1), (128, 2), (128, 4)]
if streams <= 64:
    OPTS = [(16, 1), (32, 1), (32, 2), (32, 4), (48, 1), (48, 2), (64, 1), (64, 2), (64, 4), (64, 8)]
if os.environ.get("SWEEP_OPTS"):
    OPTS = [tuple(int(v) for v in o.split(",")) for o in os.environ["SWEEP_OPTS"].split(";")]
only = os.environ.get("SWEEP_SHAPES")
def run(plan):
    env = dict(os.environ)
    if plan: env["B2TTS_GEMM_PLAN"] = plan
    out = subprocess.run([sys.executable, "bench.py", "--no-e2e", "--no-cpu-baseline", "--no-icl", "--no-reference-kernels",
                          "--streams", str(streams), "--steps", "48", "--warmup", "6", "--also-streams", ""],
                         env=env, capture_output=True, text=True)
    try:
        return json.loads(out.stdout.strip().splitlines()[-1])["ms_per_step"]
    except Exception:
        return float("nan")
ref = run(base)
print(f"streams {streams} base plan '{base}': {ref:.4f} ms", flush=True)
for name, (n, k) in SHAPES.items():
    if only and name not in only.split(","): continue
    res = []
    for ch, sp in OPTS:
        plan = f"{n},{k},{streams},{streams},{ch},{sp}" + (";" + base if base else "")
        res.append((run(plan), ch, sp))
    best = min(res)
    print(name, " ".join(f"({ch},{sp}):{t:.3f}" for t, ch, sp in res), f" best ({best[1]},{best[2]}) {best[0] - ref:+.3f}", flush=True)
