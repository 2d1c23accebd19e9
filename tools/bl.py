"""Print the frame times of bench lines: python tools/bl.py gpurun_out/x.log [...]"""
import json, sys
for f in sys.argv[1:]:
    d = json.loads(open(f).read().strip().splitlines()[-1])
    print(f, d["ms_per_step"], {k: v["ms_per_step"] for k, v in d.get("other_workloads", {}).items()},
          "gemm talker", d["roofline"]["launch_ms"], "predictor", [r["launch_ms"] for r in d["roofline_kernels"][:2]])
