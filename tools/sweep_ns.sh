for ns in 0 2 3 5 6 8; do
  if [ $ns = 0 ]; then unset B2TTS_FA_NSPLIT; else export B2TTS_FA_NSPLIT=$ns; fi
  python bench.py --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --steps 128 --warmup 8 --also-streams "" > gpurun_out/ns_$ns.log 2>&1
  python - <<P
import json
d=json.loads(open("gpurun_out/ns_$ns.log").read().strip().splitlines()[-1])
print("nsplit $ns", d["ms_per_step"], [(r["frac"], r["launch_ms"]) for r in d["roofline_kernels"][2:5]])
P
done
