# ncu evidence of the final build (run under gpurun AFTER the plain command exited 0): launch list of a short bench,
# one --set full capture of the dominant kernel (gemm_stream_kernel) and of the decode attention kernel.
set -x
BARGS="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --also-streams \"\""
python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --also-streams "" > gpurun_out/r2l_plain.log 2>&1 || exit 1
tail -c 300 gpurun_out/r2l_plain.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 6000 -c 2400 --csv --log-file gpurun_out/r2l_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --also-streams "" > gpurun_out/r2l_ncu1.log 2>&1
python tools/summarize_launches.py gpurun_out/r2l_launches.csv > gpurun_out/r2l_launches_summary.txt; head -14 gpurun_out/r2l_launches_summary.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_stream_kernel -s 2000 -c 4 -o gpurun_out/r2l_gemm_full -f \
  python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --also-streams "" > gpurun_out/r2l_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep
