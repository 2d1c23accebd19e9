B="python bench.py --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --steps 64 --warmup 8"
python -m pytest tests/test_gpu_gemm.py -x -q -m gpu 2>&1 | tail -1
B2TTS_GEMM_PLAIN_RULE=0 $B --streams 64 --also-streams 96,128,192,256,384,512 > gpurun_out/r_off.log 2>&1; python tools/bl.py gpurun_out/r_off.log
$B --streams 64 --also-streams 96,128,192,256,384,512 > gpurun_out/r_on.log 2>&1; python tools/bl.py gpurun_out/r_on.log
