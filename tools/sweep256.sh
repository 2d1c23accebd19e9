B="python bench.py --no-e2e --no-cpu-baseline --no-icl --no-reference-kernels --streams 256 --steps 64 --warmup 8 --also-streams 512"
run() { name=$1; shift; env "$@" $B > gpurun_out/s_$name.log 2>&1; echo -n "$name: "; python tools/bl.py gpurun_out/s_$name.log; }
run base A=1
run ns2 B2TTS_FA_NSPLIT=2
run ns4 B2TTS_FA_NSPLIT=4
run ch32 B2TTS_GEMM_CHUNK_ROWS=32
run ch128 B2TTS_GEMM_CHUNK_ROWS=128
run ch48 B2TTS_GEMM_CHUNK_ROWS=48
