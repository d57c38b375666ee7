python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shape.py -x -q 2>&1 | tail -3
SRN_DUMP_FWD=1 python tools/timeline.py 100 2>&1 | grep "fwd launch" | sort -u
bash tools/debug/quick_env.sh
