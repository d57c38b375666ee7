python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shape.py -x -q 2>&1 | tail -3
bash tools/debug/quick_env.sh
