# quick check of a kernel change on the GPU box: parity tests, then the device-timed step with per-category serialised times
python -m pytest tests/test_gpu_parity.py tests/test_gpu_bench_shape.py -x -q 2>&1 | tail -3
python bench.py --no-e2e --no-cpu --no-pop1000 --steps 20 --warmup 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['roofline']['categories']
        print('ms/step %.3f'%d['ms_per_step'], 'serial %.2f'%d['roofline']['step']['serialised_kernel_ms'], ' '.join('%s %.3f'%(k,c[k]['ms_per_step']) for k in c))
"
