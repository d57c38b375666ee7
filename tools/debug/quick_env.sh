# device-timed step under a few environment switches (A/B experiments)
for e in "SRN_X=0" "SRN_SLAB=1" "SRN_SPLIT_TSTAR=100000" "SRN_TMA_MASK=0"; do
  env $e python bench.py --no-e2e --no-cpu --no-pop1000 --steps 10 --warmup 3 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); c=d['roofline']['categories']
        print('$e', 'ms/step %.3f'%d['ms_per_step'], 'serial %.2f'%d['roofline']['step']['serialised_kernel_ms'], ' '.join('%s %.3f'%(k,c[k]['ms_per_step']) for k in ('gemm_fwd','gemm_dgrad','gemm_wgrad')))
"
done
