"""Condenses an ncu launch list (--csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum)
of a bench.py run into per-kernel totals of ONE training step and the per-category DRAM traffic bench.py reads:
python tools/ncu_summary.py launches.csv profiles/rNN_ncu_launches_step.csv profiles/rNN_dram_traffic_per_step.json
Training steps are counted by the k_adam launches (one per step); launches outside the steps (set-up) are kept in
the totals only if they repeat every step."""
import csv, json, re, sys, collections

CATEGORY = [('k_gemm_tc<0', 'gemm_fwd'), ('k_splitk', 'gemm_fwd'), ('k_gemm_tc<1', 'gemm_dgrad'), ('k_gemm_tc<2', 'gemm_wgrad'), ('k_conv_fwd', 'gemm_fwd'), ('k_dense_k1', 'gemm_fwd'),
            ('k_conv_dgrad', 'gemm_dgrad'), ('k_conv_wgrad', 'gemm_wgrad'), ('k_act_bwd', 'act_bwd'), ('k_bn', 'bn'),
            ('k_maxpool', 'pool'), ('k_copy', 'concat'), ('k_adam', 'adam'), ('k_loss', 'loss'), ('k_im2col', 'gather'),
            ('k_gather', 'gather')]


def main(src, out_csv, out_json):
    rows = [r for r in csv.reader(open(src, newline='')) if len(r) > 8]
    hdr = rows[0]
    iname, imetric, ival, iunit, iid = (hdr.index(k) for k in ('Kernel Name', 'Metric Name', 'Metric Value', 'Metric Unit', 'ID'))
    per = collections.OrderedDict()
    for r in rows[1:]:
        name = re.sub(r'^void (srn::)?', '', r[iname])
        name = re.sub(r'\(.*$', '', name).replace('(int)', '').replace('(bool)', '')
        d = per.setdefault(name, {'ids': set(), 'ms': 0.0, 'bytes': 0.0})
        d['ids'].add(r[iid])
        v = float(r[ival].replace(',', ''))
        if r[imetric] == 'gpu__time_duration.sum':
            d['ms'] += v * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}[r[iunit]]
        elif r[imetric].startswith('dram__bytes'):
            d['bytes'] += v * {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[r[iunit]]
    steps = max(1, len(per.get('k_adam', {'ids': ()})['ids']))
    total = sum(d['ms'] for d in per.values()) / steps
    with open(out_csv, 'w') as f:
        f.write(f'# one training step = totals over {steps} steps / {steps}; {sum(len(d["ids"]) for d in per.values()) // steps} launches, '
                f'{total:.2f} ms summed kernel time (cold-cache, serialised: compare shares)\n')
        f.write('kernel,launches,total_ms,share,dram_MB\n')
        for name, d in sorted(per.items(), key=lambda kv: -kv[1]['ms']):
            f.write(f'{name},{len(d["ids"]) // steps},{d["ms"] / steps:.3f},{d["ms"] / steps / total:.4f},{d["bytes"] / steps / 1e6:.1f}\n')
    traffic = collections.OrderedDict()
    for name, d in per.items():
        for prefix, cat in CATEGORY:
            if name.startswith(prefix):
                traffic[cat] = traffic.get(cat, 0.0) + d['bytes'] / steps
                break
    json.dump(traffic, open(out_json, 'w'), indent=1)
    print(open(out_csv).read())
    print(json.dumps(traffic))


if __name__ == '__main__':
    main(*sys.argv[1:4])
