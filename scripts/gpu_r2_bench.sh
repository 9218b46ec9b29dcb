#!/bin/bash
# Both bench arms at N=1 exactly as the driver runs them (--steps 20 --warmup 5) + our default run.
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "ref rc=$?"
grep -c libsplat gpurun_out/bench_ref_n1.err
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ours_n1.json 2> gpurun_out/bench_ours_n1.err; echo "ours rc=$?"
tail -3 gpurun_out/bench_ours_n1.err
python - <<'P'
import json
for f in ('bench_ref_n1','bench_ours_n1'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
    except Exception as e:
        print(f, 'unparsable', e); continue
    print(f, 'value %.1f ms %.4f'%(d['value'], d['ms_per_step']), 'e2e', d['e2e'] and round(d['e2e']['value'],1), 'clocks', d['clocks'])
    print('   mapping', d.get('mapping'))
    print('   kernels', d.get('kernel_ms_per_step'))
P
