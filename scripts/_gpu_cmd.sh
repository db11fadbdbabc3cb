mkdir -p gpurun_out/pipe10
(timeout 900 python -m pytest tests/test_gpu_parity.py -m "gpu and not slow" -x -q -k "pipeline or path_mode or full_default or cfg3 or dropin" 2>&1 | tail -15) > gpurun_out/pipe10/pytest.log; tail -4 gpurun_out/pipe10/pytest.log
for wl in cfg1_default_4032_waters cfg2_8000_waters_intracellular_kick; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/pipe10/bench_${wl}.json 2>> gpurun_out/pipe10/bench.err
  python -c "import json,sys; d=json.load(open('gpurun_out/pipe10/bench_${wl}.json')); print('$wl',d['value'],d['e2e']['value'],d['ms_per_step'],d['config'].get('pipeline_sprints'),d['gpu_launches'])"
done
timeout 600 python bench.py --workload cfg1_default_4032_waters --walkers-per-gpu 32768 --path-mode pipeline --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/pipe10/bench_n32768.json 2>> gpurun_out/pipe10/bench.err
python -c "import json,sys; d=json.load(open('gpurun_out/pipe10/bench_n32768.json')); print('n=32768 pipeline',d['value'],d['ms_per_step'])"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/pipe10/launches_cfg1_auto.csv python bench.py --workload cfg1_default_4032_waters --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/pipe10/ncu_cfg1.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/pipe10/launches_cfg1_auto.csv')))
for i,r in enumerate(rows):
    if r and r[0]=='ID': h=i;break
hdr=rows[h]; ix={k:j for j,k in enumerate(hdr)}
agg=collections.OrderedDict()
for r in rows[h+1:]:
    if len(r)<len(hdr): continue
    name=r[ix['Kernel Name']].split('(')[0].split('::')[-1][:40]
    a=agg.setdefault(name,[0,0.0]); a[0]+=1; a[1]+=float(r[ix['Metric Value']])
for k,v in agg.items(): print(f"{k:42s} n={v[0]:4d} total={v[1]/1e6:9.3f} ms")
PY
