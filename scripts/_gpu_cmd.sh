mkdir -p gpurun_out/v13
(timeout 900 python -m pytest tests/test_octree_gpu.py -m gpu -x -q 2>&1 | tail -6) > gpurun_out/v13/pytest_octree.log; tail -3 gpurun_out/v13/pytest_octree.log
timeout 900 python scripts/octree_bench.py default dense2000 --no-host dense8000 > gpurun_out/v13/octree_bench.jsonl 2> gpurun_out/v13/octree_bench.err; cut -c1-700 gpurun_out/v13/octree_bench.jsonl; tail -3 gpurun_out/v13/octree_bench.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v13/octree_launches.csv python scripts/octree_profile.py > gpurun_out/v13/ncu_launches.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/v13/octree_launches.csv')))
for i,r in enumerate(rows):
    if r and r[0]=='ID': h=i;break
hdr=rows[h]; ix={k:j for j,k in enumerate(hdr)}
for r in rows[h+1:]:
    if len(r)<len(hdr): continue
    name=r[ix['Kernel Name']].split('(')[0].split('::')[-1][:30]
    if 'criterion' in name or 'leaf_field' in name: print(name, r[ix['Grid Size']], round(float(r[ix['Metric Value']])/1e6,3),'ms')
PY
