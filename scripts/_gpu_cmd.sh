mkdir -p gpurun_out/v8c
python scripts/profile_case.py 1048576 1000 1 0 | tail -2
for r in 10 30 125; do
RESORT=$r ncu --metrics smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,gpu__time_duration.sum --clock-control none -k regex:walker_sprint -s 3 -c 2 --csv --log-file gpurun_out/v8c/lanes_$r.csv python scripts/profile_case.py 1048576 1000 0 0 > /dev/null 2>&1
echo "== resort $r"; tail -6 gpurun_out/v8c/lanes_$r.csv | cut -d, -f 5,12-
done
