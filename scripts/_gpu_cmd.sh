mkdir -p gpurun_out/san
python scripts/sanitize_case.py > gpurun_out/san/plain.log 2>&1; echo "plain exit $?"; tail -3 gpurun_out/san/plain.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python scripts/sanitize_case.py > gpurun_out/san/memcheck.log 2>&1; echo "memcheck exit $?"; tail -4 gpurun_out/san/memcheck.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 7 python scripts/sanitize_case.py > gpurun_out/san/racecheck.log 2>&1; echo "racecheck exit $?"; tail -4 gpurun_out/san/racecheck.log
timeout 900 compute-sanitizer --tool synccheck --error-exitcode 7 python scripts/sanitize_case.py > gpurun_out/san/synccheck.log 2>&1; echo "synccheck exit $?"; tail -3 gpurun_out/san/synccheck.log
