#!/bin/bash
# Round-2 evidence run on one B200 (one gpurun call): [GPU tests,] smoke, both bench arms, the ncu launch list of the bench command,
# --set full captures of both tendency sweeps and of one substep (summarised ON THE BOX: the reports themselves exceed what
# gpurun_out/ may carry back), the A/B of the two sweeps.  Outputs under gpurun_out/.   SKIP_TESTS=1 skips pytest.
mkdir -p gpurun_out; rm -f gpurun_out/parity_errors.jsonl
if [ -z "$SKIP_TESTS" ]; then
  timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -3 gpurun_out/pytest_gpu.log
fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 300 gpurun_out/bench_n1.json
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-targets > gpurun_out/ncu_launch.log 2>&1
python tools/summarize_launches.py gpurun_out/launches.csv > gpurun_out/launches.md 2>/dev/null
T=/tmp/ncu_r2; mkdir -p $T
timeout 600 ncu --set full --import-source on --clock-control none -k 'regex:tendency_fast|tendency_plane' --launch-skip 5 -c 2 \
    -f -o $T/tendency python tools/profile_tendency.py > gpurun_out/ncu_full.log 2>&1
python tools/ncu_summary.py $T/tendency.ncu-rep > gpurun_out/tendency_ncu_full.md 2>> gpurun_out/ncu_full.log
for k in tendency_fast tendency_plane; do
  ncu -i $T/tendency.ncu-rep --page source --csv --kernel-name regex:$k > $T/src_$k.csv 2>/dev/null
  echo "## $k" >> gpurun_out/tendency_sass_mix.md; python tools/sass_mix.py $T/src_$k.csv 16777216 >> gpurun_out/tendency_sass_mix.md 2>&1
done
timeout 600 ncu --set full --clock-control none -k 'regex:tendency_plane|pf_|pressure_correct_halo' \
    --launch-skip 28 -c 7 -f -o $T/step256 python tools/profile_step.py > gpurun_out/ncu_step.log 2>&1
python tools/ncu_summary.py $T/step256.ncu-rep > gpurun_out/step256_ncu_full.md 2>> gpurun_out/ncu_step.log
ZSEGS="" timeout 300 tools/ab_sweep.sh > /dev/null 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
du -sh gpurun_out
