#!/bin/bash
# One gpurun call: GPU parity tests, the default bench line, the 512^3 target size, the reference arm, the ncu launch
# list of the bench command, --set full captures of one substep and of both tendency builds, timings of the widened
# (LES) configurations.  Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; tail -c 400 gpurun_out/bench_n1.json
timeout 600 python bench.py --size 512 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_512.json 2> gpurun_out/bench_512.err
timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
if [ -z "$NO_NCU_FULL" ]; then
timeout 600 ncu --set full --import-source on --clock-control none -k regex:tendency_fast --launch-skip 5 -c 2 \
    -f -o gpurun_out/tendency python tools/profile_tendency.py > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --import-source on --clock-control none -k 'regex:tendency_fast|pf_|pressure_correct_halo' \
    --launch-skip 28 -c 7 -f -o gpurun_out/step256 python tools/profile_step.py > gpurun_out/ncu_step.log 2>&1
fi
for cfg in ppb les "les vitd" smag amd; do echo "$cfg: $(python tools/time_phases.py 256 5 $cfg 2>&1 | tail -1)"; done > gpurun_out/les_times.txt
echo "c4 512x512x256 stretched: $(python tools/time_phases.py 512 3 c4 2>&1 | tail -1)" >> gpurun_out/les_times.txt
python tools/time_solvers.py > gpurun_out/solver_times.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
