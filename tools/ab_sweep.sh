#!/bin/bash
# A/B of the two fused tendency sweeps (plane-marching TMA kernel vs tile kernel), both arithmetic modes: phase times per substep
# from nh_time_step_profiled at 256^3 and 512^3.  Run under gpurun; output in gpurun_out/ab_sweep.txt
mkdir -p gpurun_out
out=gpurun_out/ab_sweep.txt; : > $out
for N in 256 512; do
  for mode in "" fma; do
    echo "N=$N ${mode:-exact} plane (zseg auto): $(python tools/time_phases.py $N 5 $mode 2>&1 | tail -1)" >> $out
    for z in $ZSEGS; do echo "N=$N ${mode:-exact} plane zseg=$z: $(NHSTEP_PLANE_ZSEG=$z python tools/time_phases.py $N 5 $mode 2>&1 | tail -1)" >> $out; done
    echo "N=$N ${mode:-exact} tile: $(NHSTEP_TILE_SWEEP=1 python tools/time_phases.py $N 5 $mode 2>&1 | tail -1)" >> $out
  done
done
cat $out
