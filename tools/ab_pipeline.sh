#!/bin/bash
# A/B of tendency-kernel build switches on the GPU box (one gpurun call): phase timings at 256^3 for each variant.
cd /root/repo
run() { python tools/time_phases.py 256 10 $1 2>&1 | tail -1 | cut -c1-60; }
echo "== default build contracted: $(run fma)"
bash tools/variant.sh "-DTF_DIVUNROLL=2" | tail -1
echo "== +divunroll2 contracted: $(run fma)"
bash tools/variant.sh "-DTF_DIVUNROLL=4" | tail -1
echo "== +divunroll4 contracted: $(run fma)"
bash tools/variant.sh "-DTF_DIVUNROLL=8" | tail -1
echo "== +divunroll8 contracted: $(run fma)"
bash tools/variant.sh "-DTF_DIVUNROLL=4" exact | tail -1
echo "== divunroll4 exact: $(run)"
