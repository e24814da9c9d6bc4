#!/bin/bash
# ncu of kernel 5's two halves (launch list, full capture of each half, source hot spots).  GATO_B200_LIB selects the build.
set -u
mkdir -p gpurun_out
D="python tests/checks/k5_profile_driver.py 8192 5 8192 4"
$D > gpurun_out/k5_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file gpurun_out/k5_launches.csv $D > gpurun_out/ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gato_pcg_scaled -s 2 -c 1 -o gpurun_out/prof_k5_pcg $D > gpurun_out/ncu2.log 2>&1; echo "full pcg rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gato_fused_solve_v2 -s 2 -c 1 -o gpurun_out/prof_k5_schur $D > gpurun_out/ncu3.log 2>&1; echo "full schur rc=$?"
for f in prof_k5_pcg prof_k5_schur; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null
  python scripts/ncu_hotspots.py gpurun_out/$f.ncu-rep > gpurun_out/$f.hotspots.txt 2>gpurun_out/$f.hotspots.err
done
rm -f gpurun_out/*.ncu-rep
