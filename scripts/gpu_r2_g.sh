#!/bin/bash
# Round 2, pass G: whole GPU suite, bench both arms, ncu launch list + full captures of the three kernels of the step.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_ours.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-single --parity-sample 2"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
D="python tests/checks/k5_profile_driver.py 8192 5 0 4"
ncu --set full --clock-control none --import-source on -k regex:gato_pcg_scaled -s 2 -c 1 -o gpurun_out/prof_k5_pcg $D > gpurun_out/ncu2.log 2>&1; echo "full pcg rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gato_fused_solve_v2 -s 2 -c 1 -o gpurun_out/prof_k5_schur $D > gpurun_out/ncu3.log 2>&1; echo "full schur rc=$?"
for f in prof_k5_pcg prof_k5_schur; do
  ncu -i gpurun_out/$f.ncu-rep --page raw --csv > gpurun_out/$f.raw.csv 2>/dev/null
  python scripts/ncu_hotspots.py gpurun_out/$f.ncu-rep > gpurun_out/$f.hotspots.txt 2>gpurun_out/$f.hotspots.err
done
rm -f gpurun_out/*.ncu-rep
./build/floors > gpurun_out/r02_floors.json 2>/dev/null; echo "floors rc=$?"
