#!/bin/bash
# Round 2, multi-GPU pass (N = number of GPUs on the box): torchrun bench (ours + reference arm), the single-process
# multi-GPU entry point, the 2-GPU tests.
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "multiplan" > gpurun_out/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -3 gpurun_out/pytest_multi.log
for n in $(seq 1 $N); do case $n in 1|2|4|8) ;; *) continue;; esac
if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511"; fi
timeout 900 $L bench.py --gpus $n --steps 10 --warmup 3 --no-single > gpurun_out/r02_scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "ours n=$n rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r02_scale_n$n.json").read().strip().splitlines()[-1])
print("n=$n value %.3fM e2e %.3fM (probe ceiling %.3fM, %.1f GB/s) shared %.3fM parity %s"%(d["value"]/1e6,d["e2e"]["value"]/1e6,d["e2e"]["pinned_copy_probe"]["solves_per_s_ceiling"]/1e6,d["e2e"]["pinned_copy_probe"]["aggregate_gbs"],d["e2e_shared_matrix"]["value"]/1e6,d["parity_sample"]["pass"]))
PY
timeout 900 $L bench.py --impl reference --gpus $n --steps 2 --warmup 1 > gpurun_out/r02_scale_ref_n$n.json 2> gpurun_out/scale_ref_n$n.err; echo "ref n=$n rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_scale_ref_n$n.json').read().strip().splitlines()[-1]); print('ref n=$n', d['value'], d.get('per_gpu_values'))"
timeout 600 python bench.py --single-process-gpus $n --steps 8 --warmup 3 > gpurun_out/r02_single_process_n$n.json 2> gpurun_out/sp_n$n.err; echo "single-process n=$n rc=$?"; cut -c1-400 gpurun_out/r02_single_process_n$n.json
done
