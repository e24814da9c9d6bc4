#!/bin/bash
# Round 2, 8-GPU pass: bench at 1/2/4/8 GPUs (ours), the reference arm and the single-process driver at 8.
set -u
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for n in 1 2 4 8; do
if [ $n -eq 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511"; fi
timeout 600 $L bench.py --gpus $n --steps 10 --warmup 3 --no-single --no-cpu-baseline > gpurun_out/r02_scale8_n$n.json 2> gpurun_out/scale8_n$n.err; echo "ours n=$n rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r02_scale8_n$n.json").read().strip().splitlines()[-1])
print("n=$n value %.3fM ms %.3f e2e %.3fM (probe ceiling %.3fM, %.1f GB/s) shared %.3fM parity %s"%(d["value"]/1e6,d["ms_per_step"],d["e2e"]["value"]/1e6,d["e2e"]["pinned_copy_probe"]["solves_per_s_ceiling"]/1e6,d["e2e"]["pinned_copy_probe"]["aggregate_gbs"],d["e2e_shared_matrix"]["value"]/1e6,d["parity_sample"]["pass"]))
PY
done
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $L bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/r02_scale8_ref_n8.json 2> gpurun_out/scale8_ref.err; echo "ref n=8 rc=$?"; python -c "
import json; d=json.loads(open('gpurun_out/r02_scale8_ref_n8.json').read().strip().splitlines()[-1]); print('ref n=8', d['value'], d.get('per_gpu_values'))"
timeout 600 python bench.py --single-process-gpus 8 --steps 8 --warmup 3 > gpurun_out/r02_single_process_n8.json 2> gpurun_out/sp8.err; echo "single-process n=8 rc=$?"; cut -c1-330 gpurun_out/r02_single_process_n8.json
