#!/bin/bash
# One multi-GPU session: collective tests, comm sweep, every mode's bench line, an in-step --profile log.
#   bash bench_tools/run_8gpu.sh N      (N = number of GPUs; outputs under gpurun_out/)
N=${1:-8}
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 python -m pytest tests/test_gpu_comm.py -m gpu -q -x -s 2>&1 | grep -E "cos |passed|failed|Error|error" | tail -15 > $O/t_comm_${N}gpu.log
timeout 300 $TR --master-port 29533 bench_tools/comm_sweep.py --out $O/comm_sweep_${N}gpu.json > $O/comm_sweep_${N}gpu.log 2>&1
for m in "--mode ddp_amp" "--mode ddp_amp --sync-bn" "--mode ddp"; do
  timeout 240 $TR --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 $m 2>&1 | grep "^{" >> $O/bench_${N}gpu.jsonl
done
timeout 240 python bench.py --gpus $N --mode dp --steps 20 --warmup 5 2>&1 | grep "^{" >> $O/bench_${N}gpu.jsonl
timeout 240 python resnet_ddp_apex.py --gpus $N --epochs 1 --steps 12 --val-steps 1 --profile --log-jsonl $O/step_profile_${N}gpu.jsonl > $O/step_profile_${N}gpu.log 2>&1
if [ "$2" = "ref" ]; then
  for m in ddp dp; do
    if [ $m = dp ]; then L="python"; else L="$TR --master-port 29512"; fi
    timeout 400 $L bench.py --impl reference --gpus $N --mode $m --steps 5 --warmup 3 2>&1 | grep "^{" >> $O/bench_ref_${N}gpu.jsonl
  done
fi
python - <<PY
import json
for f in ("$O/bench_${N}gpu.jsonl", "$O/bench_ref_${N}gpu.jsonl"):
    try:
        for l in open(f):
            d = json.loads(l)
            print(d["impl"], d["config"]["mode"], d["config"].get("sync_bn"), d["dtype"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["config"].get("comm"))
    except FileNotFoundError:
        pass
PY
cat $O/t_comm_${N}gpu.log
