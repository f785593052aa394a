#!/bin/bash
# bench.py lines of every mode at N GPUs (and optionally the reference arms): bash bench_tools/run_bench_modes.sh N [ref]
N=${1:-4}
O=gpurun_out
mkdir -p $O
rm -f $O/bench_${N}gpu.jsonl $O/bench_ref_${N}gpu.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "1" ]; then TR="python"; TRP=""; else TRP="--master-port 29511"; fi
for m in "--mode ddp_amp" "--mode ddp_amp --sync-bn" "--mode ddp"; do
  timeout 240 $TR $TRP bench.py --gpus $N --steps 20 --warmup 5 $m 2>&1 | grep "^{" >> $O/bench_${N}gpu.jsonl
done
timeout 240 python bench.py --gpus $N --mode dp --steps 20 --warmup 5 2>&1 | grep "^{" >> $O/bench_${N}gpu.jsonl
if [ "$2" = "ref" ]; then
  for m in ddp_amp ddp dp; do
    if [ $m = dp ] || [ "$N" = "1" ]; then L="python"; else L="$TR --master-port 29512"; fi
    timeout 500 $L bench.py --impl reference --gpus $N --mode $m --steps 5 --warmup 3 2>&1 | grep "^{" >> $O/bench_ref_${N}gpu.jsonl
  done
fi
python - <<PY
import json
for f in ("$O/bench_${N}gpu.jsonl", "$O/bench_ref_${N}gpu.jsonl"):
    try:
        for l in open(f):
            d = json.loads(l)
            print(d["impl"], d["config"]["mode"], d["config"].get("sync_bn"), d["dtype"], d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["config"].get("comm"), "util", d.get("gpu_util_pct"), "mem", d.get("peak_mem_gb"))
    except FileNotFoundError:
        pass
PY
