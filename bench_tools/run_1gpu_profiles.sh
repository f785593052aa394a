#!/bin/bash
# One single-GPU session for the evidence under profiles/: cuDNN comparison, per-layer times, ncu captures, launch census.
O=gpurun_out
mkdir -p $O
rm -f $O/CONV_VS_CUDNN.md
timeout 400 python bench_tools/conv_vs_cudnn.py --dtype bf16 --out $O/conv_vs_cudnn_bf16.json --md $O/CONV_VS_CUDNN.md > $O/conv_vs_cudnn_bf16.log 2>&1
timeout 400 python bench_tools/conv_vs_cudnn.py --dtype fp32 --out $O/conv_vs_cudnn_fp32.json --md $O/CONV_VS_CUDNN.md > $O/conv_vs_cudnn_fp32.log 2>&1
timeout 200 python bench_tools/layer_times.py --dtype fp16 --out $O/layer_times_r2_fp16.json > $O/layer_times_fp16.log 2>&1
timeout 200 python bench_tools/layer_times.py --dtype fp32 --out $O/layer_times_r2_tf32.json > $O/layer_times_tf32.log 2>&1
# the real script path end to end (synthetic loader -> prefetcher -> step), one rank: per-step device times in the JSONL
rm -f $O/script_e2e_1gpu.jsonl
timeout 300 python resnet_ddp_apex.py --gpus 1 --epochs 1 --steps 40 --val-steps 1 --profile --log-jsonl $O/script_e2e_1gpu.jsonl > $O/script_e2e_1gpu.log 2>&1
# ncu: launch census of the comm kernels in loopback, then full captures
for what in comm fprop_tf32 wgrad_tf32; do
  timeout 120 python bench_tools/ncu_target.py $what > $O/ncu_plain_$what.log 2>&1 || { echo "plain run of $what failed" >> $O/ncu.log; continue; }
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_comm.csv python bench_tools/ncu_target.py comm > $O/ncu_comm_census.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:allreduce_kernel -s 1 -c 1 -f -o $O/prof_allreduce python bench_tools/ncu_target.py comm >> $O/ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_fprop_kernel -s 2 -c 1 -f -o $O/prof_fprop_tf32 python bench_tools/ncu_target.py fprop_tf32 >> $O/ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:conv_wgrad_kernel -s 2 -c 1 -f -o $O/prof_wgrad_tf32 python bench_tools/ncu_target.py wgrad_tf32 >> $O/ncu.log 2>&1
tail -3 $O/conv_vs_cudnn_bf16.log $O/conv_vs_cudnn_fp32.log $O/layer_times_fp16.log $O/ncu.log | cut -c1-300
python - <<PY
import json
try:
    r=[json.loads(l) for l in open("$O/script_e2e_1gpu.jsonl")]
    ms=[x["ms"] for x in r if x.get("event")=="step_time"]
    print("script e2e step ms:", [round(m,1) for m in ms[:6]], "...", [round(m,1) for m in ms[-6:]])
except Exception as e:
    print("script e2e:", e)
PY
ls -la $O/*.ncu-rep 2>/dev/null
