#!/bin/bash
# A/B of the shared view pass at N ranks (run under gpurun --gpus N): ranges-only (default), first version, not shared
n=${1:-2}; tag=${2:-r02j}; out=gpurun_out; mkdir -p $out; port=29600
timeout 300 python -m pytest tests/test_gpu_boundary.py tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus" 2>&1 | tail -2
for mode in ranges stamps noshare; do
  flag="--share"
  [ $mode = noshare ] && flag=""
  [ $mode = stamps ] && export ATRAY_SHARE_MODE=0 || unset ATRAY_SHARE_MODE
  port=$((port+1))
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 50 --warmup 5 $flag > $out/${tag}_bench_n${n}_${mode}.json 2> $out/${tag}_bench_n${n}_${mode}.err
  python - <<PY
import json
try:
    d = json.loads(open("$out/${tag}_bench_n${n}_${mode}.json").read().strip().splitlines()[-1])
    print("N=$n $mode value %.2f %s  ms/step %.4f  e2e %.2f  kernels %s gathered_ok %s" % (d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], {k: round(v * 1e3, 1) for k, v in d.get("kernel_ms_rank0").items()}, d.get("gathered_frame_equals_one_gpu_frame")))
except Exception as e:
    print("N=$n $mode failed:", e)
PY
done
