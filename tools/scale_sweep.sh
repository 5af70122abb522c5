#!/bin/bash
# Multi-GPU sweep on one box (run under `gpurun --gpus 8`): bench.py at N = 2, 4, 8 with the shared view pass and without,
# C5 at N = 8, the single-process path (app_headless --gpus N), the 2-GPU tests. Output: gpurun_out/<tag>_*.json / .txt
tag=${1:-r02i}
out=gpurun_out
mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv,noheader > $out/${tag}_gpus.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_boundary.py tests/test_gpu_parity.py -m gpu -x -q -k "two_gpus" > $out/${tag}_two_gpu_tests.txt 2>&1
tail -3 $out/${tag}_two_gpu_tests.txt
port=29500
for n in 1 2 4 8; do
  for share in noshare share; do
    [ $n = 1 ] && [ $share = share ] && continue
    flag="--share"; [ $share = noshare ] && flag=""
    port=$((port+1))
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n --steps 100 --warmup 5 $flag > $out/${tag}_bench_n${n}_${share}.json 2> $out/${tag}_bench_n${n}_${share}.err
    python - <<PY
import json
try:
    d = json.loads(open("$out/${tag}_bench_n${n}_${share}.json").read().strip().splitlines()[-1])
    print("N=$n $share value %.2f %s  ms/step %.4f  e2e %.2f  kernels %s gathered_ok %s" % (d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d.get("kernel_ms_rank0"), d.get("gathered_frame_equals_one_gpu_frame")))
except Exception as e:
    print("N=$n $share failed:", e)
PY
  done
done
port=$((port+1))
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $port bench.py --config c5 --gpus 8 > $out/${tag}_bench_c5_n8.json 2> $out/${tag}_bench_c5_n8.err
tail -c 600 $out/${tag}_bench_c5_n8.json; echo
obj=/tmp/atray_meshes/standin_272x256.obj
for n in 1 2 4 8; do
  timeout 120 Source/engine/app_headless --obj $obj --res 3840x2160 --gpus $n --gather bgrx --frames 30 2>&1 | tail -1 | tee -a $out/${tag}_app_headless_gpus.txt
  ATRAY_MULTI_SHARE=1 timeout 120 Source/engine/app_headless --obj $obj --res 3840x2160 --gpus $n --gather bgrx --frames 30 2>&1 | tail -1 | sed 's/^/shared pass: /' | tee -a $out/${tag}_app_headless_gpus.txt
done
