#!/bin/bash
# round 2, multi-GPU visit: usage  bash tools/gpu_r2_multi.sh <n_gpus> [config5]
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | wc -l > gpurun_out/ngpu.txt
timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_dropin.py -m gpu -q -rs > gpurun_out/pytest_multi_n$N.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_multi_n$N.log
tail -n 25 gpurun_out/pytest_multi_n$N.log
for n in $N $(( N > 2 ? N / 2 : 0 )) $(( N > 4 ? N / 4 : 0 )); do
  [ "$n" -ge 2 ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) bench.py --gpus $n --steps 5 --warmup 3 --cpu-sample 0 --no-tree > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n=$n exit $?"; cat gpurun_out/bench_n$n.json; tail -n 3 gpurun_out/bench_n$n.err
done
if [ "$2" = "config5" ]; then
  for n in $N $(( N / 2 )) $(( N / 4 )) 1; do
    [ "$n" -ge 1 ] || continue
    if [ "$n" -eq 1 ]; then
      timeout 900 python bench.py --config 5 --gpus 1 --steps 3 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n1.json 2> gpurun_out/config5_n1.err
    else
      timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --config 5 --gpus $n --steps 5 --warmup 3 --cpu-sample 0 > gpurun_out/config5_n$n.json 2> gpurun_out/config5_n$n.err
    fi
    echo "config5 n=$n exit $?"; cat gpurun_out/config5_n$n.json; tail -n 2 gpurun_out/config5_n$n.err
  done
fi
