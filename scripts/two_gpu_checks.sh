mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_reference_drivers.py -x -q -m gpu -k "distributed_two_gpus" > gpurun_out/pytest_2gpu_train.log 2>&1; echo ddp-train rc=$?; tail -5 gpurun_out/pytest_2gpu_train.log
for mode in "" "--ddp" "--stock-caller"; do
  tag=$(echo "x$mode" | tr -d ' -')
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 20 --warmup 3 --no-cpu-baseline --no-eager-baseline --no-hbm-kernels $mode > gpurun_out/bench_2gpu_$tag.json 2> gpurun_out/bench_2gpu_$tag.err; echo "bench $mode rc=$?"; tail -c 600 gpurun_out/bench_2gpu_$tag.json | head -c 400; echo
done
