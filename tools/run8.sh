export PCCL_DEVICE_TIMEOUT_MS=8000
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k "8" > gpurun_out/gpu8_tests.log 2>&1; echo "pytest rc=$?" >> gpurun_out/gpu8_tests.log; tail -5 gpurun_out/gpu8_tests.log | cut -c1-1500
timeout 300 $TR --master-port 29511 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench8.json 2> gpurun_out/bench8.err; echo "bench rc=$?"; cat gpurun_out/bench8.json
timeout 400 $TR --master-port 29512 benchmarks/sweep.py --iters 10 --concurrency --quant > gpurun_out/sweep8.jsonl 2> gpurun_out/sweep8.err; echo "sweep rc=$?"; tail -2 gpurun_out/sweep8.err | cut -c1-300
for q in fp8 minmax_u8; do timeout 300 $TR --master-port 29513 bench.py --gpus 8 --steps 6 --warmup 3 --params 1000000000 --quant $q > gpurun_out/bench8_1B_$q.json 2> gpurun_out/bq8.err; echo "quant $q rc=$?"; cut -c1-900 gpurun_out/bench8_1B_$q.json; done
timeout 300 $TR --master-port 29514 benchmarks/shared_state_llama8b.py > gpurun_out/shared_state_8gpu.json 2> gpurun_out/ss8.err; echo "ss rc=$?"; cat gpurun_out/shared_state_8gpu.json
timeout 200 $TR --master-port 29515 benchmarks/nvlink_counters.py > gpurun_out/nvlink_counters8.jsonl 2> gpurun_out/nvc8.err; echo "nvc rc=$?"; cat gpurun_out/nvlink_counters8.jsonl
timeout 200 $TR --master-port 29516 tools/tune_nvls.py 1048576 16777216 600000000 > gpurun_out/tune_nvls8.jsonl 2> gpurun_out/tn8.err; echo "tune rc=$?"; cat gpurun_out/tune_nvls8.jsonl
