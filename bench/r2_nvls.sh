# native NVLS group + device-timed all-reduce latency table (2 or more GPUs)
N=$(python -c "import torch; print(min(torch.cuda.device_count(), 8))")
timeout 600 python -m pytest tests/test_gpu_multiproc.py -x -q -k "nvls" 2>&1 | tail -5
timeout 300 bin/mpiexec -n $N python bench/allreduce_latency.py 400 > gpurun_out/allreduce_latency_n$N.json 2> gpurun_out/allreduce_latency_n$N.err
cat gpurun_out/allreduce_latency_n$N.json; tail -3 gpurun_out/allreduce_latency_n$N.err
