timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 400 python -m pytest tests/test_gpu_multiproc.py -x -q 2>&1 | tail -3
mkdir -p gpurun_out/log
timeout 200 bin/mpiexec -n 2 python data_parallelism_train.py --nb-proc 2 --batch-size 16 --synthetic --epochs 5 --log-dir gpurun_out/log --jsonl gpurun_out/log/dp2_bs16.jsonl 2>&1 | tail -4
ls gpurun_out/log
