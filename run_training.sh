#!/bin/sh
# Batch-size sweep of the reference (run_training.sh:1-4), at 4 processes and -- new -- at 8.
# Uses a real mpiexec when one is installed, else the bundled spawner.
MPIEXEC=$(command -v mpiexec || echo "$(dirname "$0")/bin/mpiexec")
for np in ${NPROCS:-4 8}
do
  for bs in 1 2 4 8 16 32 64
  do
    "$MPIEXEC" -n $np python data_parallelism_train.py --nb-proc $np --batch-size $bs "$@"
  done
done
# the two tables of the reference's project report (process sweep, batch-size sweep) from the files just written
python -m distributed_neural_network_b200.utils.report "${LOG_DIR:-log}" --topology symmetric
