#!/bin/bash
# the reference's unchanged ij on 2 ranks (hypre's own distributed setup over oracle/mpishim), solve on 2 GPUs
mkdir -p gpurun_out
export OMP_NUM_THREADS=4 HYPRE_B200_VERBOSE=1
for r in 18 16; do
  timeout 22 oracle/_ref/mpirun_shim -np 2 hypre_b200/lib/ij_b200_mp -n 40 40 40 -P 2 1 1 -solver 1 -rlx $r > gpurun_out/mp2_rlx$r.log 2>&1
  echo "rlx $r rc=$?"; grep -E "^Iterations|Final Relative|hypre_b200" gpurun_out/mp2_rlx$r.log | head -n 8
done
