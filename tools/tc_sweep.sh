#!/bin/bash
# Tensor-core table GEMM: step / kernel time over (cluster, products, drain steps); run on the GPU box from the repo root
for cl in 1 2 4 8; do for ps in "9 1" "9 4" "6 4"; do set -- $ps
  echo "cluster $cl products $1 steps $2: $(SDRX_TC_CLUSTER=$cl SDRX_TC_PRODUCTS=$1 SDRX_TC_STEPS=$2 timeout 120 python tools/diag_timeline.py --mode tc 2>&1 | tail -1)"
done; done
