#!/bin/bash
# one --set full capture of the matrix-free apply kernel (Stokes p4 and mixed Poisson p6)
mkdir -p gpurun_out
O=gpurun_out
NCU="ncu --clock-control none"
timeout 600 $NCU --set full --import-source on -k regex:element_vector_dmma_kernel -s 2 -c 1 -f -o $O/applyd4_full \
    python bench.py --steps 2 --warmup 3 --no-cpu --skip-e2e > $O/ncu_vd4.log 2>&1
ncu -i $O/applyd4_full.ncu-rep --page raw --csv > $O/applyd4_raw.csv 2>/dev/null
ncu -i $O/applyd4_full.ncu-rep --page source --csv > $O/applyd4_src.csv 2>/dev/null
