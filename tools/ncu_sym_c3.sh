#!/bin/bash
# plain run first, then one ncu --set full capture of the symmetric main pass at C3 (source-level stall samples)
set -e
python tools/prof_knn.py 100000 1536 16 2 pos
ncu --set full --import-source on --clock-control none -k regex:knn_tc_kernel --launch-skip 1 --launch-count 1 \
    -o gpurun_out/sym_c3 -f python tools/prof_knn.py 100000 1536 16 2 pos > gpurun_out/sym_c3_ncu.log 2>&1
ncu -i gpurun_out/sym_c3.ncu-rep --page details > gpurun_out/sym_c3.details.txt
ncu -i gpurun_out/sym_c3.ncu-rep --page raw --csv > gpurun_out/sym_c3.raw.csv
ncu -i gpurun_out/sym_c3.ncu-rep --page source --csv --print-source sass > gpurun_out/sym_c3.source_sass.csv 2>/dev/null || true
ncu -i gpurun_out/sym_c3.ncu-rep --page source --csv --print-source cuda > gpurun_out/sym_c3.source_cuda.csv 2>/dev/null || true
rm -f gpurun_out/sym_c3.ncu-rep
ls -la gpurun_out/sym_c3*
