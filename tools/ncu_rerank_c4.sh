#!/bin/bash
# plain run, then one ncu --set full capture (source-level stall samples) of the grouped re-rank of a C4 batch
set -e
python tools/prof_c4.py 128
ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:rerank_kernel --launch-skip 2 --launch-count 1 \
    -o gpurun_out/rr4 -f python tools/prof_c4.py 128 > gpurun_out/rr4_ncu.log 2>&1
ncu -i gpurun_out/rr4.ncu-rep --page details > gpurun_out/rr4.details.txt
ncu -i gpurun_out/rr4.ncu-rep --page raw --csv > gpurun_out/rr4.raw.csv
ncu -i gpurun_out/rr4.ncu-rep --page source --csv --print-source sass > gpurun_out/rr4.source_sass.csv 2>/dev/null || true
rm -f gpurun_out/rr4.ncu-rep
ls -la gpurun_out/rr4*
