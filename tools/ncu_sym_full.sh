#!/bin/bash
# plain runs first, then ncu --set full of the threshold pre-pass + symmetric main pass at C3 and C2 (final kernel);
# exports made on the box (the .ncu-rep files are too large to bring back)
set -e
python tools/prof_knn.py 100000 1536 16 2 pos
python tools/prof_knn.py 50000 1024 16 2
ncu --set full --import-source on --clock-control none -k regex:knn_tc_kernel --launch-skip 3 --launch-count 2 \
    -o gpurun_out/c3f -f python tools/prof_knn.py 100000 1536 16 2 pos > gpurun_out/c3f_ncu.log 2>&1
ncu --set full --clock-control none -k regex:knn_tc_kernel --launch-skip 3 --launch-count 2 \
    -o gpurun_out/c2f -f python tools/prof_knn.py 50000 1024 16 2 > gpurun_out/c2f_ncu.log 2>&1
for t in c3f c2f; do
  ncu -i gpurun_out/$t.ncu-rep --page details > gpurun_out/$t.details.txt
  ncu -i gpurun_out/$t.ncu-rep --page raw --csv > gpurun_out/$t.raw.csv
done
ncu -i gpurun_out/c3f.ncu-rep --page source --csv --print-source sass > gpurun_out/c3f.source_sass.csv 2>/dev/null || true
rm -f gpurun_out/c3f.ncu-rep gpurun_out/c2f.ncu-rep
ls -la gpurun_out/c3f* gpurun_out/c2f*
