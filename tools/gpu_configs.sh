#!/bin/bash
# Runs on the GPU box: the other BASELINE.json configs (C3, C4) and the stress scene (C5) through the C++ CLI
# at reduced spp (time scales linearly in spp); one JSON stats line each, images under gpurun_out/.
mkdir -p gpurun_out
CLI=./raytracing-weekend_b200/rtw_cli
echo "C2"; $CLI -s 0 -dx 600 -dy 600 -ns 256 -j -o gpurun_out/c2.ppm 2>&1 | tail -1
echo "C3"; $CLI -s 1 -dx 1920 -dy 1080 -ns 50 -j -o gpurun_out/c3.ppm 2>&1 | tail -1
echo "scene2"; $CLI -s 2 -dx 1200 -dy 600 -ns 64 -j -o gpurun_out/s2.ppm 2>&1 | tail -1
echo "scene3"; $CLI -s 3 -dx 600 -dy 600 -ns 64 -j -o gpurun_out/s3.ppm 2>&1 | tail -1
echo "C4"; $CLI -s 4 -dx 800 -dy 800 -ns 100 -j -o gpurun_out/c4.ppm 2>&1 | tail -1
