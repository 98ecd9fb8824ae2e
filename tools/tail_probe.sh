# ncu on a single ray-march warp tile: VK_TUNE knob 6 = x | y << 10 (diagnostics)
export VK_TUNE="5=0,6=$((192 + (276 << 10)))"
python tools/tune_bench.py --what timeline > gpurun_out/tail_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:vk_raycast -c 1 -o gpurun_out/prof_tail python tools/tune_bench.py --what timeline > gpurun_out/tail_ncu.log 2>&1
grep -A1 "long warp" gpurun_out/tail_plain.log | head -4
