export VK_TUNE="5=0,6=192"
python tools/tune_bench.py --what timeline > gpurun_out/tail_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:vk_raycast -c 1 -o gpurun_out/prof_tail python tools/tune_bench.py --what timeline > gpurun_out/tail_ncu.log 2>&1
tail -20 gpurun_out/tail_plain.log
