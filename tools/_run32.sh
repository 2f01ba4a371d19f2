timeout 200 python tools/prof_one.py conv > /dev/null 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -c 6 -o gpurun_out/r2l_gemm -f python tools/prof_one.py conv > gpurun_out/r2l_ncu.log 2>&1
timeout 200 python tools/prof_one.py attn > /dev/null 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_ -c 3 -o gpurun_out/r2l_attn -f python tools/prof_one.py attn >> gpurun_out/r2l_ncu.log 2>&1
true
