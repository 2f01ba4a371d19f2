timeout 900 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1n_pytest.log
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -c 3 -o gpurun_out/r1n_gemm -f python tools/prof_one.py conv > gpurun_out/ncu_gemm.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:attn_ -c 4 -s 4 -o gpurun_out/r1n_attn -f python tools/prof_one.py attn > gpurun_out/ncu_attn.log 2>&1
ncu --set full --clock-control none -k regex:gn_fused -c 4 -o gpurun_out/r1n_gn -f python tools/prof_one.py gn > gpurun_out/ncu_gn.log 2>&1
true
