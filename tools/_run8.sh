b() { SDB_MT=$1 SDB_SKIP=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels 2>/dev/null | python -c "import json,sys; print(json.loads(sys.stdin.read())['ms_per_step'])"; }
echo "mt0 base $(b 0 none)" > gpurun_out/r1j_ab.txt
echo "mt1 base $(b 1 none)" >> gpurun_out/r1j_ab.txt
echo "mt0 nofwd $(b 0 gemm_fwd)" >> gpurun_out/r1j_ab.txt
echo "mt0 nowgrad $(b 0 gemm_wgrad)" >> gpurun_out/r1j_ab.txt
echo "mt1 nofwd $(b 1 gemm_fwd)" >> gpurun_out/r1j_ab.txt
echo "mt1 nowgrad $(b 1 gemm_wgrad)" >> gpurun_out/r1j_ab.txt
