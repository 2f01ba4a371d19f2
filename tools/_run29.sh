NVCC_EXTRA=-DSDB_GEMM_EXPERIMENTS python -m stable_diffusion_training_b200._build --force > gpurun_out/dbg_build.log 2>&1
for dbg in 0 1 2 4 3 5 6; do
echo "== DBG=$dbg" >> gpurun_out/r2g_dbg.txt
SDB_GEMM_DBG=$dbg timeout 120 python tools/bench_ops.py conv 2>&1 | cut -c1-75 | head -9 >> gpurun_out/r2g_dbg.txt
done
true
