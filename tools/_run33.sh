timeout 200 python tools/bench_ops.py attn > gpurun_out/r2m_ops_sleep.txt 2>&1
NVCC_EXTRA=-DSDB_ISSUER_SPIN python -m stable_diffusion_training_b200._build --force > gpurun_out/spin_build.log 2>&1
timeout 200 python tools/bench_ops.py attn > gpurun_out/r2m_ops_spin.txt 2>&1
timeout 300 python -m pytest tests/test_ops_gpu.py -m gpu -q -k attention 2>&1 | tail -2 > gpurun_out/r2m_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-kernels > gpurun_out/r2m_bench_spin.json 2> gpurun_out/r2m_bench.err
true
