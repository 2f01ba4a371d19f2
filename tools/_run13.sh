timeout 900 python -m pytest tests/test_train_step_gpu.py -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1o_pytest.log
bash tools/ablate.sh gpurun_out/ablate_r1o.txt
