#!/bin/bash
# Round-2 GPU call C: `ncu --set full` captures of the dominant kernels of THIS build (tools/prof_all.py).
O=gpurun_out/r4c
mkdir -p $O
python tools/prof_all.py > $O/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc|attn_|gn_fused|lion8|ln_|geglu' -c 120 \
    -o $O/prof python tools/prof_all.py > $O/ncu.log 2>&1
echo "rc=$?"; tail -3 $O/ncu.log; ls -la $O
