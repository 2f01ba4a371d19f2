#!/bin/bash
# Marginal cost of each operator class inside the CUDA-graph replay of the full step: run bench.py with the class's
# entry points turned into no-ops (SDB_SKIP) and compare ms_per_step with the unablated run. Timing only.
out=${1:-gpurun_out/ablate.txt}
: > $out
run() {
  name=$1; shift
  ms=$(SDB_EXPERIMENTS=1 SDB_SKIP=$1 timeout 300 python bench.py --allow-experiments --steps 10 --warmup 3 --no-cpu --no-kernels 2>/dev/null | python -c "import json,sys; print(json.loads(sys.stdin.read())['ms_per_step'])")
  echo "$name $ms" >> $out
}
run base none
run attn_fwd sdb_attn_fwd
run attn_bwd sdb_attn_bwd
run gemm_fwd gemm_fwd
run gemm_wgrad gemm_wgrad
run groupnorm sdb_groupnorm_fwd,sdb_groupnorm_bwd
run layernorm sdb_layernorm_fwd,sdb_layernorm_bwd
run geglu sdb_geglu_fwd,sdb_geglu_bwd
run colsum sdb_colsum_bf16
run copies sdb_copy2d_bf16,sdb_add_bf16,sdb_upsample2x_fwd,sdb_upsample2x_bwd,sdb_silu_fwd,sdb_silu_bwd
run optimizer sdb_lion8_update,sdb_lion_fp32_update,sdb_global_norm
