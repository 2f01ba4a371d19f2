#!/bin/bash
# Round-2 GPU call C: `ncu --set full` captures of the dominant kernels of THIS build (tools/prof_all.py). The raw
# pages are exported to CSV on the box (the .ncu-rep files exceed the 64 MiB that come back).
O=gpurun_out/r4c
mkdir -p $O
echo "== attention tests (SDB_ATTN_POLY=4 default)"; timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_train_step_gpu.py -q -x > $O/pytest_attn.log 2>&1; tail -2 $O/pytest_attn.log
timeout 200 python tools/bench_ops.py attn > $O/ops_attn_poly4.txt 2>&1
for tag in poly0 poly2 poly3; do
  SDB_EXPERIMENTS=1 SDB_LIB_PATH=$PWD/stable_diffusion_training_b200/libsdb200_$tag.so timeout 200 python tools/bench_ops.py attn > $O/ops_attn_$tag.txt 2>&1
done
head -2 $O/ops_attn_poly*.txt
python tools/prof_all.py > $O/plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:'gemm_tc|attn_|gn_fused|lion8|ln_|geglu' -c 120 \
    -o /tmp/prof python tools/prof_all.py > $O/ncu.log 2>&1
echo "rc=$?"; tail -2 $O/ncu.log; ls -la /tmp/prof.ncu-rep
ncu -i /tmp/prof.ncu-rep --page raw --csv > $O/raw.csv 2> $O/raw.err
ncu -i /tmp/prof.ncu-rep --page details --csv > $O/details.csv 2>> $O/raw.err
# per-instruction pages of the second launch of the kernels under work (ids are 0-based launch indices of the report)
python - <<'PY' > /tmp/ids.txt
import csv
rows = list(csv.reader(open("gpurun_out/r4c/raw.csv")))
hdr = rows[0]
ki, ii = hdr.index("Kernel Name"), hdr.index("ID")
seen = {}
for r in rows[2:]:
    seen.setdefault(r[ki].split("(")[0], []).append(r[ii])
for k, v in seen.items():
    print(k.replace(" ", "_"), v[-1])
PY
cat /tmp/ids.txt
while read name id; do
  case "$name" in
    *attn_bwd_kernel\<0,_2\>*|*attn_bwd_kernel\<1,_2\>*|*attn_fwd2*|*gn_fused_kernel\<1,_0,_1\>*|*gn_fused_kernel\<1,_1,_1\>*|*gemm_tc_kernel\<1,_1\>*|*gemm_tc_kernel\<0,_2\>*)
      f=$(echo "$name" | tr -c 'A-Za-z0-9_\n' '_')
      ncu -i /tmp/prof.ncu-rep --page source --csv --print-source sass --launch-skip $id --launch-count 1 2>/dev/null | head -4000 > $O/source_$f.csv ;;
  esac
done < /tmp/ids.txt
ls -la $O; du -sh $O
