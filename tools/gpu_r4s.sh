#!/bin/bash
# Final-build evidence, call A: `ncu --set full` of the roofline kernel (conv3x3 320->320 @64x64, B = 8: gemm_tc_kernel<FWD, 2>)
O=gpurun_out/r4s
mkdir -p $O
timeout 300 python tools/prof_one.py conv > $O/plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -c 9 -o /tmp/conv python tools/prof_one.py conv > $O/ncu.log 2>&1
echo "ncu rc=$?"; tail -2 $O/ncu.log; ls -la /tmp/conv.ncu-rep
ncu -i /tmp/conv.ncu-rep --page raw --csv > $O/conv_raw.csv 2> $O/raw.err
ncu -i /tmp/conv.ncu-rep --page details --csv > $O/conv_details.csv 2>> $O/raw.err
python - <<'PY'
import csv
rows = list(csv.reader(open("gpurun_out/r4s/conv_raw.csv")))
h = rows[0]
want = ["ID", "Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__inst_executed_pipe_tc.sum", "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active"]
idx = [(w, h.index(w)) for w in want if w in h]
print("units:", [(w, rows[1][i]) for w, i in idx])
for r in rows[2:]:
    print([r[i][:60] for _, i in idx])
PY
du -sh $O
