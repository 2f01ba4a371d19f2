for mt in 1 2; do for bn in 80 112 160 256; do
echo "== MT=$mt BN=$bn" >> gpurun_out/r2f_tiles.txt
SDB_MT=$mt SDB_BN=$bn timeout 120 python tools/bench_ops.py conv linear 2>&1 | cut -c1-75 >> gpurun_out/r2f_tiles.txt
done; done
true
