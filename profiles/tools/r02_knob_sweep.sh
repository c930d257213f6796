set -x
for wl in C5 C5r; do
for div in 16 8 4 2 1; do
  HC_ITEM_WINDOW_DIV=$div python bench.py --workload $wl --no-cpu-baseline --no-also --steps 20 --warmup 5 > gpurun_out/k_${wl}_div$div.json 2>> gpurun_out/k_err.log
done
for rows in 4 16 32; do
  HC_WARP_ITEM_ROWS=$rows python bench.py --workload $wl --no-cpu-baseline --no-also --steps 20 --warmup 5 > gpurun_out/k_${wl}_rows$rows.json 2>> gpurun_out/k_err.log
done
done
for div in 8 32; do
  HC_ITEM_WINDOW_DIV=$div python bench.py --workload C2 --no-cpu-baseline --no-also --steps 10 --warmup 3 > gpurun_out/k_C2_div$div.json 2>> gpurun_out/k_err.log
done
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py smoke > gpurun_out/k_memcheck.log 2>&1; echo memcheck rc=$? >> gpurun_out/k_memcheck.log
tail -5 gpurun_out/k_memcheck.log
