# HC_ITEM_SIG (indel-signature grouping of the column blocks) x HC_ITEM_WINDOW_DIV on the bench workloads
mkdir -p gpurun_out
run() { # name workload steps env...
  local name=$1 wl=$2 steps=$3; shift 3
  env "$@" python bench.py --workload $wl --no-cpu-baseline --no-also --steps $steps --warmup 3 > gpurun_out/g_$name.json 2>> gpurun_out/g_err.log
  python - gpurun_out/g_$name.json $name <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d['ms_per_step'],3), d['pairs']['edges'], d['pairs']['scored'], flush=True)
except Exception as e: print(sys.argv[2], "ERR", e, flush=True)
P
}
for div in 16 8 4; do
  run C2_sig0_div$div C2 10 HC_ITEM_SIG=0 HC_ITEM_WINDOW_DIV=$div
  run C2_sig1_div$div C2 10 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=$div
done
run C2_sig1_div8_min4 C2 10 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=8 HC_ITEM_SIG_MIN=4
run C2_sig1_div8_min16 C2 10 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=8 HC_ITEM_SIG_MIN=16
for div in 16 8; do
  run C4_sig0_div$div C4 3 HC_ITEM_SIG=0 HC_ITEM_WINDOW_DIV=$div
  run C4_sig1_div$div C4 3 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=$div
done
for wl in C5 C5r; do
  run ${wl}_sig0 $wl 20 HC_ITEM_SIG=0
  run ${wl}_sig1 $wl 20 HC_ITEM_SIG=1
  run ${wl}_sig1_div4 $wl 20 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=4
done
tail -3 gpurun_out/g_err.log
