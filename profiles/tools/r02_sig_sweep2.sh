# second part of the HC_ITEM_SIG sweep: C4 at 200k fragments (window / group-size knobs), C5 and C5r
mkdir -p gpurun_out
run() { # name workload nfrag steps env...
  local name=$1 wl=$2 nf=$3 steps=$4; shift 4
  env "$@" timeout 120 python bench.py --workload $wl --n-frag $nf --no-cpu-baseline --no-also --steps $steps --warmup 3 > gpurun_out/h_$name.json 2>> gpurun_out/h_err.log
  python - gpurun_out/h_$name.json $name <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[2], round(d['ms_per_step'],3), d['pairs']['edges'], d['pairs']['scored'], flush=True)
except Exception as e: print(sys.argv[2], "ERR", e, flush=True)
P
}
run C4_sig0 C4 200000 3 HC_ITEM_SIG=0
run C4_sig1 C4 200000 3 HC_ITEM_SIG=1
run C4_sig1_div8 C4 200000 3 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=8
run C4_sig1_div32 C4 200000 3 HC_ITEM_SIG=1 HC_ITEM_WINDOW_DIV=32
run C4_sig1_min4 C4 200000 3 HC_ITEM_SIG=1 HC_ITEM_SIG_MIN=4
run C4_sig1_min32 C4 200000 3 HC_ITEM_SIG=1 HC_ITEM_SIG_MIN=32
for wl in C5 C5r; do
  run ${wl}_sig0 $wl 0 20 HC_ITEM_SIG=0
  run ${wl}_sig1 $wl 0 20 HC_ITEM_SIG=1
done
tail -3 gpurun_out/h_err.log
