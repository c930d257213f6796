set -x
mkdir -p gpurun_out
python profiles/tools/fuzz_parity.py 150 9000 > gpurun_out/r02_fuzz.json 2> gpurun_out/r02_fuzz.err; echo fuzz rc=$?
python profiles/tools/one_pass.py C5 3 > gpurun_out/c5_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_tile_kernel -s 1 -c 1 -o gpurun_out/r02_c5_streamed_v2 python profiles/tools/one_pass.py C5 3 > gpurun_out/c5_ncu.log 2>&1
echo ncu rc=$?
tail -3 gpurun_out/c5_plain.log gpurun_out/r02_fuzz.json
