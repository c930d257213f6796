# final state of round 2: full GPU suite, default bench line, then one ncu capture of a C4 window launch
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo rc=$? >> gpurun_out/f_pytest.log; tail -3 gpurun_out/f_pytest.log
python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo bench rc=$?
python profiles/tools/one_pass.py C4 2 200000 > gpurun_out/f_c4_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_tile2_kernel -s 20 -c 1 -o gpurun_out/r02_c4_tile2_v2 python profiles/tools/one_pass.py C4 2 200000 > gpurun_out/f_c4_ncu.log 2>&1
echo ncu rc=$?; tail -2 gpurun_out/f_c4_plain.log
