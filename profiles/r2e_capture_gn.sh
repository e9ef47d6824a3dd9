# ncu --set full of the GN kernels after the accumulate instruction diet.  Outputs: gpurun_out/r2e/
mkdir -p gpurun_out/r2e
python profiles/prof_driver.py gn 2 2 > gpurun_out/r2e/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'gn_accum' -c 3 -o gpurun_out/r2e/gn python profiles/prof_driver.py gn 2 2 > gpurun_out/r2e/ncu.log 2>&1
tail -1 gpurun_out/r2e/plain.log; tail -1 gpurun_out/r2e/ncu.log
