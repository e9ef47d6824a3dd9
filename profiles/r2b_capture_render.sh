# ncu --set full of the viewer splat kernels (one GPU).  Outputs: gpurun_out/r2b/
mkdir -p gpurun_out/r2b
python profiles/prof_driver.py render 2 > gpurun_out/r2b/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'render_' -c 8 -o gpurun_out/r2b/render python profiles/prof_driver.py render 2 > gpurun_out/r2b/ncu.log 2>&1
tail -2 gpurun_out/r2b/plain.log; tail -2 gpurun_out/r2b/ncu.log
