# one --set full capture of kernels matching $1 from `prof_driver.py $2 1 1` into gpurun_out/$3
mkdir -p gpurun_out/$3
python profiles/prof_driver.py $2 1 1 > gpurun_out/$3/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"$1" -c 2 -o gpurun_out/$3/prof python profiles/prof_driver.py $2 1 1 > gpurun_out/$3/ncu.log 2>&1
tail -2 gpurun_out/$3/ncu.log
