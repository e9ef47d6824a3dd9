# compute-sanitizer passes over the library (run under gpurun; one GPU).  Outputs: gpurun_out/r2n/
#   memcheck.log  : --tool memcheck over the whole `pytest -m gpu` suite; torch's caching allocator is
#                   switched off so that every tensor is its own cudaMalloc and an out-of-bounds
#                   access of one of our kernels cannot land inside a neighbouring pooled block
#   racecheck.log : --tool racecheck (shared-memory hazards) over __graft_entry__.smoke()
#   initcheck.log : --tool initcheck (reads of never-written global memory) over smoke()
# Each pass sits under its own `timeout`: the tile Cholesky spins on flags and must not be left hanging.
mkdir -p gpurun_out/r2n
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
timeout 420 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file gpurun_out/r2n/memcheck.log \
  python -m pytest tests -m gpu -x -q > gpurun_out/r2n/memcheck_pytest.log 2>&1
echo "memcheck rc=$?"; tail -2 gpurun_out/r2n/memcheck_pytest.log; tail -3 gpurun_out/r2n/memcheck.log
timeout 240 compute-sanitizer --tool racecheck --error-exitcode 7 --log-file gpurun_out/r2n/racecheck.log \
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2n/racecheck_smoke.log 2>&1
echo "racecheck rc=$?"; tail -1 gpurun_out/r2n/racecheck_smoke.log; tail -3 gpurun_out/r2n/racecheck.log
timeout 240 compute-sanitizer --tool initcheck --error-exitcode 7 --log-file gpurun_out/r2n/initcheck.log \
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2n/initcheck_smoke.log 2>&1
echo "initcheck rc=$?"; tail -1 gpurun_out/r2n/initcheck_smoke.log; tail -3 gpurun_out/r2n/initcheck.log
