"""Host-side mirror of the reference's caller API for the hot path (same module and function
names as /root/reference/mast3r_slam/{matching,tracker,global_opt,geometry,config}.py), so the
reference's callers — and parity tests that read like them — run unchanged on top of the
B200 kernels in `mast3r_slam_backends`.  Nothing here computes on the CPU: every per-pixel /
per-point operation is a CUDA kernel behind the C ABI; torch only allocates and indexes."""
