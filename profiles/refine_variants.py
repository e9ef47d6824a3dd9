"""Tile-shape / code-variant sweep for refine_matches (measurement tooling).

Build (container):  python profiles/refine_variants.py build TY=8 TY=16 TY=24 ...
Run (GPU box):      python profiles/refine_variants.py run [b]
Each variant is m3r_refine.cu + m3r_abi.cu compiled with extra -D flags into
mast3r-slam_b200/lib/variants/libref_<tag>.so; `run` times m3r_match_refine_fused of every variant
with CUDA events and checks its output against the shipped library's plain refine_matches entry.
"""
import ctypes as C
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mast3r-slam_b200")
VDIR = os.path.join(PKG, "lib", "variants")


def build(specs):
    os.makedirs(VDIR, exist_ok=True)
    for spec in specs:
        tag = spec.replace("=", "").replace(",", "_")
        defs = ["-D" + ("M3R_REFINE_" + d) for d in spec.split(",")]
        out = os.path.join(VDIR, f"libref_{tag}.so")
        cmd = ["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
               "-Xcompiler", "-fPIC", "-shared", "-cudart", "static", "-Xptxas", "-v"] + defs + [
               os.path.join(PKG, "csrc", "m3r_refine.cu"), os.path.join(PKG, "csrc", "m3r_abi.cu"), "-o", out]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            print(r.stderr)
            raise SystemExit(1)
        regs = [l for l in r.stderr.splitlines() if "Used" in l]
        print(tag, "fused<3> regs:", regs[1] if len(regs) > 1 else regs)


def run(b):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, PKG)
    import torch
    import mast3r_slam_backends as be
    from tests import synth

    dev = torch.device("cuda", 0)
    H, W = 384, 512
    m = synth.make_match_inputs(b, H, W, 24, seed=1235, device=dev)
    X11, X21, D11, D21 = m["X11"], m["X21"], m["D11"], m["D21"]
    rays, D11h = be.match_prep(X11, D11)
    p1, _ = be.match_iter_proj_fused(rays, X11, X21, None, 10, 1e-8, 1e-6, 0.1)
    (ref,) = be.refine_matches(D11.half(), D21.view(b, H * W, 24).half(), p1, 3, 5)
    want = ref[..., 0] + W * ref[..., 1]
    st = torch.cuda.current_stream().cuda_stream
    for path in [None] + sorted(glob.glob(os.path.join(VDIR, "libref_*.so"))):
        if path is None:
            name = "shipped"
            fn = lambda out: be._lib.m3r_match_refine_fused(D11h.data_ptr(), D21.data_ptr(), p1.data_ptr(), b, H, W,
                                                            24, 3, 5, out.data_ptr(), st)
        else:
            name = os.path.basename(path)[7:-3]
            lib = C.CDLL(path)
            f = lib.m3r_match_refine_fused
            f.argtypes = [C.c_void_p] * 3 + [C.c_int] * 6 + [C.c_void_p] * 2
            fn = lambda out, f=f: f(D11h.data_ptr(), D21.data_ptr(), p1.data_ptr(), b, H, W, 24, 3, 5,
                                    out.data_ptr(), st)
        out = torch.empty(b, H * W, dtype=torch.long, device=dev)
        for _ in range(3):
            rc = fn(out)
        assert rc == 0, rc
        a = torch.cuda.Event(enable_timing=True); z = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); a.record()
        for _ in range(20):
            fn(out)
        z.record(); torch.cuda.synchronize()
        print(f"{name:24s} {a.elapsed_time(z) / 20:.4f} ms  equal={bool(torch.equal(out, want))}", flush=True)


if __name__ == "__main__":
    if sys.argv[1] == "build":
        build(sys.argv[2:])
    else:
        run(int(sys.argv[2]) if len(sys.argv) > 2 else 8)
