"""Compile oracle/m3r_oracle.c and oracle/m3r_render_oracle.c (the CPU restatements; TEST
INFRASTRUCTURE ONLY) with gcc into one shared library."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "m3r_oracle.c")
SRCS = [SRC, os.path.join(HERE, "m3r_render_oracle.c")]
LIB = os.path.join(HERE, "libm3r_oracle.so")


def build(force=False):
    if (not force and os.path.exists(LIB)
            and all(os.path.getmtime(LIB) >= os.path.getmtime(s) for s in SRCS)):
        return LIB
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fopenmp", "-shared", "-fPIC",
           "-Wall", "-o", LIB] + SRCS + ["-lm"]
    subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build(force=True))
