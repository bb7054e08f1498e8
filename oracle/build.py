"""Compile the C oracle (test infrastructure) into oracle/libudp_oracle.so with gcc.

-ffp-contract=off is required: the restatement uses fmaf() exactly where OpenCV fuses and
nowhere else.  Called by __graft_entry__.build(); building the checker is not using it.
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
SRCS = ["udp_oracle.c", "udp_oracle_encode.c", "udp_oracle_decode.c"]
OUT = os.path.join(HERE, "libudp_oracle.so")


def build(force=False):
    srcs = [os.path.join(HERE, s) for s in SRCS]
    if (not force and os.path.exists(OUT)
            and all(os.path.getmtime(OUT) >= os.path.getmtime(s) for s in srcs)):
        return OUT
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-mfma", "-fPIC", "-shared", "-fopenmp",
           "-o", OUT] + srcs + ["-lm"]
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force=True))
