"""Build libudp_codec.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m udp_pose_b200.build [--force]

Flags that are part of the numerical contract: -fmad=false (no implicit FMA contraction;
fused ops appear only where fmaf()/fma() is written) and no fast-math.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SOURCES = ["affine.cu", "crop.cu", "encode.cu", "decode.cu", "decode_offset.cu", "evaluate.cu", "nms.cu", "legacy.cu", "aid.cu"]
HEADERS = ["common.cuh", "blur.cuh", "blur_fast.cuh", "decode_pipe.cuh", "decode_offset_prune.cuh", "decode_offset_warp.cuh", "tma_host.cuh", os.path.join("..", "..", "include", "udp_codec.h")]
OUT = os.path.join(HERE, "libudp_codec.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-fmad=false", "-DUDP_SM_ARCH=100", "--expt-relaxed-constexpr", "--extended-lambda",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not needs_build():
        return OUT
    objs = []
    procs = []
    for s in SOURCES:
        o = os.path.join(CSRC, s.replace(".cu", ".o"))
        objs.append(o)
        procs.append((s, subprocess.Popen([NVCC] + FLAGS + ["-c", os.path.join(CSRC, s), "-o", o],
                                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for s, p in procs:
        out, _ = p.communicate()
        log.append("== %s ==\n%s" % (s, out))
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s:\n%s" % (s, out))
    # the same -gencode at link time: nvcc's default device-link target would add an empty sm_52 image to the library
    subprocess.check_call([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs + ["-lcudart"])
    with open(os.path.join(CSRC, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
