"""Build libpaffy_b200.so in-tree with nvcc for sm_100a (no JIT cache; the .so travels to the GPU box)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpaffy_b200.so")
SOURCES = ["pafgpu.cu", "paf_host.cpp", "paf_compat.cpp", "paf_api.cpp"]
CLI = os.path.join(HERE, "paffy_b200_cli")
STPAF = os.path.join(HERE, "libpaffy_b200_stpaf.a")
HEADERS = ["kernels.cuh", "encode.cuh", "serialise.cuh", "trim.cuh", "paftext.cuh", "internal.h", "paf_text.h", "paf_sonlib.c", "paffy_main.c", os.path.join("..", "..", "include", "paf_b200_compat.h"), os.path.join("..", "..", "include", "paffy_b200.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math=false",
              "-Xcompiler", "-fPIC,-O3,-Wall,-fvisibility=default", "-Xptxas", "-v", "-shared", "-cudart", "shared",
              "-Xlinker", "-Bsymbolic"]


def nvcc():
    for p in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if p and (os.path.sep not in p or os.path.exists(p)):
            return p
    raise RuntimeError("nvcc not found")


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    files = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.exists(f) and os.path.getmtime(f) > t for f in files)


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    flags = [f for f in NVCC_FLAGS if f != "--use_fast_math=false"]
    flags = flags + os.environ.get("PAFGPU_NVCC_DEFS", "").split()  # experiments only (e.g. -DENC_MINB=5)
    cmd = [nvcc()] + flags + ["-I", os.path.join(HERE, "..", "include"), "-x", "cu"] + srcs + ["-o", LIB]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode:
        raise RuntimeError("nvcc failed")
    with open(os.path.join(HERE, "build_ptxas.log"), "w") as fh:
        fh.write(r.stdout + r.stderr)
    # the command line front end (C, links the library by relative rpath)
    r = subprocess.run(["gcc", "-O2", "-o", CLI, os.path.join(CSRC, "paffy_main.c"), "-L", HERE, "-lpaffy_b200",
                        "-Wl,-rpath,$ORIGIN"], capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("building the CLI failed")
    # the sonLib-container functions of paf.h (read_pafs, write_pafs, paf_shatter ...) as a static archive for host
    # programs that bring their own sonLib (include/paf_b200_compat.h)
    obj = os.path.join(HERE, "paf_sonlib.o")
    r = subprocess.run(["gcc", "-std=gnu99", "-O2", "-fPIC", "-Wall", "-c", os.path.join(CSRC, "paf_sonlib.c"), "-o", obj],
                       capture_output=True, text=True)
    if r.returncode == 0:
        if os.path.exists(STPAF):
            os.unlink(STPAF)
        r = subprocess.run(["ar", "rcs", STPAF, obj], capture_output=True, text=True)
    if r.returncode:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("building libpaffy_b200_stpaf.a failed")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
