"""Build libhsfm_ba.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m sfm_b200.build [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libhsfm_ba.so")
SOURCES = ["ba_kernels.cu", "ba_chol.cu", "ba_index.cu", "ba_comm.cu", "ba_solver.cu", "ba_api.cu", "ba_order.cpp", "ba_noalloc.cu", "ba_scan.cu", "ba_stage.cu"]
HEADERS = ["ba_internal.h", "ba_math.cuh", os.path.join("..", "..", "include", "hsfm_ba.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unknown-pragmas",
    "-ccbin", "/usr/bin/g++",
]


FLAGS += os.environ.get("HSFM_EXTRA_NVCC", "").split()  # experiment builds, e.g. -DSCHUR_MATVEC_CTAS=5


def _stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale():
        return OUT
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, "build", os.path.splitext(src)[0] + ".o")
        cmd = [NVCC, *FLAGS, *(["-x", "cu"] if src.endswith(".cpp") else []), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src} ---\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    cmd = [NVCC, "-shared", "-o", OUT, *objs, "-ccbin", "/usr/bin/g++", "-lcudart", "-ldl"]
    subprocess.run(cmd, check=True)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
