"""Build libb200lora.so (in-tree, sm_100a) with nvcc.  No torch headers involved: the library
is a plain C-ABI shared object (include/b200lora.h) loaded through ctypes."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libb200lora.so")
SOURCES = ["api.cu", "linear_tc.cu", "linear_tc2.cu", "conv_aux.cu", "linear_simt.cu", "reduce.cu", "merge.cu", "merge_batch.cu", "geglu.cu", "optim.cu", "ti.cu", "flat.cu"]
HEADERS = ["common.cuh", "ptx.cuh", os.path.join("..", "..", "include", "b200lora.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "-cudart", "static",
]


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every .cu for sm_100a and link libb200lora.so next to this file."""
    if not force and not _stale():
        return OUT
    objs = []
    procs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    for s in SOURCES:
        o = os.path.join(HERE, "build", s.replace(".cu", ".o"))
        cmd = [NVCC, *FLAGS, *os.environ.get("B200LORA_NVCC_DEFS", "").split(), "-c", os.path.join(CSRC, s), "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas")
            cmd.insert(2, "-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
        objs.append(o)
    failed = False
    for s, p in procs:
        out, _ = p.communicate()
        if out and (verbose or p.returncode):
            sys.stderr.write(out.decode())
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed building libb200lora.so")
    subprocess.check_call([NVCC, "-shared", "-o", OUT, *objs, "-cudart", "static",
                           "-gencode", "arch=compute_100a,code=sm_100a"])
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
