"""Build libmocob200.so (CUDA sm_100a + host C++) in-tree with nvcc.  No torch involved."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libmocob200.so")
HOST_SRCS = ["xml_reader.cpp", "host_model.cpp", "osim_flatten.cpp", "examples.cpp", "layout.cpp", "table_opt.cpp", "capi_host.cpp"]
CUDA_SRCS = ["kernels.cu", "kernels_vel.cu", "kernels_warp.cu", "capi_device.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "mocob200.h"))
    objs, cmds = [], []
    for src in HOST_SRCS + CUDA_SRCS:
        s = os.path.join(CSRC, src)
        o = os.path.join(objdir, src + ".o")
        objs.append(o)
        if force or _newer(o, [s] + headers):
            cmd = [NVCC, "-std=c++17", "-O3", "-lineinfo", "-Xcompiler", "-fPIC", "-c", s, "-o", o]
            if os.environ.get("MOCOB200_PROFILE_PHASES"):  # development stopwatch, never set by build()/the driver
                cmd += ["-DMOCOB200_PROFILE_PHASES"]
            if src.endswith(".cu"):
                cmd += ARCH + ["-Xptxas", "-v"] if verbose else ARCH
            cmds.append(cmd)
    if cmds:  # translation units are independent: compile them side by side
        from concurrent.futures import ThreadPoolExecutor

        def run(cmd):
            if verbose:
                print(" ".join(cmd), flush=True)
            subprocess.check_call(cmd)

        with ThreadPoolExecutor(max_workers=min(len(cmds), os.cpu_count() or 1)) as pool:
            list(pool.map(run, cmds))
    if force or _newer(OUT, objs):
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ARCH + ["-lcudart", "-ldl"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
