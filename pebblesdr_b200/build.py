"""Builds pebblesdr_b200/libpebblecuda.so in-tree: nvcc (sm_100a) for the kernels + C ABI, g++ for the
setup-time host code. Also builds the test-only artefacts under oracle/ (the C restatement and, when
/root/reference is present, the real reference into oracle/_ref) — building the checker is not using it."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpebblecuda.so")
CU_SRCS = ["lib_common.cu", "downconvert.cu", "decimator.cu", "fastfir.cu", "fastfir_long.cu", "spectrum.cu", "demod.cu", "receiver.cu", "channelizer.cu", "chz_poly.cu", "resampler.cu", "cpxutil.cu", "prefilter.cu", "noiseblanker.cu", "rds.cu"]
CPP_SRCS = ["host_design.cpp", "host_window.cpp", "host_demod.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=default"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + "\n")
        raise RuntimeError("build failed: " + cmd[0])
    return r.stdout


def build_lib(force=False, verbose=False):
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "pebble_cuda.h"))
    objs, jobs = [], []
    for s in CU_SRCS + CPP_SRCS:
        src = os.path.join(CSRC, s)
        if not os.path.exists(src):
            continue
        obj = os.path.join(objdir, s + ".o")
        objs.append(obj)
        if force or _newer(obj, [src] + headers):
            if s.endswith(".cu"):
                jobs.append(["nvcc"] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj])
            else:
                jobs.append(["g++", "-O2", "-std=c++17", "-fPIC", "-c", src, "-o", obj])
    if jobs:   # translation units are independent: compile them side by side
        from concurrent.futures import ThreadPoolExecutor
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            for out in ex.map(_run, jobs):
                if verbose:
                    print(out)
    if force or _newer(LIB, objs):
        _run(["nvcc", "-shared", "-o", LIB] + objs + ["-lquadmath"])
    # stand-alone FP64-peak microbenchmark (tools/measure_fp64_peak.py runs it on the GPU box)
    peak_src, peak_exe = os.path.join(ROOT, "tools", "fp64_peak.cu"), os.path.join(ROOT, "tools", "fp64_peak")
    if os.path.exists(peak_src) and (force or _newer(peak_exe, [peak_src])):
        _run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", peak_exe, peak_src])
    return LIB


def build_oracle():
    """C restatement (oracle/libpebbleoracle.so) and, in the build container, the real reference."""
    odir = os.path.join(ROOT, "oracle")
    src = os.path.join(odir, "pebble_oracle.c")
    out = os.path.join(odir, "libpebbleoracle.so")
    if os.path.exists(src) and _newer(out, [src]):
        _run(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", out, src, "-lm"])
    ref = os.path.join(odir, "_ref", "libpebbleref.so")
    ref_long = os.path.join(odir, "_ref", "libpebbleref_long.so")   # config 5 sizes (131072/65537)
    others = [os.path.join(odir, "_ref", n) for n in ("libpebbleref_app.so", "libpebbleref_dec12.so", "libpebbleref_cute.so")]
    deps = [os.path.join(odir, "ref_shim.cpp"), os.path.join(odir, "ref_app_shim.cpp"), os.path.join(odir, "build_ref.sh")]
    deps += [os.path.join(odir, "qt_shim_app", f) for f in os.listdir(os.path.join(odir, "qt_shim_app"))]
    if os.path.isdir("/root/reference/pebblelib") and any(_newer(t, deps) for t in [ref, ref_long] + others):
        _run(["bash", os.path.join(odir, "build_ref.sh")])


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose="-v" in sys.argv))
    build_oracle()
