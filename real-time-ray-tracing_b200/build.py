"""Build recipe for the C-ABI shared library (librtrt.so) — explicit nvcc, sm_100a only, in-tree output.

    python real-time-ray-tracing_b200/build.py [--force]

--fmad=false: the bit-exact stages (AABB/Morton/box/triangle tests) must not contract a*b+c into FMA;
the oracle is built with -ffp-contract=off for the same reason.  fmaf() is used explicitly where the
reference calls FMA (src/linearMath.h:41-98).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.environ.get("RTRT_BUILD_OUT") or os.path.join(HERE, "librtrt.so")      # RTRT_BUILD_OUT: A/B variants (experiments), with their own object directory
OBJDIR = os.path.join(HERE, "build" + ("_" + os.path.basename(OUT) if os.environ.get("RTRT_BUILD_OUT") else ""))

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "--fmad=false", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function",
    "-Xptxas", "-v" if os.environ.get("RTRT_PTXAS_V") else "-O3",
] + os.environ.get("RTRT_NVCC_EXTRA", "").split()      # experiment knobs, e.g. RTRT_NVCC_EXTRA="-DRTRT_GATHER_UNROLL=8"


# Per-file override (experiment knob, off by default): RTRT_DENOISE_FMAD=1 compiles the denoiser with --fmad=true like the reference's
# own build.  Measured on B200: 0.500 -> 0.486 ms only, and one of the GPU parity tests failed with it (not investigated further),
# so everything stays --fmad=false.
FMAD_TRUE = {"denoise.cu"} if os.environ.get("RTRT_DENOISE_FMAD", "0") == "1" else set()


def flags_for(src):
    f = list(NVCC_FLAGS)
    if os.path.basename(src) in FMAD_TRUE:
        f[f.index("--fmad=false")] = "--fmad=true"
    return f


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def deps():
    d = sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    d.append(os.path.join(HERE, "..", "include", "rtrt.h"))
    d.append(os.path.abspath(__file__))
    return d


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(p) > t for p in deps())


def build(force=False, verbose=True):
    if not force and not needs_build():
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    os.makedirs(OBJDIR, exist_ok=True)
    for src in sources():
        obj = os.path.join(OBJDIR, os.path.basename(src) + ".o")
        objs.append(obj)
        cmd = [nvcc] + flags_for(src) + ["-x", "cu" if src.endswith(".cu") else "c++", "-c", src, "-o", obj]
        if verbose:
            print(" ".join(cmd), flush=True)
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if out.strip() and verbose:
            print(out)
        if p.returncode != 0:
            failed = True
            print(f"nvcc failed for {src}:\n{out}", file=sys.stderr)
    if failed:
        raise RuntimeError("nvcc compilation failed")
    cmd = [nvcc, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv)
    print("built", OUT)
