"""Builds reportree_b200/libreportree_hc.so with nvcc for sm_100a (in-tree, no torch needed)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libreportree_hc.so")
SOURCES = ["api.cu", "recode.cu", "collide.cu", "distance.cu", "distance_ws.cu", "mst.cu", "cut.cu", "linkage.cu", "nnchain.cu", "probe.cu", "zoom.cu", "stability.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
         "-Xcompiler", "-fPIC", "-Xcompiler", "-O3", "-fmad=false"]
FLAGS += os.environ.get("RTHC_EXTRA_NVCC_FLAGS", "").split()


HOSTIO_SRC = os.path.join(CSRC, "hostio.cpp")
HOSTIO_OUT = os.path.join(HERE, "librthc_hostio.so")
CXX = os.environ.get("CXX", "g++")


def build_hostio(force=False):
    """librthc_hostio.so: plain C++ host text I/O (include/rthio.h), no CUDA involved."""
    deps = [HOSTIO_SRC, os.path.join(HERE, "..", "include", "rthio.h")]
    if not force and os.path.exists(HOSTIO_OUT) and all(os.path.getmtime(d) <= os.path.getmtime(HOSTIO_OUT) for d in deps):
        return HOSTIO_OUT
    subprocess.check_call([CXX, "-O3", "-std=c++17", "-fPIC", "-shared", "-pthread", "-Wall", "-o", HOSTIO_OUT, HOSTIO_SRC])
    return HOSTIO_OUT


CGMLST_SRC = os.path.join(CSRC, "cgmlst_dists_main.cpp")
CGMLST_OUT = os.path.join(HERE, "bin", "cgmlst-dists")


def build_cgmlst_dists(force=False):
    """bin/cgmlst-dists: the external program's CLI (seam B2) on top of libreportree_hc.so."""
    deps = [CGMLST_SRC, OUT]
    if not force and os.path.exists(CGMLST_OUT) and all(os.path.getmtime(d) <= os.path.getmtime(CGMLST_OUT) for d in deps):
        return CGMLST_OUT
    os.makedirs(os.path.dirname(CGMLST_OUT), exist_ok=True)
    subprocess.check_call([CXX, "-O3", "-std=c++17", "-pthread", "-Wall", "-o", CGMLST_OUT, CGMLST_SRC,
                           "-L" + HERE, "-lreportree_hc", "-Wl,-rpath,$ORIGIN/..", "-ldl", "-lrt"])
    return CGMLST_OUT


def needs_build():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f not in ("hostio.cpp", "cgmlst_dists_main.cpp")] + [os.path.join(HERE, "..", "include", "rthc.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    build_hostio(force)
    if not force and not needs_build():
        build_cgmlst_dists(force)
        return OUT
    objs = []
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, "build", src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            sys.stderr.write(out)
        if p.returncode:
            raise RuntimeError("nvcc failed on " + src)
    # link next to the target and rename: a snapshot (gpurun) or a concurrent dlopen never sees a partial library
    subprocess.check_call([NVCC, "-shared", "-o", OUT + ".tmp"] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                                             "-cudart", "static"])
    os.replace(OUT + ".tmp", OUT)
    build_cgmlst_dists(True)
    return OUT


def build_checked():
    """build/lib_checked.so: the same sources with -DRTHC_CHECKED (device-side invariants, see rthc_internal.cuh);
    load it with RTHC_LIB_PATH=reportree_b200/build/lib_checked.so."""
    out = os.path.join(HERE, "build", "lib_checked.so")
    odir = os.path.join(HERE, "build", "checked")
    os.makedirs(odir, exist_ok=True)
    procs, objs = [], []
    for src in SOURCES:
        obj = os.path.join(odir, src.replace(".cu", ".o"))
        procs.append((src, subprocess.Popen([NVCC] + FLAGS + ["-DRTHC_CHECKED", "-c", os.path.join(CSRC, src), "-o", obj],
                                            stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    for src, p in procs:
        text, _ = p.communicate()
        if p.returncode:
            sys.stderr.write(text)
            raise RuntimeError("nvcc failed on " + src)
    subprocess.check_call([NVCC, "-shared", "-o", out + ".tmp"] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                                             "-cudart", "static"])
    os.replace(out + ".tmp", out)
    return out


if __name__ == "__main__":
    if "--checked" in sys.argv:
        print(build_checked())
    else:
        build(force=True, verbose="-v" in sys.argv)
        print(OUT)
