"""Build libdba.so (the C-ABI library with the sm_100a kernels) in-tree with nvcc.

`python -m cut3r_slam_b200.build` or `__graft_entry__.build()`.  nvcc cross-compiles sm_100a without
a GPU; the resulting cut3r_slam_b200/libdba.so is git-ignored but travels to the GPU box.
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libdba.so")
STAMP = os.path.join(HERE, "csrc", ".build_stamp")

SOURCES = ["dba_api.cu", "dba_linearize.cu", "dba_linearize_tc.cu", "dba_linearize_v2.cu", "dba_assemble.cu", "dba_chol.cu", "dba_chol_fused.cu", "dba_refine.cu", "dba_trisolve_cluster.cu", "dba_p2p.cu", "dba_pgba.cu", "dba_companion.cu", "dba_jdsa.cu", "dba_glue.cu", "dba_cov.cu", "dba_relpose.cu", "dba_plan.cpp", "dba_graph.cpp"]
HEADERS = ["dba_kernels.h", "dba_math.cuh", "dba_linearize.cuh", "dba_umma.cuh", "dba_plan.h", os.path.join("..", "..", "include", "dba.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O3",
]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest():
    h = hashlib.sha256()
    for f in SOURCES + HEADERS:
        with open(os.path.join(CSRC, f), "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    dig = _digest()
    if not force and os.path.exists(OUT) and os.path.exists(STAMP) and open(STAMP).read().strip() == dig:
        return OUT
    nvcc = _nvcc()
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    objs = []
    for s in SOURCES:
        o = os.path.join(objdir, os.path.splitext(s)[0] + ".o")
        objs.append(o)
        cmd = [nvcc] + NVCC_FLAGS + ["-x", "cu", "-c", os.path.join(CSRC, s), "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((s, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for s, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(out)
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed on {s}")
    link = [nvcc, "-shared", "-o", OUT] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    subprocess.run(link, check=True)
    with open(STAMP, "w") as fh:
        fh.write(dig)
    return OUT


EXT_DIR = os.path.join(HERE, "ext")
EXT_SRC = os.path.join(CSRC, "droid_backends_ext.cpp")
EXT_STAMP = os.path.join(CSRC, ".ext_build_stamp")


def ext_path():
    import sysconfig
    return os.path.join(EXT_DIR, "droid_backends" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_ext(force=False):
    """The compiled `droid_backends` torch extension (csrc/droid_backends_ext.cpp) over libdba, in-tree at
    cut3r_slam_b200/ext/ (import it with that directory on sys.path: the reference's module name).  One g++ call with
    torch's own include / library paths; rpath $ORIGIN/.. finds libdba.so, torch's libraries are already loaded by
    `import torch`."""
    import sysconfig
    from torch.utils import cpp_extension as ce
    build()
    out = ext_path()
    h = hashlib.sha256()
    for f in (EXT_SRC, os.path.join(HERE, "..", "include", "dba.h")):
        with open(f, "rb") as fh:
            h.update(fh.read())
    import torch
    h.update(torch.__version__.encode())
    dig = h.hexdigest()
    if not force and os.path.exists(out) and os.path.exists(EXT_STAMP) and open(EXT_STAMP).read().strip() == dig:
        return out
    os.makedirs(EXT_DIR, exist_ok=True)
    cxx = os.environ.get("CXX", "g++")
    cmd = [cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-DTORCH_EXTENSION_NAME=droid_backends",
           "-DTORCH_API_INCLUDE_EXTENSION_H", "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    for inc in ce.include_paths("cuda") + [sysconfig.get_paths()["include"], os.path.join(HERE, "..", "include")]:
        cmd += ["-I", inc]
    cmd += [EXT_SRC, "-o", out]
    for lp in ce.library_paths("cuda"):
        cmd += ["-L", lp]
    cmd += ["-L", HERE, "-ldba", "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda", "-lcudart",
            "-Wl,-rpath,$ORIGIN/..", "-Wl,--no-as-needed"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout)
        raise RuntimeError("building the droid_backends extension failed")
    with open(EXT_STAMP, "w") as fh:
        fh.write(dig)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    if "--ext" in sys.argv:
        print(build_ext(force="--force" in sys.argv))
