"""TEST INFRASTRUCTURE ONLY. Build the reference's own lietorch CPU backend into oracle/_ref/.

The DBA hot path's only executable statement in the reference tree is the pure-torch
hislam2/geom/{ba,chol,projective_ops,pinhole}.py, which stands on lietorch's SE3 ops
(thirdparty/lietorch/lietorch/src/lietorch.cpp, lietorch_cpu.cpp, include/*.h + vendored Eigen).
This recipe compiles those two C++ files WHERE THEY LIE under /root/reference (nothing is copied
into the repo) together with oracle/ref_gpu_stubs.cpp (ours; stands in for lietorch_gpu.cu so the
build needs no nvcc pass) and writes lietorch_backends.so into oracle/_ref/ (git-ignored).

Used by oracle/ref_harness.py (validation of the restatement + golden-vector generation, in the
build container only: /root/reference does not exist on the GPU box).
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("DBA_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")


def build(verbose=False):
    lt = os.path.join(REF, "thirdparty", "lietorch")
    if not os.path.isdir(lt):
        raise FileNotFoundError(f"reference not present at {REF}")
    os.makedirs(OUT, exist_ok=True)
    so = os.path.join(OUT, "lietorch_backends.so")
    if os.path.exists(so):
        return so
    from torch.utils.cpp_extension import load
    os.environ.setdefault("MAX_JOBS", "8")
    load(
        name="lietorch_backends",
        sources=[
            os.path.join(lt, "lietorch", "src", "lietorch.cpp"),
            os.path.join(lt, "lietorch", "src", "lietorch_cpu.cpp"),
            os.path.join(HERE, "ref_gpu_stubs.cpp"),
        ],
        extra_include_paths=[
            os.path.join(lt, "lietorch", "include"),
            os.path.join(REF, "thirdparty", "eigen"),
            "/usr/local/cuda/include",
        ],
        extra_cflags=["-O2"],
        build_directory=OUT,
        with_cuda=False,
        is_python_module=False,
        verbose=verbose,
    )
    return so


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
