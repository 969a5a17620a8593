"""oracle/build_ref.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Compiles the UNMODIFIED reference WKV6 CUDA kernels from where they lie
(/root/reference/wenet/rwkv_v6/cuda/wkv6_cuda.cu) for sm_100a, with the reference's own JIT flags
(wenet/rwkv_v6/src/model.py:105-106: --use_fast_math -O3 -Xptxas -O3 --extra-device-vectorization
-D_N_=64 -D_T_=<ctx_len>), plus oracle/ref_exports.cpp (extern "C" doors), into
oracle/_ref/libwkv6_ref.so.  The .so is git-ignored but travels to the GPU box with gpurun, where
tests/golden/make_refcuda_golden.py and tests/test_gpu_vs_reference_cuda.py run it.
Nothing is copied from the reference tree; without /root/reference this script is a no-op.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_CU = "/root/reference/wenet/rwkv_v6/cuda/wkv6_cuda.cu"
OUT_DIR = os.path.join(HERE, "_ref")
OUT = os.path.join(OUT_DIR, "libwkv6_ref.so")
CTX_LEN = 4096          # _T_: upper bound on T for the reference backward (wkv6_cuda.cu:207)


def main() -> int:
    if not os.path.exists(REF_CU):
        print("build_ref: reference tree absent, nothing to do")
        return 0
    os.makedirs(OUT_DIR, exist_ok=True)
    from torch.utils import cpp_extension as ce
    inc = []
    for p in ce.include_paths():
        inc += ["-isystem", p]
    nvcc = "/usr/local/cuda/bin/nvcc" if os.path.exists("/usr/local/cuda/bin/nvcc") else "nvcc"
    obj_ref = os.path.join(OUT_DIR, "wkv6_cuda_ref.o")
    obj_exp = os.path.join(OUT_DIR, "ref_exports.o")
    common = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-Xcompiler", "-fPIC", *inc]
    ref_flags = ["--use_fast_math", "-O3", "-Xptxas", "-O3", "--extra-device-vectorization",
                 "-D_N_=64", f"-D_T_={CTX_LEN}"]
    subprocess.run([nvcc, *common, *ref_flags, "-c", REF_CU, "-o", obj_ref], check=True)
    subprocess.run([nvcc, *common, f"-DWKV6_REF_T={CTX_LEN}", "-x", "cu", "-c", os.path.join(HERE, "ref_exports.cpp"),
                    "-o", obj_exp], check=True)
    subprocess.run([nvcc, "-shared", "-o", OUT, obj_ref, obj_exp, "-gencode", "arch=compute_100a,code=sm_100a",
                    "-lcudart"], check=True)
    print("built", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
