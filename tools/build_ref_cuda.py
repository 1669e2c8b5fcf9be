#!/usr/bin/env python
"""Build the reference's OWN CUDA extension (pdvc/ops/src, unmodified kernels) for sm_100 as the same-box GPU comparison
(SURVEY 8d "Same-box GPU comparison").  Runs in the build container only: the sources are compiled where they lie in
/root/reference through a scratch copy under /tmp that carries the 2-token torch-2.x patch of SURVEY finding 0.7
(`value.type()` -> `value.scalar_type()` at ms_deform_attn_cuda.cu:64,:134); nothing of the reference is copied into the
repo.  Output: baseline/_ref/ref_cuda/MultiScaleDeformableAttention_refcuda.so (git-ignored, travels to the GPU box).

bench.py --ref-cuda loads that .so (if present) and times ms_deform_attn_forward/backward on the unfused boundary."""
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/pdvc/ops/src"
OUT = os.path.join(ROOT, "baseline", "_ref", "ref_cuda")
NAME = "MultiScaleDeformableAttention_refcuda"


def main():
    if not os.path.isdir(SRC):
        print("reference sources not mounted; nothing to build")
        return 0
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    scratch = "/tmp/ref_cuda_src"
    shutil.rmtree(scratch, ignore_errors=True)
    shutil.copytree(SRC, scratch)
    cu = os.path.join(scratch, "cuda", "ms_deform_attn_cuda.cu")
    with open(cu) as f:
        text = f.read()
    old, new = "AT_DISPATCH_FLOATING_TYPES(value.type()", "AT_DISPATCH_FLOATING_TYPES(value.scalar_type()"
    assert text.count(old) == 2, "unexpected reference source"
    with open(cu, "w") as f:
        f.write(text.replace(old, new))
    os.makedirs(OUT, exist_ok=True)
    from torch.utils.cpp_extension import load
    load(name=NAME,
         sources=[os.path.join(scratch, "vision.cpp"), os.path.join(scratch, "cpu", "ms_deform_attn_cpu.cpp"), cu],
         extra_include_paths=[scratch], extra_cflags=["-DWITH_CUDA"],
         extra_cuda_cflags=["-DWITH_CUDA", "-DCUDA_HAS_FP16=1", "-D__CUDA_NO_HALF_OPERATORS__",
                            "-D__CUDA_NO_HALF_CONVERSIONS__", "-D__CUDA_NO_HALF2_OPERATORS__", "-lineinfo"],
         build_directory=OUT, is_python_module=False, verbose=False)
    so = os.path.join(OUT, NAME + ".so")
    assert os.path.exists(so), so
    for junk in os.listdir(OUT):
        if not junk.endswith(".so"):
            p = os.path.join(OUT, junk)
            shutil.rmtree(p) if os.path.isdir(p) else os.remove(p)
    print("built", so)
    return 0


if __name__ == "__main__":
    sys.exit(main())
