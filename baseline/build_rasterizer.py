#!/usr/bin/env python
"""Builds the reference's UNCHANGED diff-gaussian-rasterization (the consumer of the solver's output,
main.py:108-157) for sm_100a into baseline/_ref/ (git-ignored; travels to the GPU box with gpurun).

Nothing of it is copied into the repository: the sources are compiled where they lie under
/root/reference/gaussian_splatting/submodules/diff-gaussian-rasterization, through a scratch copy under /tmp because
setuptools writes build products next to the sources and /root/reference is read-only.  The one deviation from the
submodule's setup.py: `-include cstdint` / `-include cfloat` on the nvcc command line -- rasterizer_impl.h uses
uint32_t / uint64_t and FLT_MAX without including the headers that declare them, which gcc >= 13 no longer lets
through transitively.  No source file is edited.

    python baseline/build_rasterizer.py          # -> baseline/_ref/diff_gaussian_rasterization/{__init__.py,_C*.so}
"""
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = "/root/reference/gaussian_splatting/submodules/diff-gaussian-rasterization"
OUT = os.path.join(ROOT, "baseline", "_ref")

SETUP = r'''
import os
from setuptools import setup
from torch.utils.cpp_extension import CUDAExtension, BuildExtension
here = os.path.dirname(os.path.abspath(__file__))
setup(name="diff_gaussian_rasterization", packages=["diff_gaussian_rasterization"],
      ext_modules=[CUDAExtension(name="diff_gaussian_rasterization._C",
                                 sources=["cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu", "cuda_rasterizer/backward.cu",
                                          "rasterize_points.cu", "ext.cpp"],
                                 extra_compile_args={"nvcc": ["-I" + os.path.join(here, "third_party/glm/"), "-include", "cstdint", "-include", "cfloat",
                                                              "-gencode", "arch=compute_100a,code=sm_100a"],
                                                     "cxx": ["-include", "cstdint"]})],
      cmdclass={"build_ext": BuildExtension})
'''


def built():
    d = os.path.join(OUT, "diff_gaussian_rasterization")
    return os.path.isdir(d) and any(f.startswith("_C") and f.endswith(".so") for f in os.listdir(d))


def main():
    if built() and "--force" not in sys.argv:
        print("baseline/_ref/diff_gaussian_rasterization already built")
        return 0
    if not os.path.isdir(SRC):
        print("reference tree not present (GPU box): using the prebuilt baseline/_ref if any")
        return 0 if built() else 1
    tmp = tempfile.mkdtemp(prefix="dgr_build_")
    work = os.path.join(tmp, "dgr")
    shutil.copytree(SRC, work)
    with open(os.path.join(work, "setup_b200.py"), "w") as f:
        f.write(SETUP)
    env = dict(os.environ)
    env["TORCH_CUDA_ARCH_LIST"] = "10.0a"      # torch's own arch flags; the explicit -gencode above is the one that matters
    env.setdefault("MAX_JOBS", "8")
    os.makedirs(OUT, exist_ok=True)
    subprocess.check_call([sys.executable, "setup_b200.py", "build_ext", "--build-lib", OUT, "--build-temp", os.path.join(tmp, "obj")], cwd=work, env=env)
    # the package's Python wrapper (autograd.Function + settings tuple) next to the extension, as an install would place it
    shutil.copy(os.path.join(work, "diff_gaussian_rasterization", "__init__.py"), os.path.join(OUT, "diff_gaussian_rasterization", "__init__.py"))
    shutil.rmtree(tmp, ignore_errors=True)
    print("built", os.path.join(OUT, "diff_gaussian_rasterization"))
    return 0


if __name__ == "__main__":
    sys.exit(main())
