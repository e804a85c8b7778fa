"""In-tree build of the CUDA library (sm_100a only): `python -m kdl_b200.build` or `build_lib()`."""
import os
import subprocess

CSRC = os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc")
LIB = os.path.join(CSRC, "libkdl.so")


def build_lib(force: bool = False, verbose: bool = False) -> str:
    """nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo ... (see csrc/Makefile).  Cross-compiles
    without a GPU.  Returns the path of libkdl.so."""
    args = ["make", "-C", CSRC] + (["-B"] if force else [])
    r = subprocess.run(args, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout)
        print(r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc build of libkdl.so failed (see output above)")
    return LIB


if __name__ == "__main__":
    print(build_lib(verbose=True))
