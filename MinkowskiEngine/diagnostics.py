"""`ME.print_diagnostics` (reference: diagnostics.py:22-66): environment report."""
import os
import platform
import subprocess
import sys


def print_diagnostics():
    print("==========System==========")
    print(platform.platform())
    print(sys.version)
    print("==========Pytorch==========")
    try:
        import torch
        print(torch.__version__)
        print(f"torch.cuda.is_available(): {torch.cuda.is_available()}")
    except ImportError:
        print("torch not installed")
    print("==========NVIDIA-SMI==========")
    try:
        print(subprocess.run(["nvidia-smi"], capture_output=True, text=True).stdout)
    except OSError:
        print("nvidia-smi not found")
    print("==========NVCC==========")
    try:
        print(subprocess.run(["nvcc", "--version"], capture_output=True, text=True).stdout)
    except OSError:
        print("nvcc not found")
    print("==========MinkowskiEngine (B200 engine)==========")
    import MinkowskiEngine as ME
    from mink_octtree_stablediffusion_b200 import _lib
    print(ME.__version__)
    print(f"library: {_lib.LIB_PATH} exists={os.path.exists(_lib.LIB_PATH)}")
    print(f"MinkowskiEngine compiled with CUDA Support: {ME.is_cuda_available()}")
    print(f"NVCC version MinkowskiEngine is compiled: {ME.cuda_version()}")
    print(f"CUDART version MinkowskiEngine is compiled: {ME.cudart_version()}")


if __name__ == "__main__":
    print_diagnostics()
