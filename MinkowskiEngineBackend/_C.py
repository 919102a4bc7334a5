"""`MinkowskiEngineBackend._C` — same module-level names as the pybind module
defined at pybind/minkowski.cu:36-68 / pybind/extern.hpp:579-838, implemented
by mink_octtree_stablediffusion_b200 (CUDA sm_100a, C-ABI library)."""
from mink_octtree_stablediffusion_b200.backend import (  # noqa: F401
    BroadcastMode, ConvolutionMode, CoordinateMapKey, CoordinateMapManagerGPU_c10, CoordinateMapType,
    CUDAKernelMapMode, GPUMemoryAllocatorType, MinkowskiAlgorithm, PoolingMode, RegionType,
    cuda_version, cudart_version, get_gpu_memory_info, is_cuda_available)
from mink_octtree_stablediffusion_b200.backend_ops import (  # noqa: F401
    ConvolutionBackwardGPU, ConvolutionForwardGPU, ConvolutionTransposeBackwardGPU,
    ConvolutionTransposeForwardGPU, CoordinateMapManagerCPU, CoordinateMapManagerGPU_default)
from mink_octtree_stablediffusion_b200.backend_pool import (  # noqa: F401
    BroadcastBackwardGPU, BroadcastForwardGPU, GlobalPoolingBackwardGPU, GlobalPoolingForwardGPU,
    LocalPoolingBackwardGPU, LocalPoolingForwardGPU, LocalPoolingTransposeBackwardGPU,
    LocalPoolingTransposeForwardGPU, PruningBackwardGPU, PruningForwardGPU)

# pybind enums are created with .export_values(): members are module attributes too
for _enum in (GPUMemoryAllocatorType, CUDAKernelMapMode, MinkowskiAlgorithm, CoordinateMapType, RegionType,
              PoolingMode, BroadcastMode, ConvolutionMode):
    for _member in _enum:
        globals().setdefault(_member.name, _member)
del _enum, _member
