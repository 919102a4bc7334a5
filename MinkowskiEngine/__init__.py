"""MinkowskiEngine-compatible Python API on the B200-native sparse-convolution
engine (`mink_octtree_stablediffusion_b200`).

Export list follows the reference's `MinkowskiEngine/__init__.py:53-221`, so
`import MinkowskiEngine as ME` keeps working for the octree VAE / diffusion
model code.  Continuous-coordinate features (TensorField, interpolation,
channel-wise convolution, SPMM quantisation) are outside this engine's scope
 and raise NotImplementedError when touched.
"""
__version__ = "0.5.4"

import os
import sys

# the reference is a flat-module package: its modules (and third-party model
# code) also import siblings by bare name, e.g. `from MinkowskiSparseTensor import ...`
_file_dir = os.path.dirname(__file__)
_repo_root = os.path.dirname(_file_dir)
if _repo_root not in sys.path:
    sys.path.insert(0, _repo_root)

import torch  # noqa: E402,F401

from MinkowskiEngineBackend._C import (  # noqa: E402,F401
    BroadcastMode, ConvolutionMode, CoordinateMapKey, CoordinateMapType, GPUMemoryAllocatorType,
    MinkowskiAlgorithm, PoolingMode, RegionType, cuda_version, cudart_version, get_gpu_memory_info,
    is_cuda_available)

from .diagnostics import print_diagnostics  # noqa: E402,F401
from .MinkowskiKernelGenerator import KernelGenerator, KernelRegion, get_kernel_volume  # noqa: E402,F401
from .MinkowskiTensor import (  # noqa: E402,F401
    SparseTensorOperationMode, SparseTensorQuantizationMode, clear_global_coordinate_manager,
    global_coordinate_manager, set_global_coordinate_manager, set_sparse_tensor_operation_mode,
    sparse_tensor_operation_mode)
from .MinkowskiSparseTensor import SparseTensor  # noqa: E402,F401
from .MinkowskiCommon import MinkowskiModuleBase, convert_to_int_tensor  # noqa: E402,F401
from .MinkowskiCoordinateManager import (  # noqa: E402,F401
    CoordinateManager, CoordsManager, set_gpu_allocator, set_memory_manager_backend)
from .MinkowskiConvolution import (  # noqa: E402,F401
    MinkowskiConvolution, MinkowskiConvolutionFunction, MinkowskiConvolutionTranspose,
    MinkowskiConvolutionTransposeFunction, MinkowskiGenerativeConvolutionTranspose,
    MinkowskiUpsampleInterpolate)
from .MinkowskiPooling import (  # noqa: E402,F401
    MinkowskiAvgPooling, MinkowskiGlobalAvgPooling, MinkowskiGlobalMaxPooling, MinkowskiGlobalPooling,
    MinkowskiGlobalPoolingFunction, MinkowskiGlobalSumPooling, MinkowskiLocalPoolingFunction,
    MinkowskiLocalPoolingTransposeFunction, MinkowskiMaxPooling, MinkowskiPoolingTranspose,
    MinkowskiSumPooling)
from .MinkowskiBroadcast import (  # noqa: E402,F401
    MinkowskiBroadcast, MinkowskiBroadcastAddition, MinkowskiBroadcastConcatenation,
    MinkowskiBroadcastFunction, MinkowskiBroadcastMultiplication)
from .MinkowskiNonlinearity import *  # noqa: E402,F401,F403
from .MinkowskiNonlinearity import MinkowskiSinusoidal  # noqa: E402,F401
from .MinkowskiNormalization import (  # noqa: E402,F401
    AdaStableInstanceNorm, HjmGroupNorm, HjmInstanceNorm, MinkowskiBatchNorm, MinkowskiInstanceNorm,
    MinkowskiStableGroupNorm, MinkowskiStableInstanceNorm, MinkowskiSyncBatchNorm)
from .MinkowskiPruning import MinkowskiPruning, MinkowskiPruningFunction  # noqa: E402,F401
from .MinkowskiUnion import MinkowskiUnion  # noqa: E402,F401
from .MinkowskiNetwork import MinkowskiNetwork  # noqa: E402,F401
from . import MinkowskiOps  # noqa: E402,F401
from .MinkowskiOps import (  # noqa: E402,F401
    MinkowskiLinear, MinkowskiStackCat, MinkowskiStackMean, MinkowskiStackSum, MinkowskiStackVar,
    MinkowskiToDenseTensor, MinkowskiToFeature, MinkowskiToSparseTensor, cat, dense_coordinates, mean,
    to_sparse, to_sparse_all, var)
from .MinkowskiOps import _sum as sum  # noqa: E402,F401
from . import MinkowskiFunctional  # noqa: E402,F401
from . import utils  # noqa: E402,F401
from . import modules  # noqa: E402,F401


def _out_of_scope(name):
    class _Missing:
        def __init__(self, *args, **kwargs):
            raise NotImplementedError(f"ME.{name} is outside this engine's scope (SURVEY.md §2.1 row 5)")
    _Missing.__name__ = name
    return _Missing


TensorField = _out_of_scope("TensorField")
MinkowskiInterpolation = _out_of_scope("MinkowskiInterpolation")
MinkowskiChannelwiseConvolution = _out_of_scope("MinkowskiChannelwiseConvolution")
