"""nbody_b200 -- B200-native force-evaluation + Velocity-Verlet path behind LessUp/n-body's
ForceCalculator / ParticleSystem / SimulationConfig API.

Layout:  csrc/        hand-written CUDA for sm_100a + the C ABI (include/nbody_b200.h)
         _lib.py      ctypes loader (fails loudly when libnbody_b200.so is missing)
         api.py       Python mirror of the reference's classes over the C ABI
         ics.py       synthetic initial conditions (Plummer, uniform sphere/box, disk)
         serialization.py  the reference's .nbody checkpoint format
         render.py    headless Camera / ColorMapper / Renderer over the CUDA splat kernels
         dist.py      one-process-per-GPU target sharding for direct N^2 (torch.distributed)
"""
from .api import (  # noqa: F401
    BarnesHutCalculator, Context, CudaException, CudaGLInterop, DirectForceCalculator, ForceCalculator, ForceMethod,
    InitDistribution, Integrator, ParticleData, ParticleDataManager, ParticleSystem, ResourceException,
    SimulationConfig, SimulationState, SpatialHashCalculator, ValidationException, createForceCalculator,
    get_context, validateSimulationConfig,
)
from .serialization import Serializer  # noqa: F401
from .render import Camera, ColorMapper, ColorMode, Renderer  # noqa: F401

__version__ = "0.1.0"
