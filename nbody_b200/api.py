"""Python mirror of the reference's hot-path API, over the C ABI (include/nbody_b200.h).

Same class and method names, argument meaning and error behaviour as the reference's
`ForceCalculator` strategy family (include/nbody/force_calculator.hpp), `Integrator`
(include/nbody/integrator.hpp), `ParticleData`/`SimulationConfig` (include/nbody/types.hpp),
`ParticleDataManager` (include/nbody/particle_data.hpp) and `ParticleSystem`
(include/nbody/particle_system.hpp), so the parity tests read like the reference's own tests.
The C++20 twin of this file is include/nbody/*.hpp.

PyTorch is plumbing only: device tensors own the 13 SoA arrays and supply the stream handle.
Every compute call goes to the CUDA extension; there is no CPU or torch fallback.
"""
from __future__ import annotations

import ctypes as C
import enum
import math
from dataclasses import dataclass, replace

import numpy as np

from . import _lib
from ._lib import CudaException, NbParticles, ResourceException, ValidationException, check  # noqa: F401
from .ics import FIELDS


class ForceMethod(enum.IntEnum):  # include/nbody/types.hpp:66-70
    DIRECT_N2 = 0
    BARNES_HUT = 1
    SPATIAL_HASH = 2


class InitDistribution(enum.IntEnum):  # include/nbody/types.hpp:82-86
    UNIFORM = 0
    SPHERICAL = 1
    DISK = 2


@dataclass
class SimulationConfig:  # include/nbody/types.hpp:301-313
    particle_count: int = 10000
    init_distribution: InitDistribution = InitDistribution.SPHERICAL
    force_method: ForceMethod = ForceMethod.DIRECT_N2
    dt: float = 0.001
    G: float = 1.0
    softening: float = 0.1
    barnes_hut_theta: float = 0.5
    spatial_hash_cell_size: float = 1.0
    spatial_hash_cutoff: float = 2.0
    cuda_block_size: int = 256


# ---- validation: src/utils/error_handling.cpp:46-123 -------------------------------------------
def _finite(x):
    return not (math.isnan(x) or math.isinf(x))


def validateParticleCountRange(count):
    if count == 0:
        raise ValidationException("Particle count must be greater than 0")
    if count > 100000000:
        raise ValidationException("Particle count exceeds maximum supported (100M)")


def validateTimeStep(dt):
    if dt <= 0:
        raise ValidationException("Time step must be positive")
    if not _finite(dt):
        raise ValidationException("Time step must be a finite number")
    if dt > 1.0:
        raise ValidationException("Time step is too large (max 1.0)")


def validateSoftening(eps):
    if eps < 0:
        raise ValidationException("Softening parameter must be non-negative")
    if not _finite(eps):
        raise ValidationException("Softening parameter must be a finite number")


def validateTheta(theta):
    if theta < 0 or theta > 2.0:
        raise ValidationException("Barnes-Hut theta must be between 0 and 2")
    if not _finite(theta):
        raise ValidationException("Barnes-Hut theta must be a finite number")


def validateSimulationConfig(cfg: SimulationConfig):
    validateParticleCountRange(cfg.particle_count)
    validateTimeStep(cfg.dt)
    validateSoftening(cfg.softening)
    if cfg.force_method == ForceMethod.BARNES_HUT:
        validateTheta(cfg.barnes_hut_theta)
    if cfg.G <= 0 or not _finite(cfg.G):
        raise ValidationException("Gravitational constant must be positive and finite")
    if cfg.force_method == ForceMethod.SPATIAL_HASH:
        if cfg.spatial_hash_cell_size <= 0 or not _finite(cfg.spatial_hash_cell_size):
            raise ValidationException("Spatial hash cell size must be positive and finite")
        if cfg.spatial_hash_cutoff <= 0 or not _finite(cfg.spatial_hash_cutoff):
            raise ValidationException("Spatial hash cutoff must be positive and finite")
    if cfg.cuda_block_size <= 0 or cfg.cuda_block_size > 1024:
        raise ValidationException("CUDA block size must be between 1 and 1024")


def validateResourceRequirements(particle_count, device=0):
    """src/utils/error_handling.cpp:25-44: 2 x 52 B/body must fit in 80 % of device memory."""
    total = C.c_size_t(0)
    check(_lib.lib().nb_device_info(device, None, None, None, C.byref(total), None))
    required = particle_count * 4 * 13 * 2
    if required > total.value * 0.8:
        raise ResourceException(f"Insufficient GPU memory - Required: {required} bytes, "
                                f"Available: {int(total.value * 0.8)} bytes")


# ---- context ----------------------------------------------------------------------------------
_contexts: dict = {}


class Context:
    """One nb_context per device (owns all scratch)."""

    def __init__(self, device: int = 0):
        self.device = device
        h = C.c_void_p()
        check(_lib.lib().nb_context_create(device, 0, C.byref(h)))
        self.handle = h

    def close(self):
        if self.handle:
            _lib.lib().nb_context_destroy(self.handle)
            self.handle = None

    @property
    def sm_count(self):
        return _lib.lib().nb_context_sm_count(self.handle)

    @property
    def launch_count(self):
        return int(_lib.lib().nb_launch_count(self.handle))


def get_context(device: int | None = None) -> Context:
    if device is None:
        import torch
        device = torch.cuda.current_device()
    ctx = _contexts.get(device)
    if ctx is None:
        ctx = _contexts[device] = Context(device)
    return ctx


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


# ---- ParticleData -------------------------------------------------------------------------------
class ParticleData:
    """13-array SoA (include/nbody/types.hpp:234-276).  Arrays are torch CUDA tensors (device
    data) or numpy arrays (host data); `.c` is the matching nb_particles / ParticleData struct."""

    def __init__(self):
        self.arrays: dict = {}
        self.count = 0
        self.on_device = False
        self.c = NbParticles()

    def _refresh(self):
        for k in FIELDS:
            a = self.arrays.get(k)
            if a is None:
                ptr = None
            elif self.on_device:
                ptr = a.data_ptr()
            else:
                ptr = a.ctypes.data
            setattr(self.c, k, ptr)
        self.c.count = self.count

    def __getattr__(self, name):
        arrays = self.__dict__.get("arrays", {})
        if name in arrays:
            return arrays[name]
        raise AttributeError(name)

    def numpy(self) -> dict:
        if self.on_device:
            return {k: v.detach().cpu().numpy() for k, v in self.arrays.items()}
        return {k: np.array(v, copy=True) for k, v in self.arrays.items()}


class ParticleDataManager:
    """src/cuda/particle_init.cu:143-283."""

    @staticmethod
    def allocateDevice(data: ParticleData, count: int, device=None):
        import torch
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        data.arrays = {k: torch.zeros(max(count, 1), dtype=torch.float32, device=dev)[:count] for k in FIELDS}
        data.count = count
        data.on_device = True
        data._refresh()

    @staticmethod
    def freeDevice(data: ParticleData):
        data.arrays = {}
        data.count = 0
        data._refresh()

    @staticmethod
    def allocateHost(data: ParticleData, count: int, pinned: bool = False):
        if pinned:
            import torch
            data._pinned = {k: torch.zeros(max(count, 1), dtype=torch.float32).pin_memory() for k in FIELDS}
            data.arrays = {k: v.numpy()[:count] for k, v in data._pinned.items()}
        else:
            data.arrays = {k: np.zeros(count, np.float32) for k in FIELDS}
        data.count = count
        data.on_device = False
        data._refresh()

    @staticmethod
    def freeHost(data: ParticleData):
        data.arrays = {}
        data.count = 0
        data._refresh()

    @staticmethod
    def fromState(state: dict, pinned: bool = False) -> ParticleData:
        h = ParticleData()
        n = state["pos_x"].size
        ParticleDataManager.allocateHost(h, n, pinned)
        for k in FIELDS:
            if k in state:
                h.arrays[k][:] = state[k]
        return h

    @staticmethod
    def copyToDevice(d_data: ParticleData, h_data: ParticleData):
        """pos, vel, acc, mass -- acc_old is not copied (particle_init.cu:257-269)."""
        ctx = get_context()
        check(_lib.lib().nb_particles_upload(ctx.handle, C.byref(d_data.c), C.byref(h_data.c), _stream()))
        check(_lib.lib().nb_synchronize(ctx.handle, _stream()))

    @staticmethod
    def copyToHost(h_data: ParticleData, d_data: ParticleData):
        ctx = get_context()
        check(_lib.lib().nb_particles_download(ctx.handle, C.byref(h_data.c), C.byref(d_data.c), _stream()))

    @staticmethod
    def copyPositionsToHost(h_pos_x, h_pos_y, h_pos_z, d_data: ParticleData):
        """Positions only, device -> three host float32 arrays (particle_data.hpp:29-31)."""
        ctx, nbytes = get_context(), d_data.count * 4
        for dst, key in ((h_pos_x, "pos_x"), (h_pos_y, "pos_y"), (h_pos_z, "pos_z")):
            assert dst.dtype == np.float32 and dst.flags["C_CONTIGUOUS"] and dst.size >= d_data.count
            check(_lib.lib().nb_device_download(ctx.handle, dst.ctypes.data_as(C.c_void_p),
                                                C.c_void_p(getattr(d_data.c, key)), nbytes, _stream()))

    @staticmethod
    def copyPositionsToDevice(d_data: ParticleData, h_pos_x, h_pos_y, h_pos_z):
        """Positions only, three host float32 arrays -> device (particle_data.hpp:33-35)."""
        ctx, nbytes = get_context(), d_data.count * 4
        for src, key in ((h_pos_x, "pos_x"), (h_pos_y, "pos_y"), (h_pos_z, "pos_z")):
            src = np.ascontiguousarray(src, np.float32)
            assert src.size >= d_data.count
            check(_lib.lib().nb_device_upload(ctx.handle, C.c_void_p(getattr(d_data.c, key)),
                                              src.ctypes.data_as(C.c_void_p), nbytes, _stream()))


# ---- force calculators ----------------------------------------------------------------------------
class ForceCalculator:
    """include/nbody/force_calculator.hpp:36-89 (defaults eps = 0.01, G = 1)."""

    def __init__(self):
        self.softening_eps_ = np.float32(0.01)
        self.softening_eps2_ = np.float32(0.0001)
        self.G_ = np.float32(1.0)

    def computeForces(self, d_particles: ParticleData):
        raise NotImplementedError

    def getMethod(self) -> ForceMethod:
        raise NotImplementedError

    def setSofteningParameter(self, eps):
        self.softening_eps_ = np.float32(eps)
        self.softening_eps2_ = np.float32(self.softening_eps_ * self.softening_eps_)

    def setGravitationalConstant(self, G):
        self.G_ = np.float32(G)

    def getSofteningParameter(self):
        return float(self.softening_eps_)

    def getGravitationalConstant(self):
        return float(self.G_)


class DirectForceCalculator(ForceCalculator):
    """include/nbody/force_calculator.hpp:99-121; computeForces = src/cuda/force_direct.cu:103-106."""

    def __init__(self, block_size: int = 256):
        super().__init__()
        self.block_size_ = block_size  # accepted for API compatibility; a hint only

    def computeForces(self, d_particles: ParticleData):
        ctx = get_context()
        check(_lib.lib().nb_direct_forces(ctx.handle, C.byref(d_particles.c), float(self.G_),
                                          float(self.softening_eps2_), _stream()))

    def getMethod(self):
        return ForceMethod.DIRECT_N2

    def setBlockSize(self, size):
        self.block_size_ = size

    def getBlockSize(self):
        return self.block_size_


class BarnesHutTreeView:
    """What BarnesHutCalculator::getTree() exposes, read back from the device."""

    def __init__(self, n):
        self.n = n

    def export(self):
        ctx = get_context()
        n = self.n
        ni = max(n - 1, 1)
        out = dict(keys=np.zeros(n, np.uint32), perm=np.zeros(n, np.int32), left=np.full(ni, -1, np.int32),
                   right=np.full(ni, -1, np.int32), parent=np.full(2 * n - 1, -1, np.int32),
                   prefix=np.zeros(ni, np.int32), node_com=np.zeros((2 * n - 1, 4), np.float32),
                   bbox=np.zeros(6, np.float32))
        check(_lib.lib().nb_bh_export(ctx.handle, *[out[k].ctypes.data_as(C.c_void_p) for k in
                                                    ("keys", "perm", "left", "right", "parent", "prefix",
                                                     "node_com", "bbox")]))
        return out

    def getNodeCount(self):
        cnt, inter = C.c_int(0), C.c_uint64(0)
        check(_lib.lib().nb_bh_stats(get_context().handle, C.byref(inter), C.byref(cnt)))
        return cnt.value

    def interactions(self):
        cnt, inter = C.c_int(0), C.c_uint64(0)
        check(_lib.lib().nb_bh_stats(get_context().handle, C.byref(inter), C.byref(cnt)))
        return inter.value


class BarnesHutCalculator(ForceCalculator):
    """include/nbody/force_calculator.hpp:130-166; computeForces = src/cuda/force_barnes_hut.cu:526-532."""

    def __init__(self, theta: float = 0.5):
        super().__init__()
        self.theta_ = np.float32(theta)
        self.tree_ = None

    def computeForces(self, d_particles: ParticleData):
        ctx = get_context()
        if self.tree_ is None or self.tree_.n != d_particles.count:
            self.tree_ = BarnesHutTreeView(d_particles.count)
        L = _lib.lib()
        check(L.nb_bh_build(ctx.handle, C.byref(d_particles.c), _stream()))
        check(L.nb_bh_forces(ctx.handle, C.byref(d_particles.c), float(self.theta_), float(self.G_),
                             float(self.softening_eps_), _stream()))

    def getMethod(self):
        return ForceMethod.BARNES_HUT

    def setTheta(self, theta):
        self.theta_ = np.float32(theta)

    def getTheta(self):
        return float(self.theta_)

    def getTree(self):
        return self.tree_


class SpatialHashGridView:
    def __init__(self, n):
        self.n = n

    def getTotalCells(self):
        t = C.c_int(0)
        check(_lib.lib().nb_sh_total_cells(get_context().handle, C.byref(t)))
        return t.value

    def copyCellDataToHost(self):
        """src/cuda/force_spatial_hash.cu:318-335 (+ dims and padded bbox)."""
        total = self.getTotalCells()
        out = dict(dims=np.zeros(3, np.int32), bbox=np.zeros(6, np.float32),
                   particle_cell=np.zeros(self.n, np.int32), cell_start=np.zeros(total, np.int32),
                   cell_end=np.zeros(total, np.int32), sorted_indices=np.zeros(self.n, np.int32))
        check(_lib.lib().nb_sh_export(get_context().handle, *[out[k].ctypes.data_as(C.c_void_p) for k in
                                                              ("dims", "bbox", "particle_cell", "cell_start",
                                                               "cell_end", "sorted_indices")]))
        out["total_cells"] = total
        return out


class SpatialHashCalculator(ForceCalculator):
    """include/nbody/force_calculator.hpp:176-221; computeForces = src/cuda/force_spatial_hash.cu:370-377."""

    def __init__(self, cell_size: float = 1.0, cutoff_radius: float = 2.0):
        super().__init__()
        self.cell_size_ = np.float32(cell_size)
        self.cutoff_radius_ = np.float32(cutoff_radius)
        self.grid_ = None

    def computeForces(self, d_particles: ParticleData):
        ctx = get_context()
        if self.grid_ is None or self.grid_.n != d_particles.count:
            self.grid_ = SpatialHashGridView(d_particles.count)
        L = _lib.lib()
        check(L.nb_sh_build(ctx.handle, C.byref(d_particles.c), float(self.cell_size_), _stream()))
        check(L.nb_sh_forces(ctx.handle, C.byref(d_particles.c), float(self.cutoff_radius_), float(self.G_),
                             float(self.softening_eps_), _stream()))

    def getMethod(self):
        return ForceMethod.SPATIAL_HASH

    def setCellSize(self, size):
        self.cell_size_ = np.float32(size)

    def setCutoffRadius(self, radius):
        self.cutoff_radius_ = np.float32(radius)

    def getCellSize(self):
        return float(self.cell_size_)

    def getCutoffRadius(self):
        return float(self.cutoff_radius_)

    def getGrid(self):
        return self.grid_


def createForceCalculator(method: ForceMethod, config: SimulationConfig) -> ForceCalculator:
    """src/cuda/force_spatial_hash.cu:380-401 -- note the bare factory leaves Direct at G=1, eps=0.01."""
    if method == ForceMethod.BARNES_HUT:
        calc = BarnesHutCalculator(config.barnes_hut_theta)
        calc.setGravitationalConstant(config.G)
        calc.setSofteningParameter(config.softening)
        return calc
    if method == ForceMethod.SPATIAL_HASH:
        calc = SpatialHashCalculator(config.spatial_hash_cell_size, config.spatial_hash_cutoff)
        calc.setGravitationalConstant(config.G)
        calc.setSofteningParameter(config.softening)
        return calc
    return DirectForceCalculator(config.cuda_block_size)


# ---- integrator -----------------------------------------------------------------------------------
class _Diagnostics(C.Structure):  # nb_diagnostics_t (include/nbody_b200.h)
    _fields_ = [("total_mass", C.c_double), ("center_of_mass", C.c_double * 3), ("momentum", C.c_double * 3),
                ("angular_momentum", C.c_double * 3), ("kinetic", C.c_double), ("potential", C.c_double),
                ("total_energy", C.c_double), ("virial_ratio", C.c_double), ("has_potential", C.c_int)]


class Integrator:
    """include/nbody/integrator.hpp:33-150, src/cuda/integrator.cu:205-293."""

    def __init__(self, block_size: int = 256):
        self.block_size_ = block_size
        self.fuse = True  # use nb_step_direct when the calculator is our DirectForceCalculator

    def integrate(self, d_particles: ParticleData, force_calc: ForceCalculator, dt: float):
        if self.fuse and type(force_calc) is DirectForceCalculator:
            ctx = get_context()
            check(_lib.lib().nb_step_direct(ctx.handle, C.byref(d_particles.c), float(np.float32(dt)),
                                            float(force_calc.G_), float(force_calc.softening_eps2_), _stream()))
            return
        # generic 4-call sequence (integrator.cu:224-238): works with any ForceCalculator
        self.storeOldAccelerations(d_particles)
        self.updatePositions(d_particles, dt)
        force_calc.computeForces(d_particles)
        self.updateVelocities(d_particles, dt)

    def updatePositions(self, d_particles, dt):
        check(_lib.lib().nb_update_positions(get_context().handle, C.byref(d_particles.c), float(np.float32(dt)),
                                             _stream()))

    def updateVelocities(self, d_particles, dt):
        check(_lib.lib().nb_update_velocities(get_context().handle, C.byref(d_particles.c), float(np.float32(dt)),
                                              _stream()))

    def storeOldAccelerations(self, d_particles):
        check(_lib.lib().nb_store_accelerations(get_context().handle, C.byref(d_particles.c), _stream()))

    def computeKineticEnergy(self, d_particles) -> float:
        ke = C.c_double(0)
        check(_lib.lib().nb_kinetic_energy(get_context().handle, C.byref(d_particles.c), C.byref(ke), _stream()))
        return ke.value

    def computePotentialEnergy(self, d_particles, G, eps) -> float:
        pe = C.c_double(0)
        check(_lib.lib().nb_potential_energy(get_context().handle, C.byref(d_particles.c), float(G), float(eps),
                                             C.byref(pe), _stream()))
        return pe.value

    def computeTotalEnergy(self, d_particles, G, eps) -> float:
        return self.computeKineticEnergy(d_particles) + self.computePotentialEnergy(d_particles, G, eps)

    def computeDiagnostics(self, d_particles, G=1.0, eps=0.01, with_potential=True) -> dict:
        """nb_diagnostics: total mass, centre of mass, linear and angular momentum, KE and -- if asked -- PE, total
        energy and the virial ratio 2 KE / |PE|, all fp64 sums over the device arrays (no reference equivalent)."""
        out = _Diagnostics()
        check(_lib.lib().nb_diagnostics(get_context().handle, C.byref(d_particles.c), float(G), float(eps),
                                        1 if with_potential else 0, C.byref(out), _stream()))
        return {
            "total_mass": out.total_mass,
            "center_of_mass": np.array(out.center_of_mass[:]),
            "momentum": np.array(out.momentum[:]),
            "angular_momentum": np.array(out.angular_momentum[:]),
            "kinetic": out.kinetic,
            "potential": out.potential if out.has_potential else None,
            "total_energy": out.total_energy if out.has_potential else None,
            "virial_ratio": out.virial_ratio if out.has_potential else None,
        }

    def setBlockSize(self, size):
        self.block_size_ = size

    def getBlockSize(self):
        return self.block_size_

    def ensureScratchBuffer(self, particle_count):
        pass  # scratch is owned by the context


# ---- CudaGLInterop, headless (include/nbody/cuda_gl_interop.hpp, src/render/cuda_gl_interop.cpp) ------
class CudaGLInterop:
    """The reference's two GL vertex buffers as plain device buffers of interleaved xyz floats: no GL names
    (getPositionVBO() is 0, like the reference's own no-graphics build), "mapping" hands out the device tensor,
    updatePositions / updateVelocities are the reference's copy kernels (:7-28) as one nb_interop_pack launch."""

    def __init__(self):
        self.particle_count_ = 0
        self.pos_ = self.vel_ = None
        self.is_mapped_ = self.is_velocity_mapped_ = False

    def initialize(self, particle_count: int):
        import torch
        self.cleanup()
        self.particle_count_ = int(particle_count)
        if particle_count:
            self.pos_ = torch.empty(3 * particle_count, dtype=torch.float32, device="cuda")
            self.vel_ = torch.empty(3 * particle_count, dtype=torch.float32, device="cuda")

    def cleanup(self):
        self.pos_ = self.vel_ = None
        self.is_mapped_ = self.is_velocity_mapped_ = False
        self.particle_count_ = 0

    def mapPositionBuffer(self):
        self.is_mapped_ = self.pos_ is not None
        return self.pos_

    def mapVelocityBuffer(self):
        self.is_velocity_mapped_ = self.vel_ is not None
        return self.vel_

    def unmapPositionBuffer(self):
        self.is_mapped_ = False

    def unmapVelocityBuffer(self):
        self.is_velocity_mapped_ = False

    def unmapAllBuffers(self):
        self.is_mapped_ = self.is_velocity_mapped_ = False

    def isMapped(self):
        return self.is_mapped_

    def isVelocityMapped(self):
        return self.is_velocity_mapped_

    def getPositionVBO(self):
        return 0

    def getVelocityVBO(self):
        return 0

    def getParticleCount(self):
        return self.particle_count_

    def _pack(self, d_particles, pos, vel):
        if self.pos_ is None or d_particles.count != self.particle_count_:
            raise ValidationException("interop buffers are not sized for these particles")
        check(_lib.lib().nb_interop_pack(get_context().handle, C.byref(d_particles.c),
                                         C.c_void_p(pos.data_ptr()) if pos is not None else None,
                                         C.c_void_p(vel.data_ptr()) if vel is not None else None, _stream()))

    def updatePositions(self, d_particles):
        self._pack(d_particles, self.pos_, None)

    def updateVelocities(self, d_particles):
        self._pack(d_particles, None, self.vel_)

    def verifyDataIntegrity(self, expected, count=None) -> bool:
        """Bit-exact comparison of the position buffer with `count` floats of interleaved expected data."""
        if self.pos_ is None:
            return False
        exp = np.ascontiguousarray(expected, dtype=np.float32).reshape(-1)
        count = exp.size if count is None else int(count)
        if count > 3 * self.particle_count_:
            return False
        return bool(np.array_equal(self.pos_[:count].cpu().numpy().view(np.uint32), exp[:count].view(np.uint32)))


# ---- ParticleSystem facade ------------------------------------------------------------------------
@dataclass
class SimulationState:  # include/nbody/simulation_state.hpp
    particle_count: int
    simulation_time: float
    dt: float
    G: float
    softening: float
    force_method: ForceMethod
    arrays: dict


class ParticleSystem:
    """src/core/particle_system.cpp (the orchestration around the hot path)."""

    def __init__(self):
        self.particle_count_ = 0
        self.dt_ = 0.001
        self.G_ = 1.0
        self.softening_ = 0.01
        self.simulation_time_ = 0.0
        self.force_method_ = ForceMethod.DIRECT_N2
        self.is_paused_ = False
        self.is_initialized_ = False
        self.interop_ = None
        self.use_graph_ = False
        self.graph_, self.graph_key_, self.graph_warm_key_, self.graph_stream_ = None, None, None, None
        self.config_ = SimulationConfig()
        self.d_particles_ = ParticleData()
        self.h_particles_ = ParticleData()
        self.force_calculator_ = None
        self.integrator_ = None

    # particle_system.cpp:40-100
    def initialize(self, config: SimulationConfig):
        from . import ics
        validateSimulationConfig(config)
        validateResourceRequirements(config.particle_count)
        self.config_ = replace(config)
        self.dt_, self.G_, self.softening_ = config.dt, config.G, config.softening
        self.force_method_ = config.force_method
        self.is_paused_ = False
        n = config.particle_count
        if config.init_distribution == InitDistribution.UNIFORM:
            st = ics.uniform_box(n, -10.0, 10.0)
        elif config.init_distribution == InitDistribution.SPHERICAL:
            st = ics.uniform_sphere(n, 10.0)
        else:
            st = ics.disk(n, 10.0, 1.0, 0.5)
        self._install(st)
        self.simulation_time_ = 0.0
        self._resize_interop()  # particle_system.cpp:96-99

    def initializeWithDistribution(self, particle_count: int, dist: InitDistribution):  # :102-107
        cfg = SimulationConfig()
        cfg.particle_count = particle_count
        cfg.init_distribution = dist
        self.initialize(cfg)

    # rendering interop (particle_system.cpp:320-345), headless: device buffers instead of GL buffer objects
    def getInterop(self):
        return self.interop_

    def initializeInterop(self):
        if not self.is_initialized_:
            raise ValidationException("ParticleSystem is not initialized")
        self.interop_ = CudaGLInterop()
        self.interop_.initialize(self.particle_count_)
        self.updateInteropBuffer()

    def updateInteropBuffer(self):
        if self.interop_ is None:
            return
        self.interop_.updatePositions(self.d_particles_)
        self.interop_.updateVelocities(self.d_particles_)

    def _resize_interop(self):
        if self.interop_ is not None:
            self.interop_.initialize(self.particle_count_)
            self.updateInteropBuffer()

    def getForceMethod(self):
        return self.force_method_

    def getTimeStep(self):
        return self.dt_

    def _install(self, st: dict):
        n = st["pos_x"].size
        self.particle_count_ = n
        self.h_particles_ = ParticleDataManager.fromState(st)
        self.d_particles_ = ParticleData()
        ParticleDataManager.allocateDevice(self.d_particles_, n)
        ParticleDataManager.copyToDevice(self.d_particles_, self.h_particles_)
        self.createForceCalculator()
        self.integrator_ = Integrator(self.config_.cuda_block_size)
        self.force_calculator_.computeForces(self.d_particles_)  # initial a(0), particle_system.cpp:91
        self.is_initialized_ = True

    def createForceCalculator(self):  # particle_system.cpp:109-113
        self.force_calculator_ = createForceCalculator(self.force_method_, self.config_)
        self.force_calculator_.setGravitationalConstant(self.G_)
        self.force_calculator_.setSofteningParameter(self.softening_)

    def update(self, dt: float):  # particle_system.cpp:115-127
        if not self.is_initialized_ or self.is_paused_:
            return
        if self.use_graph_ and self.force_method_ in (ForceMethod.DIRECT_N2, ForceMethod.BARNES_HUT):
            self._update_graphed(dt)
        else:
            self.integrator_.integrate(self.d_particles_, self.force_calculator_, dt)
        self.simulation_time_ += dt
        if self.interop_ is not None:
            self.updateInteropBuffer()

    # ---- CUDA-graph replay of the steady-state step --------------------
    def setUseCudaGraph(self, on: bool):
        """Replay update() as one captured graph launch: the first step with a given set of
        parameters runs eagerly (sizes all scratch), the second is captured, later ones replay."""
        self.use_graph_ = bool(on)
        if not on:
            self._drop_graph()

    def _drop_graph(self):
        if getattr(self, "graph_", None):
            check(_lib.lib().nb_graph_destroy(self.graph_))
        self.graph_, self.graph_key_, self.graph_warm_key_ = None, None, None

    def _graph_key(self, dt):
        fc, d = self.force_calculator_, self.d_particles_.c
        extra = fc.getTheta() if self.force_method_ == ForceMethod.BARNES_HUT else 0.0
        return (float(np.float32(dt)), int(self.force_method_), fc.getGravitationalConstant(),
                fc.getSofteningParameter(), extra, int(d.count), int(d.pos_x or 0), int(d.acc_x or 0), id(fc))

    def _update_graphed(self, dt):
        import torch
        L, ctx = _lib.lib(), get_context()
        key = self._graph_key(dt)
        # a stale graph (context scratch reallocated since capture, e.g. by a renderer frame with more fragments or
        # a larger second system) must not be replayed: drop it, run this step eagerly, capture again
        if self.graph_ and L.nb_graph_is_stale(self.graph_):
            self._drop_graph()
        if self.graph_ and self.graph_key_ == key:
            check(L.nb_graph_launch(self.graph_, _stream()))
            return
        if self.graph_warm_key_ != key:      # first step with these parameters: eager
            self._drop_graph()
            self.integrator_.integrate(self.d_particles_, self.force_calculator_, dt)
            self.graph_warm_key_ = key
            return
        if self.graph_stream_ is None:
            self.graph_stream_ = torch.cuda.Stream()
        with torch.cuda.stream(self.graph_stream_):
            check(L.nb_graph_begin_capture(ctx.handle, _stream()))
            try:
                self.integrator_.integrate(self.d_particles_, self.force_calculator_, dt)
            finally:
                g = C.c_void_p()
                rc = L.nb_graph_end_capture(ctx.handle, _stream(), C.byref(g))
            check(rc)
        self.graph_, self.graph_key_ = g, key
        check(L.nb_graph_launch(self.graph_, _stream()))

    def graphNodeCount(self) -> int:
        n = C.c_size_t(0)
        if self.graph_:
            check(_lib.lib().nb_graph_node_count(self.graph_, C.byref(n)))
        return n.value

    def reset(self):
        if self.is_initialized_:
            self.initialize(self.config_)

    def pause(self):
        self.is_paused_ = True

    def resume(self):
        self.is_paused_ = False

    def isPaused(self):
        return self.is_paused_

    def setForceMethod(self, method: ForceMethod):  # particle_system.cpp:137-143
        if self.force_method_ != method:
            self.force_method_ = method
            self.config_.force_method = method
            self.createForceCalculator()

    def setGravitationalConstant(self, G):
        if G <= 0 or not _finite(G):
            raise ValidationException("Gravitational constant must be positive and finite")
        self.G_ = G
        self.config_.G = G
        if self.force_calculator_:
            self.force_calculator_.setGravitationalConstant(G)

    def setSofteningParameter(self, eps):
        validateSoftening(eps)
        self.softening_ = eps
        self.config_.softening = eps
        if self.force_calculator_:
            self.force_calculator_.setSofteningParameter(eps)

    def setTimeStep(self, dt):
        validateTimeStep(dt)
        self.dt_ = dt
        self.config_.dt = dt

    def setBarnesHutTheta(self, theta):
        validateTheta(theta)
        self.config_.barnes_hut_theta = theta
        if isinstance(self.force_calculator_, BarnesHutCalculator):
            self.force_calculator_.setTheta(theta)

    def setSpatialHashCellSize(self, size):
        if size <= 0 or not _finite(size):
            raise ValidationException("Spatial hash cell size must be positive and finite")
        self.config_.spatial_hash_cell_size = size
        if isinstance(self.force_calculator_, SpatialHashCalculator):
            self.force_calculator_.setCellSize(size)

    def setSpatialHashCutoff(self, cutoff):
        if cutoff <= 0 or not _finite(cutoff):
            raise ValidationException("Spatial hash cutoff must be positive and finite")
        self.config_.spatial_hash_cutoff = cutoff
        if isinstance(self.force_calculator_, SpatialHashCalculator):
            self.force_calculator_.setCutoffRadius(cutoff)

    def getGravitationalConstant(self):
        return self.G_

    def getSofteningParameter(self):
        return self.softening_

    def getParticleCount(self):
        return self.particle_count_

    def getSimulationTime(self):
        return self.simulation_time_

    def getDeviceData(self):
        return self.d_particles_

    def getState(self) -> SimulationState:  # particle_system.cpp:213-238
        arr = self.d_particles_.numpy()
        keep = {k: arr[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")}
        return SimulationState(self.particle_count_, self.simulation_time_, self.dt_, self.G_, self.softening_,
                               self.force_method_, keep)

    def setState(self, state: SimulationState):  # particle_system.cpp:240-292
        cfg = replace(self.config_, particle_count=state.particle_count, dt=state.dt, G=state.G,
                      softening=state.softening, force_method=state.force_method)
        validateSimulationConfig(cfg)
        validateResourceRequirements(state.particle_count)
        self.config_ = cfg
        self.simulation_time_ = state.simulation_time
        self.dt_, self.G_, self.softening_ = state.dt, state.G, state.softening
        self.force_method_ = state.force_method
        st = {k: np.ascontiguousarray(v, np.float32) for k, v in state.arrays.items()}
        self._install(st)  # accelerations zeroed then recomputed (particle_system.cpp:261, 283)
        self.is_paused_ = False
        self._resize_interop()  # :288-291

    def saveState(self, filename):  # particle_system.cpp:294-297
        from .serialization import Serializer
        Serializer.save(filename, self.getState())

    def loadState(self, filename):  # particle_system.cpp:299-302
        from .serialization import Serializer
        self.setState(Serializer.load(filename))

    def computeKineticEnergy(self):
        return self.integrator_.computeKineticEnergy(self.d_particles_) if self.integrator_ else 0.0

    def computePotentialEnergy(self):
        if not self.integrator_:
            return 0.0
        return self.integrator_.computePotentialEnergy(self.d_particles_, self.G_, self.softening_)

    def computeTotalEnergy(self):
        return self.computeKineticEnergy() + self.computePotentialEnergy()

    def computeDiagnostics(self, with_potential=True):
        """Conserved quantities of the current state (see Integrator.computeDiagnostics)."""
        if not self.integrator_:
            raise ValidationException("ParticleSystem is not initialized")
        return self.integrator_.computeDiagnostics(self.d_particles_, self.G_, self.softening_, with_potential)
