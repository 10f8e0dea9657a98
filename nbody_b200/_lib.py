"""ctypes loader for libnbody_b200.so (the C ABI declared in include/nbody_b200.h).

There is no fallback: if the shared library is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("NB_LIB_PATH") or os.path.join(_HERE, "libnbody_b200.so")  # NB_LIB_PATH: A/B runs of an older build


class CudaException(RuntimeError):
    """include/nbody/error_handling.hpp:29-52"""


class ResourceException(RuntimeError):
    """include/nbody/error_handling.hpp:77-95"""


class ValidationException(RuntimeError):
    """include/nbody/error_handling.hpp:98-102 -- message prefixed 'Validation Error: '"""

    def __init__(self, msg: str):
        super().__init__("Validation Error: " + msg)


class NoDeviceError(RuntimeError):
    pass


NB_ERR_INVALID, NB_ERR_NO_DEVICE, NB_ERR_RESOURCE, NB_ERR_GRID_TOO_LARGE, NB_ERR_NCCL, NB_ERR_INTERNAL = -1, -2, -3, -4, -5, -6
NB_ERR_STALE_GRAPH = -7


class NbParticles(C.Structure):
    """nb_particles == nbody::ParticleData (include/nbody/types.hpp:234-276), 112 bytes."""

    _fields_ = [(n, C.c_void_p) for n in (
        "pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z",
        "acc_old_x", "acc_old_y", "acc_old_z", "mass")] + [("count", C.c_size_t)]


assert C.sizeof(NbParticles) == 112

_lib = None

# name -> (restype, argtypes); every symbol include/nbody_b200.h declares
_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_pp = C.POINTER(NbParticles)
SYMBOLS = {
    "nb_version": (_i, []),
    "nb_last_error": (C.c_char_p, []),
    "nb_device_info": (_i, [_i, C.POINTER(_i), C.POINTER(_i), C.POINTER(_i), C.POINTER(_sz), C.POINTER(_i)]),
    "nb_context_create": (_i, [_i, _sz, C.POINTER(_vp)]),
    "nb_context_destroy": (_i, [_vp]),
    "nb_context_sm_count": (_i, [_vp]),
    "nb_synchronize": (_i, [_vp, _vp]),
    "nb_launch_count": (C.c_uint64, [_vp]),
    "nb_device_alloc": (_i, [_vp, _sz, C.POINTER(_vp)]),
    "nb_device_free": (_i, [_vp, _vp]),
    "nb_device_download": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "nb_device_upload": (_i, [_vp, _vp, _vp, _sz, _vp]),
    "nb_stream_create": (_i, [_vp, C.POINTER(_vp)]),
    "nb_stream_destroy": (_i, [_vp, _vp]),
    "nb_graph_begin_capture": (_i, [_vp, _vp]),
    "nb_graph_end_capture": (_i, [_vp, _vp, C.POINTER(_vp)]),
    "nb_graph_launch": (_i, [_vp, _vp]),
    "nb_graph_is_stale": (_i, [_vp]),
    "nb_scratch_generation": (C.c_uint64, [_vp]),
    "nb_scratch_bytes": (_sz, [_vp]),
    "nb_probe_fp32": (_i, [_vp, _i, C.POINTER(C.c_double), C.POINTER(C.c_double), _vp]),
    "nb_profile_stages_enable": (_i, [_vp, _i]),
    "nb_profile_stages_read": (_i, [_vp, C.c_char_p, _sz]),
    "nb_graph_node_count": (_i, [_vp, C.POINTER(_sz)]),
    "nb_graph_destroy": (_i, [_vp]),
    "nb_profile_enable": (_i, [_vp, _i]),
    "nb_profile_read": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    "nb_particles_alloc_device": (_i, [_pp, _sz]),
    "nb_particles_free_device": (_i, [_pp]),
    "nb_particles_upload": (_i, [_vp, _pp, _pp, _vp]),
    "nb_particles_download": (_i, [_vp, _pp, _pp, _vp]),
    "nb_direct_forces": (_i, [_vp, _pp, _f, _f, _vp]),
    "nb_direct_forces_static": (_i, [_vp, _pp, _f, _f, _vp]),
    "nb_store_accelerations": (_i, [_vp, _pp, _vp]),
    "nb_update_positions": (_i, [_vp, _pp, _f, _vp]),
    "nb_update_velocities": (_i, [_vp, _pp, _f, _vp]),
    "nb_step_direct": (_i, [_vp, _pp, _f, _f, _f, _vp]),
    "nb_kinetic_energy": (_i, [_vp, _pp, C.POINTER(C.c_double), _vp]),
    "nb_potential_energy": (_i, [_vp, _pp, _f, _f, C.POINTER(C.c_double), _vp]),
    "nb_diagnostics": (_i, [_vp, _pp, _f, _f, _i, _vp, _vp]),
    "nb_host_direct_forces": (_i, [_vp, _pp, _f, _f]),
    "nb_host_step_direct": (_i, [_vp, _pp, _f, _f, _f, _i, _i]),
    "nb_bh_build": (_i, [_vp, _pp, _vp]),
    "nb_bh_forces": (_i, [_vp, _pp, _f, _f, _f, _vp]),
    "nb_bh_export": (_i, [_vp] + [_vp] * 8),
    "nb_bh_stats": (_i, [_vp, C.POINTER(C.c_uint64), C.POINTER(_i)]),
    "nb_bh_max_stack": (_i, [_vp, C.POINTER(_i), C.POINTER(_i)]),
    "nb_sh_build": (_i, [_vp, _pp, _f, _vp]),
    "nb_sh_forces": (_i, [_vp, _pp, _f, _f, _f, _vp]),
    "nb_sh_export": (_i, [_vp] + [_vp] * 6),
    "nb_sh_total_cells": (_i, [_vp, C.POINTER(_i)]),
    "nb_sh_stats": (_i, [_vp, C.POINTER(C.c_uint64)]),
    "nb_render_splats": (_i, [_vp, _pp, _vp, _vp, _vp]),
    "nb_render_stats": (_i, [_vp, C.POINTER(C.c_uint64)]),
    "nb_interop_pack": (_i, [_vp, _pp, _vp, _vp, _vp]),
    "nb_sort_pairs": (_i, [_vp, _vp, _vp, _sz, _i, _vp]),
    "nb_sort_passes": (_i, [_i]),
    "nb_bounding_box": (_i, [_vp, _pp, _vp, _vp]),
    "nb_mgpu_layout": (_i, [_vp, _sz, _i, _i, C.POINTER(_sz), C.POINTER(_sz), C.POINTER(_sz), C.POINTER(_sz)]),
    "nb_mgpu_pack_local": (_i, [_vp, _pp, _vp, _f, _f, _i, _vp]),
    "nb_mgpu_forces_local": (_i, [_vp, _vp, _f, _vp]),
    "nb_mgpu_forces_remote": (_i, [_vp, _vp, _f, _vp]),
    "nb_mgpu_finish": (_i, [_vp, _pp, _f, _i, _vp]),
    "nb_mgpu_peer_alloc": (_i, [_vp, _vp]),
    "nb_mgpu_peer_open": (_i, [_vp, _i, _vp]),
    "nb_mgpu_peer_step": (_i, [_vp, _pp, _f, _f, _f, _i, _vp]),
    "nb_mgpu_potential_partial": (_i, [_vp, _pp, _vp, _f, C.POINTER(C.c_double), _vp]),
    "nb_mgpu_peer_push": (_i, [_vp, _pp, _f, _f, _i, _vp]),
    "nb_mgpu_peer_forces": (_i, [_vp, _pp, _f, _f, _i, _vp]),
    "nb_mgpu_peer_block": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "nb_mgpu_peer_map": (_i, [_vp, _i, _vp]),
    "nb_mgpu_peer_close": (_i, [_vp]),
    "nb_mgpu_bh_alloc": (_i, [_vp, _sz, _i, _i, _vp]),
    "nb_mgpu_bh_open": (_i, [_vp, _i, _vp]),
    "nb_mgpu_bh_block": (_i, [_vp, C.POINTER(_vp), C.POINTER(_sz)]),
    "nb_mgpu_bh_map": (_i, [_vp, _i, _vp]),
    "nb_mgpu_bh_step": (_i, [_vp, _pp, _f, _f, _f, _f, _i, _vp]),
    "nb_mgpu_bh_traverse": (_i, [_vp, _pp, _f, _f, _f, _f, _i, _vp]),
    "nb_mgpu_bh_finish": (_i, [_vp, _pp, _f, _i, _vp]),
    "nb_mgpu_bh_close": (_i, [_vp]),
}


def lib():
    """Load the CUDA extension; fail loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C nbody_b200/csrc`). There is no CPU or PyTorch fallback.")
    L = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        try:
            fn = getattr(L, name)  # AttributeError if the library does not export it
        except AttributeError:
            if os.environ.get("NB_LIB_PATH"):
                continue  # A/B run of an older build: it may predate newer entry points
            raise
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def last_error() -> str:
    return (lib().nb_last_error() or b"").decode("utf-8", "replace")


def check(rc: int):
    """Translate a C-ABI return code into the reference's exception types."""
    if rc == 0:
        return
    msg = last_error()
    if rc == NB_ERR_INVALID:
        raise ValidationException(msg)
    if rc == NB_ERR_NO_DEVICE:
        raise NoDeviceError(msg)
    if rc == NB_ERR_RESOURCE:
        raise ResourceException(msg)
    if rc == NB_ERR_GRID_TOO_LARGE:
        raise RuntimeError(msg)  # std::runtime_error in the reference (force_spatial_hash.cu:252-254)
    if rc > 0:
        raise CudaException(msg)
    raise RuntimeError(f"nbody_b200 error {rc}: {msg}")
