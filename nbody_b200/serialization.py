"""The reference's `.nbody` checkpoint format (include/nbody/serialization.hpp:33-66,
src/utils/serialization.cpp:25-137) for the Python host mirror: a 56-byte little-endian header
{u32 magic "NBOD", u32 version 1, u64 particle_count, f32 simulation_time, dt, G, softening,
u32 force_method, 4 reserved u32} followed by seven fp32 arrays of particle_count values
(pos x/y/z, vel x/y/z, mass).  Files are byte-identical to the reference's for the same state
(tests/golden/ref_checkpoint.nbody) and load in either implementation."""
from __future__ import annotations

import io
import struct

import numpy as np

from ._lib import ValidationException

NBODY_MAGIC = 0x4E424F44
NBODY_VERSION = 1
MAX_PARTICLE_COUNT = 100_000_000
_HEADER = struct.Struct("<IIQffffI4I4x")  # 4 tail bytes: the C struct is 8-byte aligned
assert _HEADER.size == 56
ARRAYS = ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")


class Serializer:
    """Static save/load/validate, same names and error behaviour as nbody::Serializer."""

    @staticmethod
    def save(target, state) -> None:
        if isinstance(target, (str, bytes)):
            try:
                f = open(target, "wb")
            except OSError:
                raise RuntimeError(f"Failed to open file for writing: {target}")
            with f:
                Serializer.save(f, state)
            return
        n = int(state.particle_count)
        target.write(_HEADER.pack(NBODY_MAGIC, NBODY_VERSION, n, float(state.simulation_time), float(state.dt),
                                  float(state.G), float(state.softening), int(state.force_method), 0, 0, 0, 0))
        for k in ARRAYS:
            a = np.ascontiguousarray(state.arrays[k], dtype="<f4")
            target.write(a.tobytes())

    @staticmethod
    def _read_header(src):
        raw = src.read(_HEADER.size)
        if len(raw) != _HEADER.size:
            raise RuntimeError("Failed to read file header: file may be truncated or corrupted")
        magic, version, n, t, dt, G, eps, method = _HEADER.unpack(raw)[:8]
        if magic != NBODY_MAGIC:
            raise RuntimeError("Invalid file format: wrong magic number")
        if version != NBODY_VERSION:
            raise RuntimeError("Unsupported file version")
        return n, t, dt, G, eps, method

    @staticmethod
    def load(source):
        from .api import ForceMethod, SimulationState
        if isinstance(source, (str, bytes)) and not isinstance(source, io.IOBase):
            try:
                f = open(source, "rb")
            except OSError:
                raise RuntimeError(f"Failed to open file for reading: {source}")
            with f:
                return Serializer.load(f)
        n, t, dt, G, eps, method = Serializer._read_header(source)
        if n > MAX_PARTICLE_COUNT:
            raise ValidationException(f"Particle count ({n}) exceeds maximum allowed ({MAX_PARTICLE_COUNT})")
        arrays = {}
        for k in ARRAYS:
            raw = source.read(4 * n)
            if len(raw) != 4 * n:
                raise RuntimeError("Failed to read particle data: file may be truncated or corrupted")
            arrays[k] = np.frombuffer(raw, dtype="<f4").astype(np.float32, copy=True)
        return SimulationState(n, t, dt, G, eps, ForceMethod(method), arrays)

    @staticmethod
    def validateFile(filename) -> bool:
        try:
            with open(filename, "rb") as f:
                return Serializer.validateStream(f)
        except OSError:
            return False

    @staticmethod
    def validateStream(src) -> bool:
        try:
            Serializer._read_header(src)
            return True
        except Exception:
            return False
