"""Headless mirror of the reference's render classes (include/nbody/camera.hpp, renderer.hpp):
Camera, ColorMapper and a Renderer whose render() runs the CUDA splat kernels of csrc/render.cu into
a device RGBA8 framebuffer instead of issuing glDrawArrays.  Windows, swapchains and CUDA-GL interop
are left out (no graphics stack on this box); images go to a PPM file where one is wanted."""
from __future__ import annotations

import ctypes as C
import enum
import math

import numpy as np

from . import _lib

f32 = np.float32


class ColorMode(enum.IntEnum):  # include/nbody/types.hpp
    DEPTH = 0
    VELOCITY = 1
    DENSITY = 2


class RenderParams(C.Structure):
    """nb_render_params (include/nbody_b200.h)"""
    _fields_ = [("view", C.c_float * 16), ("proj", C.c_float * 16), ("point_size", C.c_float),
                ("max_depth", C.c_float), ("max_velocity", C.c_float), ("color_mode", C.c_int),
                ("width", C.c_int), ("height", C.c_int), ("point_size_min", C.c_float),
                ("point_size_max", C.c_float), ("clear_color", C.c_float * 4)]


def _normalize(v):
    return (v / f32(math.sqrt(float(np.dot(v, v))))).astype(f32)


def look_at(eye, center, up) -> np.ndarray:
    """glm::lookAt (right-handed), column-major flat[col*4+row], fp32."""
    eye, center, up = (np.asarray(a, f32) for a in (eye, center, up))
    f = _normalize(center - eye)
    s = _normalize(np.cross(f, up).astype(f32))
    u = np.cross(s, f).astype(f32)
    m = np.zeros((4, 4), f32)  # m[col][row]
    m[0][0], m[1][0], m[2][0] = s
    m[0][1], m[1][1], m[2][1] = u
    m[0][2], m[1][2], m[2][2] = -f
    m[3][0], m[3][1], m[3][2] = -np.dot(s, eye), -np.dot(u, eye), np.dot(f, eye)
    m[3][3] = 1.0
    return m.reshape(16)


def perspective(fovy_rad, aspect, near, far) -> np.ndarray:
    """glm::perspective (right-handed, depth -1..1), column-major flat, fp32."""
    t = f32(math.tan(float(f32(fovy_rad)) / 2.0))
    m = np.zeros((4, 4), f32)
    m[0][0] = f32(1.0) / (f32(aspect) * t)
    m[1][1] = f32(1.0) / t
    m[2][2] = -(f32(far) + f32(near)) / (f32(far) - f32(near))
    m[2][3] = -1.0
    m[3][2] = -(f32(2.0) * f32(far) * f32(near)) / (f32(far) - f32(near))
    return m.reshape(16)


class Camera:
    """include/nbody/camera.hpp:9-88, src/render/camera.cpp:7-102"""

    def __init__(self, fov=45.0, aspect=16.0 / 9.0, near=0.1, far=1000.0):
        self.position_ = np.array([0, 0, 50], f32)
        self.target_ = np.zeros(3, f32)
        self.up_ = np.array([0, 1, 0], f32)
        self.fov_, self.aspect_, self.near_, self.far_ = fov, aspect, near, far

    def getViewMatrix(self):
        return look_at(self.position_, self.target_, self.up_)

    def getProjectionMatrix(self):
        return perspective(math.radians(self.fov_), self.aspect_, self.near_, self.far_)

    def getViewProjectionMatrix(self):  # camera.cpp:32-34; flat column-major like the two factors
        v = self.getViewMatrix().reshape(4, 4).T
        p = self.getProjectionMatrix().reshape(4, 4).T
        return (p @ v).T.reshape(16).astype(f32)

    def setPosition(self, p):
        self.position_ = np.asarray(p, f32)

    def setTarget(self, t):
        self.target_ = np.asarray(t, f32)

    def setUp(self, u):
        self.up_ = np.asarray(u, f32)

    def getPosition(self):
        return self.position_

    def getTarget(self):
        return self.target_

    def getUp(self):
        return self.up_

    def getFOV(self):
        return self.fov_

    def getAspectRatio(self):
        return self.aspect_

    def getNear(self):
        return self.near_

    def getFar(self):
        return self.far_

    def getForward(self):
        return _normalize(self.target_ - self.position_)

    def getRight(self):
        return _normalize(np.cross(self.getForward(), self.up_).astype(f32))

    def setFOV(self, fov):
        self.fov_ = fov

    def setAspectRatio(self, a):
        self.aspect_ = a

    def setNearFar(self, near, far):
        self.near_, self.far_ = near, far

    def rotate(self, yaw, pitch):  # camera.cpp:46-65
        d = self.position_ - self.target_
        dist = math.sqrt(float(np.dot(d, d)))
        theta = math.atan2(float(d[0]), float(d[2])) + yaw
        phi = min(max(math.acos(float(d[1]) / dist) + pitch, 0.1), 3.04)
        self.position_ = self.target_ + np.array([dist * math.sin(phi) * math.sin(theta), dist * math.cos(phi),
                                                  dist * math.sin(phi) * math.cos(theta)], f32)

    orbit = rotate

    def pan(self, dx, dy):  # :67-75
        right = self.getRight()
        up = _normalize(np.cross(right, self.getForward()).astype(f32))
        off = right * f32(dx) + up * f32(dy)
        self.position_ = self.position_ + off
        self.target_ = self.target_ + off

    def zoom(self, delta):  # :77-85
        d = self.target_ - self.position_
        dist = math.sqrt(float(np.dot(d, d)))
        new = min(max(dist - delta, 1.0), 1000.0)
        self.position_ = self.target_ - _normalize(d) * f32(new)

    def reset(self):  # :91-96
        self.position_ = np.array([0, 0, 50], f32)
        self.target_ = np.zeros(3, f32)
        self.up_ = np.array([0, 1, 0], f32)

    def setOrbitDistance(self, distance):
        self.position_ = self.target_ + _normalize(self.position_ - self.target_) * f32(distance)

    def getOrbitDistance(self):
        d = self.position_ - self.target_
        return math.sqrt(float(np.dot(d, d)))


class ColorMapper:
    """src/render/camera.cpp:104-128"""

    @staticmethod
    def gradientMap(t, low, high):
        low, high = np.asarray(low, f32), np.asarray(high, f32)
        return low * (f32(1.0) - f32(t)) + high * f32(t)

    @staticmethod
    def velocityToColor(velocity, max_velocity):
        return ColorMapper.gradientMap(min(max(velocity / max_velocity, 0.0), 1.0), (0, 0, 1), (1, 0, 0))

    @staticmethod
    def depthToColor(depth, max_depth):
        return ColorMapper.gradientMap(min(max(depth / max_depth, 0.0), 1.0), (1, 0.5, 0), (0, 0.5, 1))

    @staticmethod
    def densityToColor(density, max_density):
        return ColorMapper.gradientMap(min(max(density / max_density, 0.0), 1.0), (0.2, 0.2, 0.8), (1, 1, 0.2))


class Renderer:
    """include/nbody/renderer.hpp: initialize / render / setters / getCamera / onResize; render()
    takes the device ParticleData (the reference takes GL buffer ids filled from it)."""

    def __init__(self):
        self.color_mode_ = ColorMode.DEPTH          # src/render/renderer.cpp:79-88
        self.point_size_ = 2.0
        self.max_depth_ = 100.0
        self.max_velocity_ = 10.0
        self.width_, self.height_ = 1280, 720
        self.point_size_range_ = (1.0, 63.0)
        self.clear_color_ = (0.02, 0.02, 0.05, 1.0)
        self.camera_ = Camera()
        self.framebuffer_ = None
        self.is_initialized_ = False

    def initialize(self, width, height):           # :93-121 without GLEW / VAO / shader compilation
        import torch
        self.width_, self.height_ = int(width), int(height)
        self.camera_.setPosition((0, 0, 50))
        self.camera_.setTarget((0, 0, 0))
        self.camera_.setAspectRatio(float(width) / float(height))
        self.framebuffer_ = torch.zeros(self.height_ * self.width_ * 4, dtype=torch.uint8, device="cuda")
        self.is_initialized_ = True

    def cleanup(self):                             # :123-140: drops the framebuffer (no GL objects to delete)
        self.framebuffer_ = None
        self.is_initialized_ = False

    def onResize(self, width, height):             # :233-238
        self.initialize(width, height) if self.framebuffer_ is None or (width, height) != (self.width_, self.height_) \
            else None
        self.camera_.setAspectRatio(float(width) / float(height))

    def setColorMode(self, mode):
        self.color_mode_ = ColorMode(mode)

    def getColorMode(self):
        return self.color_mode_

    def setPointSize(self, size):
        self.point_size_ = float(size)

    def getPointSize(self):
        return self.point_size_

    def setCamera(self, camera):
        self.camera_ = camera

    def getShaderProgram(self):
        return 0  # no GL program object: the shader math is the CUDA kernels

    def setMaxDepth(self, d):
        self.max_depth_ = float(d)

    def setMaxVelocity(self, v):
        self.max_velocity_ = float(v)

    def getCamera(self):
        return self.camera_

    def isInitialized(self):
        return self.is_initialized_

    def params(self) -> RenderParams:
        p = RenderParams()
        p.view[:] = [float(x) for x in self.camera_.getViewMatrix()]
        p.proj[:] = [float(x) for x in self.camera_.getProjectionMatrix()]
        p.point_size, p.max_depth, p.max_velocity = self.point_size_, self.max_depth_, self.max_velocity_
        p.color_mode, p.width, p.height = int(self.color_mode_), self.width_, self.height_
        p.point_size_min, p.point_size_max = self.point_size_range_
        p.clear_color[:] = self.clear_color_
        return p

    def render(self, d_particles, particle_count=None):
        """One frame into the device framebuffer (asynchronous on the current stream)."""
        if not self.is_initialized_:
            return                                  # :185-186
        import torch
        from .api import get_context
        ctx = get_context()
        p = self.params()
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(_lib.lib().nb_render_splats(ctx.handle, C.byref(d_particles.c), C.byref(p),
                                               C.c_void_p(self.framebuffer_.data_ptr()), stream))

    def fragments(self) -> int:
        from .api import get_context
        n = C.c_uint64(0)
        _lib.check(_lib.lib().nb_render_stats(get_context().handle, C.byref(n)))
        return n.value

    def readPixels(self) -> np.ndarray:
        """height x width x 4 uint8, row 0 = bottom row (glReadPixels order)."""
        return self.framebuffer_.cpu().numpy().reshape(self.height_, self.width_, 4)

    def savePPM(self, filename):
        img = self.readPixels()[::-1, :, :3]       # top row first
        with open(filename, "wb") as f:
            f.write(b"P6\n%d %d\n255\n" % (self.width_, self.height_))
            f.write(np.ascontiguousarray(img).tobytes())
