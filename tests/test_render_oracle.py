"""CPU: the splat-renderer oracle (oracle/render.c) against values that follow directly from the
reference's shader source (src/render/renderer.cpp:16-77) and camera defaults (src/render/camera.cpp),
and the host-side Camera / ColorMapper mirror.  Parity of the oracle against a GL implementation:
unpinned (no graphics stack on this box)."""
import math

import numpy as np


def _one(pos, vel=(0, 0, 0)):
    z = np.zeros(1, np.float32)
    return dict(pos_x=z + np.float32(pos[0]), pos_y=z + np.float32(pos[1]), pos_z=z + np.float32(pos[2]),
                vel_x=z + np.float32(vel[0]), vel_y=z + np.float32(vel[1]), vel_z=z + np.float32(vel[2]))


def _cam(w, h):
    from nbody_b200.render import Camera
    c = Camera()
    c.setAspectRatio(w / h)
    return c


def _at(c, w, h, xw, yw, depth):
    """World position (camera on +z looking at the origin) that projects to window (xw, yw) at `depth`."""
    p = c.getProjectionMatrix().reshape(4, 4).T
    return ((xw - w / 2) * depth / (w / 2 * p[0, 0]), (yw - h / 2) * depth / (h / 2 * p[1, 1]), 50.0 - depth)


def test_camera_matrices_match_glm_closed_forms():
    c = _cam(1280, 720)
    v = c.getViewMatrix().reshape(4, 4).T       # row-major view for reading
    assert np.allclose(v, [[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, -50], [0, 0, 0, 1]])
    p = c.getProjectionMatrix().reshape(4, 4).T
    t = math.tan(math.radians(45.0) / 2)
    assert abs(p[1, 1] - 1 / t) < 1e-6 and abs(p[0, 0] - 1 / (t * 1280 / 720)) < 1e-6
    assert abs(p[2, 2] + (1000 + 0.1) / (1000 - 0.1)) < 1e-6 and p[3, 2] == -1
    assert abs(p[2, 3] + 2 * 1000 * 0.1 / (1000 - 0.1)) < 1e-6
    c.zoom(10.0)
    assert abs(c.getOrbitDistance() - 40.0) < 1e-5
    c.reset()
    c.rotate(math.pi / 2, 0.0)                   # quarter orbit about the target keeps the distance
    assert abs(c.getOrbitDistance() - 50.0) < 1e-4 and abs(c.getPosition()[0] - 50.0) < 1e-3


def test_color_ramps():
    from nbody_b200.render import ColorMapper
    assert np.allclose(ColorMapper.depthToColor(0.0, 100.0), (1, 0.5, 0))
    assert np.allclose(ColorMapper.depthToColor(100.0, 100.0), (0, 0.5, 1))
    assert np.allclose(ColorMapper.velocityToColor(5.0, 10.0), (0.5, 0, 0.5))
    assert np.allclose(ColorMapper.densityToColor(1e9, 1.0), (1, 1, 0.2))   # clamped


def test_single_body_lands_on_the_centre_pixel_with_the_depth_colour(orc):
    w, h = 64, 36
    c = _cam(w, h)
    st = _one(_at(c, w, h, w // 2 + 0.5, h // 2 + 0.5, 50.0))   # centre of pixel (w/2, h/2)
    img, frags = orc.render_splats(st, c.getViewMatrix(), c.getProjectionMatrix(), 2.0, 100.0, 10.0, 0, w, h)
    assert frags == 1                            # 2 / 50 px clamps to the 1-px minimum
    lit = np.argwhere((img != img[0, 0]).any(axis=2))
    assert lit.tolist() == [[h // 2, w // 2]]
    assert img[0, 0].tolist() == [5, 5, 13, 255]  # clear colour 0.02, 0.02, 0.05, 1
    # depth 50 of 100 -> t = 0.5 -> mix((1,.5,0),(0,.5,1)) = (.5,.5,.5); centre distance < 0.4 px -> alpha 1
    assert img[h // 2, w // 2].tolist() == [128, 128, 128, 255]


def test_culling_and_blend_order(orc):
    w, h = 64, 36
    c = _cam(w, h)
    behind = _one((0, 0, 60.0))                   # behind the camera at z = 50
    img, frags = orc.render_splats(behind, c.getViewMatrix(), c.getProjectionMatrix(), 2.0, 100.0, 10.0, 0, w, h)
    assert frags == 0 and (img == img[0, 0]).all()
    # two bodies on one pixel, second one semi-transparent there: the result depends on draw order
    a = _at(c, w, h, w // 2 + 0.5, h // 2 + 0.5, 50.0)          # pixel centre: alpha 1
    b = _at(c, w, h, w // 2 + 0.95, h // 2 + 0.5, 40.0)         # 0.45 px off the same centre: alpha 0.5
    def two(first, second):
        st = {k: np.concatenate([_one(first)[k], _one(second)[k]]) for k in _one(first)}
        return orc.render_splats(st, c.getViewMatrix(), c.getProjectionMatrix(), 2.0, 100.0, 10.0, 2, w, h)[0]
    ab, ba = two(a, b), two(b, a)
    assert (ab != ba).any()
    # an opaque fragment drawn last hides everything under it
    assert ba[h // 2, w // 2].tolist() == two(a, a)[h // 2, w // 2].tolist()
    # a 1-px sprite whose centre is more than half a pixel from the pixel centre is discarded by the shader
    far = _one(_at(c, w, h, w // 2 + 0.01, h // 2 + 0.01, 50.0))
    img, frags = orc.render_splats(far, c.getViewMatrix(), c.getProjectionMatrix(), 2.0, 100.0, 10.0, 0, w, h)
    assert frags == 1 and (img == img[0, 0]).all()


def test_large_sprites_are_round_with_soft_edges(orc):
    w, h = 96, 96
    c = _cam(w, h)
    st = _one((0.0, 0.0, 45.0))                   # depth 5 -> 200 / 5 = 40 px sprite
    img, frags = orc.render_splats(st, c.getViewMatrix(), c.getProjectionMatrix(), 200.0, 100.0, 10.0, 0, w, h)
    assert frags == 40 * 40
    alpha_like = (img[..., :3].astype(int) - img[0, 0, :3].astype(int)).sum(axis=2)
    lit = (alpha_like != 0)
    assert abs(lit.sum() - math.pi * 20 ** 2) < 60          # a disc of radius 20 px, not a square
    assert not lit[48 - 19, 48 - 19] and lit[48, 48 - 19]    # corner discarded, edge midpoint kept
