"""CPU checks of the ray-caster restatement (oracle/render_ref.py) against closed forms: silhouette of a sphere under a
pinhole camera, head-light shading at normal incidence, occlusion order, finite planes, box faces."""
import numpy as np

from oracle import render_ref as R

I3 = np.eye(3)


def test_sphere_silhouette_and_head_light():
    # camera 2 m above the origin looking down (-z), fovy 30 deg, 65 x 65 pixels: the centre pixel looks straight down
    r, D, fovy, n = 0.1, 2.0, 30.0, 65
    img = R.render((0, 0, D), I3, fovy, n, n, [(R.SPHERE, [r, 0, 0], (0, 0, 0), I3, (1, 0, 0))])
    c = n // 2
    assert tuple(img[c, c]) == (255, 0, 0)  # normal incidence: 0.3 + 0.7 = 1
    assert img[:, :, 1:].max() == 0
    # silhouette: rays tangent to the sphere make the angle asin(r / D) with the axis
    th = np.tan(np.deg2rad(fovy / 2))
    px = np.tan(np.arcsin(r / D)) / th * (n / 2)  # radius in pixels
    hit = img[:, :, 0] > 0
    cols = np.nonzero(hit[c])[0]
    assert abs((cols.max() - cols.min() + 1) / 2 - px) < 1.0
    assert abs(hit.sum() - np.pi * px * px) < 0.08 * np.pi * px * px
    # limb darkening: the rim tends to the ambient term
    assert 0.3 * 255 - 1 <= img[c, cols.min(), 0] < 0.6 * 255


def test_occlusion_finite_plane_and_background():
    plane = (R.PLANE, [0.5, 0.5, 0.1], (0, 0, 0), I3, (0.3, 0.3, 0.3))
    box = (R.BOX, [0.1, 0.1, 0.05], (0, 0, 0.05), I3, (0, 1, 0))
    img = R.render((0, 0, 3.0), I3, 40.0, 64, 64, [plane, box])
    assert tuple(img[32, 32]) == (0, 255, 0)                       # the box top hides the plane at the centre
    v = int(np.floor(255 * 0.3 * (0.3 + 0.7 * np.cos(np.arctan(0.4 / 3.0))) + 0.5))
    assert abs(int(img[32, 32 + 9, 0]) - v) <= 1                    # plane next to the box: grey scaled by the head-light term
    assert tuple(img[0, 0]) == (0, 0, 0)                            # beyond the finite plane: background
    img2 = R.render((0, 0, 3.0), I3, 40.0, 64, 64, [box, plane])    # order of the list does not matter
    np.testing.assert_array_equal(img, img2)


def test_cylinder_and_capsule_profiles_from_the_side():
    # camera on the -y axis looking along +y: rotate so that -z_cam = +y_world, up = +z_world
    cam_mat = np.array([[1, 0, 0], [0, 0, -1], [0, 1, 0]], float)
    for gtype, extra in ((R.CYLINDER, 0.0), (R.CAPSULE, 0.05)):
        img = R.render((0, -2.0, 0), cam_mat, 20.0, 101, 101, [(gtype, [0.05, 0.2, 0], (0, 0, 0), I3, (1, 1, 1))])
        hit = img[:, :, 0] > 0
        rows = np.nonzero(hit[:, 50])[0]; cols = np.nonzero(hit[50])[0]
        th = np.tan(np.deg2rad(10.0)); ppm = 50.5 / (th * (2.0 - 0.05))  # pixels per metre at the front surface
        assert abs((rows.max() - rows.min() + 1) / 2 - (0.2 + extra) * ppm) < 2.0
        assert abs((cols.max() - cols.min() + 1) / 2 - 0.05 * ppm * (2.0 - 0.05) / 2.0 * 1.0) < 1.5
        assert img[50, 50, 0] == 255
