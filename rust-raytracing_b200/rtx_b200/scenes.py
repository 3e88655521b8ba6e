"""Canonical synthetic scenes S1-S3.

The reference ships no demo scene (no main.rs / examples/ / tests/), so BASELINE.json's "built-in sphere
demo scene" is defined here; once fixed in code these are the contract.  All random draws come from
splitmix64 with u01 = (u >> 11) * 2^-53, done in exact integer / IEEE arithmetic, so every consumer
(GPU path, CPU oracle) is fed bit-identical arrays.
"""
import math

import numpy as np

from . import Camera, Config, Material, Object, Plane, Scene, Sphere

_MASK = (1 << 64) - 1


def splitmix64_stream(seed, n):
    """n successive splitmix64 outputs as uint64 (vectorised)."""
    idx = np.arange(1, n + 1, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = np.uint64(seed & _MASK) + idx * np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def u01(u):
    return (u >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def sphere_demo(rays_per_pixel=64, max_bounces=8, device=0):
    """S1 "sphere demo" — configs C1, C2, C5.  7 objects: ground plane, 4 diffuse/glossy spheres, 2 lights."""
    cam = Camera((0, 0, 1), (1, 0, 0), math.radians(90.0))
    cfg = Config().with_rays_per_pixel(rays_per_pixel).with_max_bounces(max_bounces)
    sc = Scene(cfg, cam, device=device)
    sc.add_object(Object(Plane((0, 0, 0), (0, 0, 1)), Material((0.8, 0.8, 0.8), (0, 0, 0), 1.0)))
    sc.add_object(Object(Sphere((6, 0, 1), 1), Material((0.9, 0.2, 0.2), (0, 0, 0), 1.0)))
    sc.add_object(Object(Sphere((6, -2.5, 1), 1), Material((0.9, 0.9, 0.9), (0, 0, 0), 0.05)))
    sc.add_object(Object(Sphere((6, 2.5, 1), 1), Material((0.2, 0.9, 0.2), (0, 0, 0), 0.5)))
    sc.add_object(Object(Sphere((4, 1.2, 0.4), 0.4), Material((0.2, 0.3, 0.9), (0, 0, 0), 1.0)))
    sc.add_object(Object(Sphere((4, -1.2, 0.4), 0.4), Material.light((2, 1.6, 1))))
    sc.add_object(Object(Sphere((5, 0, 12), 6), Material.light((4, 4, 4))))
    return sc


def random_spheres(n=10_000, rays_per_pixel=16, max_bounces=8, seed=0x5EED0001, device=0):
    """S2 "10k spheres" — config C3.  Per sphere, in draw order: cx∈[8,88) cy∈[-40,40) cz∈[-40,40) r∈[0.2,1)
    base∈[0.2,0.95)^3 rough∈[0,1) e; e<0.1 → light((3,3,3)*u01) else Material(base, 0, rough)."""
    draws_per = 10
    u = u01(splitmix64_stream(seed, n * draws_per)).reshape(n, draws_per)
    centers = np.stack([8.0 + 80.0 * u[:, 0], -40.0 + 80.0 * u[:, 1], -40.0 + 80.0 * u[:, 2]], axis=1)
    radii = 0.2 + 0.8 * u[:, 3]
    base = 0.2 + 0.75 * u[:, 4:7]
    rough = u[:, 7]
    is_light = u[:, 8] < 0.1
    power = 3.0 * u[:, 9]
    mats = np.zeros((n, 7), dtype=np.float64)
    mats[:, 0:3] = np.where(is_light[:, None], 0.0, base)
    mats[:, 3:6] = np.where(is_light[:, None], power[:, None], 0.0)
    mats[:, 6] = np.where(is_light, 1.0, rough)
    cam = Camera((0, 0, 0), (1, 0, 0), math.pi / 2)
    cfg = Config().with_rays_per_pixel(rays_per_pixel).with_max_bounces(max_bounces)
    sc = Scene(cfg, cam, device=device)
    sc.add_spheres(centers, radii, mats)
    return sc


def terrain(grid=709, rays_per_pixel=1, max_bounces=4, seed=0x5EED0002, device=0):
    """S3 "1M-triangle terrain" — config C4.  Height field over a grid×grid vertex lattice on [-30,30]^2,
    z = 20 + sin(0.25x)cos(0.2y) + 0.002(u01-0.5); (grid-1)^2*2 triangles wound so the normal is +z, then
    one light sphere.  The reference's facing test n·(v0 - dir) >= 0 (triangle.rs:115) passes for every unit
    dir only if n·v0 >= 1; the generator asserts it."""
    xs = np.linspace(-30.0, 30.0, grid)
    ys = np.linspace(-30.0, 30.0, grid)
    X, Y = np.meshgrid(xs, ys, indexing="xy")  # row-major: Y varies along axis 0
    jitter = u01(splitmix64_stream(seed, grid * grid)).reshape(grid, grid)
    Z = 20.0 + 1.0 * np.sin(0.25 * X) * np.cos(0.2 * Y) + 0.002 * (jitter - 0.5)
    P = np.stack([X, Y, Z], axis=-1)  # [gy][gx][3]
    p00, p10 = P[:-1, :-1], P[:-1, 1:]
    p01, p11 = P[1:, :-1], P[1:, 1:]
    # two CCW (seen from +z) triangles per cell: (p00,p10,p11) and (p00,p11,p01)
    t1 = np.stack([p00, p10, p11], axis=2)
    t2 = np.stack([p00, p11, p01], axis=2)
    tris = np.stack([t1, t2], axis=2).reshape(-1, 3, 3)
    r, s = tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]
    nrm = np.cross(r, s)
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    assert (nrm[:, 2] > 0).all(), "winding must give +z normals"
    assert ((nrm * tris[:, 0]).sum(axis=1) >= 1.0).all(), "facing test (triangle.rs:115) would hide triangles"
    cam = Camera((-25, 0, 50), (1, 0, -1), math.pi / 2)
    cfg = Config().with_rays_per_pixel(rays_per_pixel).with_max_bounces(max_bounces)
    sc = Scene(cfg, cam, device=device)
    sc.add_triangles(tris, np.array([[0.7, 0.7, 0.7, 0, 0, 0, 1.0]]))
    sc.add_object(Object(Sphere((0, 0, 140), 40), Material.light((4, 4, 4))))
    return sc
