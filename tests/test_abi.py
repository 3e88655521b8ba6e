"""CPU tests of the drop-in boundary: librtx_cuda.so loads, exports every symbol include/rtx_cuda.h
declares, mirrors have the reference's layout, host helpers match the oracle bit for bit, and the render
entry points fail loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "rtx_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rtx_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(rtx):
    lib = rtx._abi.load()
    names = _declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in rtx_cuda.h but not exported"
    # and the Python binding declares a signature for each of them
    assert set(names) == set(rtx._abi.SIGNATURES), set(names) ^ set(rtx._abi.SIGNATURES)
    assert lib.rtx_abi_version() == 1


def test_mirror_layouts(rtx):
    a = rtx._abi
    sizes = {a.Vec3: 24, a.Mat3: 72, a.CameraC: 200, a.ConfigC: 40, a.MaterialC: 56, a.SphereC: 32, a.PlaneC: 48,
             a.TriangleC: 72}
    for t, s in sizes.items():
        assert C.sizeof(t) == s
    assert a.CameraC.position.offset == 8 and a.CameraC.to_world_space.offset == 128


def test_config_default(rtx):  # scene.rs:55-65
    c = rtx._abi.ConfigC()
    assert rtx._abi.load().rtx_config_default(C.byref(c)) == 0
    assert (c.rays_per_pixel, c.max_bounces, c.focal_length, c.focal_offset, c.non_focal_offset) == \
        (16, 10, 10.0, 1e-4, 1e-1)
    d = rtx.Config()
    assert (d.rays_per_pixel, d.max_bounces) == (16, 10)
    e = d.with_rays_per_pixel(32).with_max_bounces(3)
    assert (e.rays_per_pixel, e.max_bounces, d.rays_per_pixel) == (32, 3, 16)  # builders clone (scene.rs:29-54)


def test_camera_reference_tests_through_the_abi(rtx):  # camera.rs:79-110
    V = rtx.Vector3
    cam = rtx.Camera(V.zeros(), V.x_axis(), math.radians(90.0))
    assert cam.to_cam_space(V.x_axis()) == V.z_axis()
    assert cam.to_cam_space(V.y_axis()) == V.x_axis()
    assert cam.to_cam_space(V.z_axis()) == V.y_axis()
    assert cam.to_world_space(V.x_axis()) == V.y_axis()
    assert cam.to_world_space(V.y_axis()) == V.z_axis()
    assert cam.to_world_space(V.z_axis()) == V.x_axis()
    cam = rtx.Camera(V.zeros(), V.y_axis(), math.radians(90.0))
    assert cam.to_world_space(V.x_axis()) == -V.x_axis()
    assert cam.to_world_space(V.y_axis()) == V.z_axis()
    assert cam.to_world_space(V.z_axis()) == V.y_axis()
    assert cam.to_cam_space(V.x_axis()) == -V.x_axis()
    assert cam.to_cam_space(V.y_axis()) == V.z_axis()
    assert cam.to_cam_space(V.z_axis()) == V.y_axis()


def test_camera_helpers_match_oracle_bitwise(rtx, oracle):
    rng = np.random.default_rng(11)
    for _ in range(100):
        p, d, fov = rng.normal(size=3), rng.normal(size=3), float(rng.uniform(0.2, 2.5))
        cam = rtx.Camera(p, d, fov)
        oc = oracle.camera_new(p, d, fov)
        assert bytes(cam._c) == bytes(oc)
        d2 = rng.normal(size=3)
        cam.set_direction(d2)
        oracle.camera_set_direction(oc, d2)
        assert bytes(cam._c) == bytes(oc)
        v = rng.normal(size=3)
        assert tuple(cam.to_cam_space(v)) == tuple(oracle.camera_to_cam_space(oc, v))
        assert tuple(cam.to_world_space(v)) == tuple(oracle.camera_to_world_space(oc, v))
        assert tuple(cam.rotate_to_world_space(v)) == tuple(oracle.camera_rotate_to_world_space(oc, v))


def test_stream_state_matches_oracle(rtx, oracle):
    for seed, key in ((0, 0), (1, 1), (7, 2**33), (2**64 - 1, 2**40)):
        assert rtx.stream_state(seed, key) == oracle.stream_state(seed, key)


def test_shard_tile_count(rtx):
    # 1920x1080 → 120 x 68 = 8160 tiles
    assert rtx.shard_tile_count(1920, 1080, 0, 1) == 8160
    for n in (2, 3, 4, 8):
        counts = [rtx.shard_tile_count(1920, 1080, i, n) for i in range(n)]
        assert sum(counts) == 8160 and max(counts) - min(counts) <= 1 and counts[0] == max(counts)
    assert rtx.shard_tile_count(16, 16, 1, 2) == 0  # one tile, second shard owns none
    assert rtx.shard_tile_count(17, 1, 1, 2) == 1


def test_user_defined_shape_is_rejected(rtx):  # object.rs:53-76 cannot cross the ABI
    class Torus:
        pass
    with pytest.raises(rtx.RtxError) as e:
        rtx.Object(Torus(), rtx.Material.mirror())
    assert e.value.status == rtx._abi.RTX_ERR_UNSUPPORTED_SHAPE


def test_material_constructors(rtx):  # object.rs:92-135
    m = rtx.Material.colored((1, 1, 0))
    assert m.row() == [1, 1, 0, 0, 0, 0, 1.0]
    m = rtx.Material.light((1, 0.8, 0.5))
    assert m.row() == [0, 0, 0, 1, 0.8, 0.5, 1.0]
    assert rtx.Material.mirror().row() == [1, 1, 1, 0, 0, 0, 1.0]  # roughness 1.0 on the CPU path


def test_render_fails_loudly_without_a_gpu(rtx):
    if rtx.device_count() > 0:
        pytest.skip("a GPU is present; the loud-failure path is exercised on CPU-only boxes")
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=1)
    with pytest.raises(rtx.RtxError) as e:
        sc.render(16, 16, seed=0)
    assert e.value.status == rtx._abi.RTX_ERR_CUDA
    assert "no CPU fallback" in str(e.value)
    with pytest.raises(rtx.RtxError):
        rtx.probe_fp64_peak()
