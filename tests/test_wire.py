"""The reference's f32 wire format (GpuSerialize): decoders of the C ABI against serializers
written from the reference's layouts (raytracing/camera.rs:70-78, scene.rs:67-76, gpu/object.rs:34-42,
object/sphere.rs:36-42, plane.rs:38-46, triangle.rs:130-136).  CPU part: pure host decoding.  GPU part: a scene
ingested from wire bytes renders the same bytes as the scene built from the f32-rounded values."""
import struct

import numpy as np
import pytest


def _demo(rtx):
    from rtx_b200 import scenes
    sc = scenes.sphere_demo(rays_per_pixel=2, max_bounces=4)
    return sc, sc.objects_flat()


def test_decode_camera_and_config(rtx):
    from rtx_b200 import wire
    pos, direction, fov = (0.1, -2.5, 1.0 / 3.0), (1.0, 0.2, -0.1), 1.2345678
    data = wire.serialize_camera(pos, direction, fov)
    assert len(data) == 32 and data[12:16] == b"\0" * 4
    c = wire.decode_camera(data)
    f = lambda v: float(np.float32(v))  # noqa: E731
    assert (c.position.x, c.position.y, c.position.z) == tuple(f(v) for v in pos)
    assert (c.direction.x, c.direction.y, c.direction.z) == tuple(f(v) for v in direction)
    assert c.fov == f(fov)
    # Camera::new ran on the widened values (camera.rs:19-28): same matrices as the ABI's own constructor
    ref = rtx.Camera(tuple(f(v) for v in pos), tuple(f(v) for v in direction), f(fov))
    assert bytes(c) == bytes(ref._c)
    cfg = wire.decode_config(wire.serialize_config(64, 8, 10.0, 1e-4, 1e-1))
    assert (cfg.rays_per_pixel, cfg.max_bounces) == (64, 8)
    assert (cfg.focal_length, cfg.focal_offset, cfg.non_focal_offset) == (10.0, f(1e-4), f(1e-1))
    # usize -> u32 truncation of the reference's serializer (scene.rs:69-70)
    assert struct.unpack("<I", wire.serialize_config(2**32 + 5, 1, 1, 1, 1)[:4])[0] == 5


def test_decode_objects_roundtrip(rtx):
    from rtx_b200 import wire
    rng = np.random.default_rng(3)
    n = 50
    kinds = rng.integers(0, 3, size=n)
    shapes = rng.normal(size=(n, 9)) * 5
    mats = rng.uniform(size=(n, 7))
    ncol = {0: 4, 1: 6, 2: 9}
    records, buffers = wire.serialize_scene(kinds, shapes, mats)
    assert len(records) == 48 * n and sum(c for _, _, c in buffers) == n
    # shape ids are numbered in order of first appearance (gpu_state.rs:57-68)
    first = []
    for k in kinds.tolist():
        if k not in first:
            first.append(k)
    assert [t for t, _, _ in buffers] == [wire.TAGS[k] for k in first]
    k2, s2, m2 = wire.decode_objects(records, buffers)
    assert np.array_equal(k2, kinds.astype(np.uint32))
    for i in range(n):
        c = ncol[int(kinds[i])]
        assert np.array_equal(s2[i, :c], shapes[i, :c].astype(np.float32).astype(np.float64))
        assert not s2[i, c:].any()
    assert np.array_equal(m2, mats.astype(np.float32).astype(np.float64))


def test_decode_rejects_bad_input(rtx):
    from rtx_b200 import wire
    rec = wire.serialize_object_record(np.zeros(7), 0, 3)
    with pytest.raises(rtx.RtxError):  # index outside the shape buffer
        wire.decode_objects(rec, [(wire.TAGS[0], wire.serialize_shape(0, np.zeros(4)), 1)])
    with pytest.raises(rtx.RtxError) as e:  # a user-defined GpuShape: WGSL source, not data
        wire.decode_objects(wire.serialize_object_record(np.zeros(7), 0, 0), [("my_crate::torus", b"\0" * 16, 1)])
    assert e.value.status == 4  # RTX_ERR_UNSUPPORTED_SHAPE


@pytest.mark.gpu
def test_wire_scene_renders_like_the_f32_rounded_scene(rtx, oracle, gpu):
    import ctypes as C

    from rtx_b200 import _abi, wire
    sc, (types, shapes, mats) = _demo(rtx)
    cam, cfg = sc.camera, sc.config
    pos = (cam._c.position.x, cam._c.position.y, cam._c.position.z)
    dr = (cam._c.direction.x, cam._c.direction.y, cam._c.direction.z)
    records, buffers = wire.serialize_scene(types, shapes, mats)
    cam_b = wire.serialize_camera(pos, dr, cam._c.fov)
    cfg_b = wire.serialize_config(cfg.rays_per_pixel, cfg.max_bounces, cfg.focal_length, cfg.focal_offset,
                                  cfg.non_focal_offset)
    a = rtx.Scene.from_wire(cam_b, cfg_b, records, buffers)
    # the same scene built from structs holding the f32-rounded values
    r32 = lambda x: np.asarray(x, dtype=np.float64).astype(np.float32).astype(np.float64)  # noqa: E731
    cc = wire.decode_config(cfg_b)
    b = rtx.Scene(rtx.Config(cc.rays_per_pixel, cc.max_bounces, cc.focal_length, cc.focal_offset, cc.non_focal_offset),
                  rtx.Camera(tuple(r32(pos)), tuple(r32(dr)), float(np.float32(cam._c.fov))))
    ncol = {0: 4, 1: 6, 2: 9}
    for k, s, m in zip(types.tolist(), shapes, mats):
        b._append(int(k), r32(s[None, :ncol[int(k)]]), r32(m[None, :]))
    ia, ida = a.render(200, 120, seed=4, want_ids=True)
    ib, idb = b.render(200, 120, seed=4, want_ids=True)
    assert ia.tobytes() == ib.tobytes() and np.array_equal(ida, idb) and np.isfinite(ia).all() and ia.any()
    # ... and both equal the CPU oracle's render of the f32-rounded scene (each scene's own host-side object list is
    # what the oracle is fed): ids, ray count, every radiance byte
    for s_ in (a, b):
        ref = oracle.render_scene(s_, 200, 120, seed=4)
        assert np.array_equal(ida, ref["ids"]) and ia.tobytes() == ref["rgb"].tobytes()
        assert s_.last_stats.rays == ref["rays"]
    # and straight through the C entry points on a raw handle
    lib = _abi.load()
    h = C.c_void_p()
    dc, dcfg = wire.decode_camera(cam_b), wire.decode_config(cfg_b)
    _abi.check(lib.rtx_scene_create(C.byref(dc), C.byref(dcfg), 0, C.byref(h)))
    try:
        arr, keep = wire._shape_buffers(buffers)
        _abi.check(lib.rtx_scene_set_wire_camera(h, cam_b), h)
        _abi.check(lib.rtx_scene_set_wire_config(h, cfg_b), h)
        _abi.check(lib.rtx_scene_add_wire_objects(h, records, len(records) // 48, arr, len(buffers)), h)
        out = np.empty((120, 200, 3), dtype=np.float64)
        _abi.check(lib.rtx_scene_render(h, 200, 120, 4, 0, out.ctypes.data, None, None), h)
        assert out.tobytes() == ia.tobytes()
    finally:
        lib.rtx_scene_destroy(h)


@pytest.mark.gpu
def test_wire_scene_previews_as_the_wgpu_backend_would(rtx, oracle, gpu):
    """The reference's own use of these bytes: gpu::Scene uploads them and base_shader.wgsl renders them.  A scene
    ingested from the f32 wire format and rendered with rtx_scene_render_preview_f32 equals the WGSL oracle's render of
    those f32 values — narrowing the widened doubles back to f32 is exact, so nothing is rounded twice."""
    from rtx_b200 import wire
    sc, (types, shapes, mats) = _demo(rtx)
    cam, cfg = sc.camera, sc.config
    pos = (cam._c.position.x, cam._c.position.y, cam._c.position.z)
    dr = (cam._c.direction.x, cam._c.direction.y, cam._c.direction.z)
    records, buffers = wire.serialize_scene(types, shapes, mats)
    a = rtx.Scene.from_wire(wire.serialize_camera(pos, dr, cam._c.fov),
                            wire.serialize_config(cfg.rays_per_pixel, cfg.max_bounces, cfg.focal_length, cfg.focal_offset,
                                                  cfg.non_focal_offset), records, buffers)
    rgb, ids = a.render_preview(160, 96, want_ids=True)
    ref = oracle.render_preview_scene(a, 160, 96)
    assert np.array_equal(ids, ref["ids"]) and rgb.tobytes() == ref["rgb"].tobytes() and a.last_stats.rays == ref["rays"]
    # the struct-built scene holds the unrounded doubles: its preview narrows them itself — to the same f32 values
    rgb2 = sc.render_preview(160, 96)
    assert rgb2.tobytes() == rgb.tobytes()
