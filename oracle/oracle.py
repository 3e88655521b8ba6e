"""ctypes loader for the CPU oracle (oracle/rtx_oracle.cpp → oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: import from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs — never from rust-raytracing_b200/.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")

RNG_PIXEL_SAMPLE, RNG_ROW = 0, 1
SHAPE_SPHERE, SHAPE_PLANE, SHAPE_TRIANGLE = 0, 1, 2
MISS_ID = 0xFFFFFFFF


class OCamera(C.Structure):
    _fields_ = [("fov", C.c_double), ("position", C.c_double * 3), ("direction", C.c_double * 3),
                ("to_cam_space", C.c_double * 9), ("to_world_space", C.c_double * 9)]


class OConfig(C.Structure):
    _fields_ = [("rays_per_pixel", C.c_uint64), ("max_bounces", C.c_uint64), ("focal_length", C.c_double),
                ("focal_offset", C.c_double), ("non_focal_offset", C.c_double)]


OBJECT_DTYPE = np.dtype([("type", np.int32), ("pad", np.int32), ("shape", np.float64, 9),
                         ("base_color", np.float64, 3), ("emission_color", np.float64, 3),
                         ("roughness", np.float64)], align=True)

_lib = None


WGSL_LIB_PATH = os.path.join(_HERE, "libwgsl_oracle.so")
_wgsl = None
WGSL_OBJECT_DTYPE = np.dtype([("base", "<f4", 3), ("roughness", "<f4"), ("emission", "<f4", 3), ("type", "<u4"),
                              ("shape", "<f4", 9)])


def build(force=False):
    for lib, src in ((LIB_PATH, "rtx_oracle.cpp"), (WGSL_LIB_PATH, "wgsl_oracle.cpp")):
        if force or not os.path.exists(lib) or os.path.getmtime(lib) < os.path.getmtime(os.path.join(_HERE, src)):
            subprocess.check_call(["make", "-C", _HERE, "-s", "-B", os.path.basename(lib)])
    return LIB_PATH


def load_wgsl():
    """The f32 restatement of the wgpu backend's shader arithmetic (wgsl_oracle.cpp)."""
    global _wgsl
    if _wgsl is None:
        build()
        lib = C.CDLL(WGSL_LIB_PATH)
        lib.wgsl_render.restype = C.c_int
        lib.wgsl_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32,
                                    C.c_int32, C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        lib.wgsl_rng_probe.argtypes = [C.c_float, C.c_float, C.c_uint32, C.POINTER(C.c_uint32), C.c_void_p, C.c_void_p]
        lib.wgsl_sizeof_object.restype = C.c_uint64
        assert lib.wgsl_sizeof_object() == WGSL_OBJECT_DTYPE.itemsize
        _wgsl = lib
    return _wgsl


def wgsl_rng_probe(ux, uy, n):
    """(seed, random_int draws uint32[n], random_float draws float32[n]) of the pixel whose uv is (ux, uy)."""
    lib = load_wgsl()
    seed = C.c_uint32()
    u = np.empty(n, dtype=np.uint32)
    f = np.empty(n, dtype=np.float32)
    lib.wgsl_rng_probe(float(ux), float(uy), n, C.byref(seed), u.ctypes.data, f.ctypes.data)
    return seed.value, u, f


def render_preview_scene(scene, width, height, nthreads=0):
    """wgsl_oracle render of an rtx_b200.Scene narrowed to f32 as GpuSerialize does (`as f32`): dict(rgb float32
    [h][w][3] row 0 = bottom, ids uint32 [h][w], rays)."""
    lib = load_wgsl()
    types, shapes, mats = scene.objects_flat()
    objs = np.zeros(len(types), dtype=WGSL_OBJECT_DTYPE)
    with np.errstate(over="ignore"):
        objs["type"] = types
        objs["shape"] = shapes.astype(np.float32)
        objs["base"] = mats[:, 0:3].astype(np.float32)
        objs["emission"] = mats[:, 3:6].astype(np.float32)
        objs["roughness"] = mats[:, 6].astype(np.float32)
        cam = scene.camera
        camv = np.array([*cam.position, *cam.get_direction(), cam.fov], dtype=np.float64).astype(np.float32)
        c = scene.config
        cfg_f = np.array([c.focal_length, c.focal_offset, c.non_focal_offset], dtype=np.float64).astype(np.float32)
    cfg_u = np.array([c.rays_per_pixel, c.max_bounces], dtype=np.uint32)
    rgb = np.empty((height, width, 3), dtype=np.float32)
    ids = np.empty((height, width), dtype=np.uint32)
    rays = C.c_uint64()
    rc = lib.wgsl_render(camv.ctypes.data, cfg_u.ctypes.data, cfg_f.ctypes.data, objs.ctypes.data if len(objs) else None,
                         len(objs), width, height, nthreads, rgb.ctypes.data, ids.ctypes.data, C.byref(rays))
    assert rc == 0
    return dict(rgb=rgb, ids=ids, rays=rays.value)


def load():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        build()
    lib = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    lib.oracle_render.restype = C.c_int
    lib.oracle_render.argtypes = [C.POINTER(OCamera), C.POINTER(OConfig), vp, C.c_uint64, C.c_uint32, C.c_uint32,
                                  C.c_uint64, C.c_int32, C.c_int32, C.c_uint32, C.c_uint32, vp, vp, vp,
                                  C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    lib.oracle_to_rgb8.argtypes = [vp, C.c_uint32, C.c_uint32, vp]
    lib.oracle_hardware_threads.restype = C.c_uint
    lib.oracle_wyrand_u64.argtypes = [C.POINTER(C.c_uint64), C.c_uint64, vp]
    lib.oracle_wyrand_f64.argtypes = [C.POINTER(C.c_uint64), C.c_uint64, vp]
    lib.oracle_stream_state.restype = C.c_uint64
    lib.oracle_stream_state.argtypes = [C.c_uint64, C.c_uint64]
    lib.oracle_camera_new.argtypes = [vp, vp, C.c_double, C.POINTER(OCamera)]
    lib.oracle_camera_set_direction.argtypes = [C.POINTER(OCamera), vp]
    for f in ("oracle_camera_to_cam_space", "oracle_camera_to_world_space", "oracle_camera_rotate_to_world_space"):
        getattr(lib, f).argtypes = [C.POINTER(OCamera), vp, vp]
    lib.oracle_get_ray_dir.argtypes = [C.POINTER(OCamera), C.c_double, C.c_double, C.c_double, vp]
    lib.oracle_mat_inverse.argtypes = [vp, vp]
    lib.oracle_mat_determinant.restype = C.c_double
    lib.oracle_mat_determinant.argtypes = [vp]
    lib.oracle_object_distance.argtypes = [vp, vp, C.c_uint64, vp, vp]
    lib.oracle_object_normal_at.argtypes = [vp, vp, C.c_uint64, vp]
    lib.oracle_closest_object.argtypes = [vp, C.c_uint64, vp, C.c_uint64, vp, vp]
    lib.oracle_random_bounce_dir.argtypes = [vp, vp, C.c_uint64, vp]
    lib.oracle_random_direction.argtypes = [C.POINTER(C.c_uint64), C.c_uint64, vp]
    for f in ("oracle_sizeof_object", "oracle_sizeof_camera", "oracle_sizeof_config"):
        getattr(lib, f).restype = C.c_uint64
    assert lib.oracle_sizeof_object() == OBJECT_DTYPE.itemsize, (lib.oracle_sizeof_object(), OBJECT_DTYPE.itemsize)
    assert lib.oracle_sizeof_camera() == C.sizeof(OCamera) and lib.oracle_sizeof_config() == C.sizeof(OConfig)
    _lib = lib
    return lib


def _arr3(v):
    return np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(3))


def camera_new(position, direction, fov):
    cam = OCamera()
    p, d = _arr3(position), _arr3(direction)
    load().oracle_camera_new(p.ctypes.data, d.ctypes.data, float(fov), C.byref(cam))
    return cam


def camera_set_direction(cam, direction):
    d = _arr3(direction)
    load().oracle_camera_set_direction(C.byref(cam), d.ctypes.data)


def _cam_xf(name, cam, v):
    v, out = _arr3(v), np.empty(3)
    getattr(load(), name)(C.byref(cam), v.ctypes.data, out.ctypes.data)
    return out


def camera_to_cam_space(cam, v):
    return _cam_xf("oracle_camera_to_cam_space", cam, v)


def camera_to_world_space(cam, v):
    return _cam_xf("oracle_camera_to_world_space", cam, v)


def camera_rotate_to_world_space(cam, v):
    return _cam_xf("oracle_camera_rotate_to_world_space", cam, v)


def camera_from_bytes(raw):
    """OCamera from the 200 bytes of an rtx_camera (identical layout)."""
    cam = OCamera()
    assert len(raw) == C.sizeof(OCamera)
    C.memmove(C.byref(cam), raw, len(raw))
    return cam


def make_objects(types, shapes, materials):
    """types[n], shapes[n,9], materials[n,7] → structured OObject array."""
    n = len(types)
    objs = np.zeros(n, dtype=OBJECT_DTYPE)
    objs["type"] = types
    objs["shape"] = shapes
    objs["base_color"] = materials[:, 0:3]
    objs["emission_color"] = materials[:, 3:6]
    objs["roughness"] = materials[:, 6]
    return objs


def render(cam, cfg, objs, width, height, seed=0, rng_mode=RNG_PIXEL_SAMPLE, nthreads=0, xs=1, ys=1):
    """Returns dict(rgb[h',w',3], ids[h',w'], rays_per_pixel[h',w'], rays, panics)."""
    lib = load()
    ow, oh = (width + xs - 1) // xs, (height + ys - 1) // ys
    rgb = np.empty((oh, ow, 3), dtype=np.float64)
    ids = np.empty((oh, ow), dtype=np.uint32)
    rpp = np.empty((oh, ow), dtype=np.uint32)
    total, panics = C.c_uint64(), C.c_uint64()
    objs = np.ascontiguousarray(objs)
    rc = lib.oracle_render(C.byref(cam), C.byref(cfg), objs.ctypes.data if len(objs) else None, len(objs), width,
                           height, seed, rng_mode, nthreads, xs, ys, rgb.ctypes.data, ids.ctypes.data,
                           rpp.ctypes.data, C.byref(total), C.byref(panics))
    assert rc == 0
    return dict(rgb=rgb, ids=ids, rays_per_pixel=rpp, rays=total.value, panics=panics.value)


def render_scene(scene, width, height, seed=0, **kw):
    """Oracle render of an rtx_b200.Scene (same camera struct bytes, config and object list)."""
    cam = camera_from_bytes(bytes(scene.camera._c))
    c = scene.config
    cfg = OConfig(c.rays_per_pixel, c.max_bounces, c.focal_length, c.focal_offset, c.non_focal_offset)
    objs = make_objects(*scene.objects_flat())
    return render(cam, cfg, objs, width, height, seed, **kw)


def to_rgb8(rgb):
    h, w, _ = rgb.shape
    rgb = np.ascontiguousarray(rgb, dtype=np.float64)
    out = np.empty((h, w, 3), dtype=np.uint8)
    load().oracle_to_rgb8(rgb.ctypes.data, w, h, out.ctypes.data)
    return out


def wyrand_u64(state, n):
    st = C.c_uint64(state)
    out = np.empty(n, dtype=np.uint64)
    load().oracle_wyrand_u64(C.byref(st), n, out.ctypes.data)
    return out, st.value


def wyrand_f64(state, n):
    st = C.c_uint64(state)
    out = np.empty(n, dtype=np.float64)
    load().oracle_wyrand_f64(C.byref(st), n, out.ctypes.data)
    return out, st.value


def stream_state(seed, key):
    return load().oracle_stream_state(seed, key)


def object_distance(obj, rays):
    """obj: 1-element OObject array; rays[n,6] → (t[n], some[n] bool)."""
    rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
    t = np.empty(len(rays))
    some = np.empty(len(rays), dtype=np.uint8)
    load().oracle_object_distance(obj.ctypes.data, rays.ctypes.data, len(rays), t.ctypes.data, some.ctypes.data)
    return t, some.astype(bool)


def object_normal_at(obj, positions):
    positions = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
    out = np.empty_like(positions)
    load().oracle_object_normal_at(obj.ctypes.data, positions.ctypes.data, len(positions), out.ctypes.data)
    return out


def closest_object(objs, rays):
    rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
    objs = np.ascontiguousarray(objs)
    t = np.empty(len(rays))
    ids = np.empty(len(rays), dtype=np.uint32)
    load().oracle_closest_object(objs.ctypes.data, len(objs), rays.ctypes.data, len(rays), t.ctypes.data,
                                 ids.ctypes.data)
    return t, ids


def random_bounce_dir(states, inputs):
    states = np.ascontiguousarray(states, dtype=np.uint64).copy()
    inputs = np.ascontiguousarray(inputs, dtype=np.float64).reshape(-1, 7)
    out = np.empty((len(inputs), 3))
    load().oracle_random_bounce_dir(states.ctypes.data, inputs.ctypes.data, len(inputs), out.ctypes.data)
    return out, states


def random_direction(state, n):
    st = C.c_uint64(state)
    out = np.empty((n, 3))
    load().oracle_random_direction(C.byref(st), n, out.ctypes.data)
    return out, st.value


def hardware_threads():
    return load().oracle_hardware_threads()
