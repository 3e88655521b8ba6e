"""ctypes binding of the C ABI in include/rtx_cuda.h (librtx_cuda.so).

This module is the Python analogue of the `rtx-cuda-sys` Rust crate described in INTEGRATION.md: struct
mirrors + the extern "C" declarations, nothing else.  It fails loudly when the CUDA library has not been
built — there is no CPU fallback behind it.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "..", "lib", "librtx_cuda.so")


class Vec3(C.Structure):  # math/vector.rs:12-20
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double)]


class Mat3(C.Structure):  # math/mat.rs:10-18 (rows)
    _fields_ = [("x", Vec3), ("y", Vec3), ("z", Vec3)]


class CameraC(C.Structure):  # raytracing/camera.rs:7-15
    _fields_ = [("fov", C.c_double), ("position", Vec3), ("direction", Vec3),
                ("to_cam_space", Mat3), ("to_world_space", Mat3)]


class ConfigC(C.Structure):  # raytracing/scene.rs:16-28
    _fields_ = [("rays_per_pixel", C.c_uint64), ("max_bounces", C.c_uint64),
                ("focal_length", C.c_double), ("focal_offset", C.c_double), ("non_focal_offset", C.c_double)]


class MaterialC(C.Structure):  # raytracing/object.rs:79-89
    _fields_ = [("base_color", Vec3), ("emission_color", Vec3), ("roughness", C.c_double)]


class SphereC(C.Structure):  # object/sphere.rs:9-12
    _fields_ = [("position", Vec3), ("radius", C.c_double)]


class PlaneC(C.Structure):  # object/plane.rs:9-12
    _fields_ = [("position", Vec3), ("normal", Vec3)]


class TriangleC(C.Structure):  # object/triangle.rs:9-11
    _fields_ = [("vertices", Vec3 * 3)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("paths", C.c_uint64), ("pixels", C.c_uint64),
                ("triangle_panics", C.c_uint64), ("device_ms", C.c_float), ("total_ms", C.c_float),
                ("kernel_launches", C.c_uint32), ("accel_used", C.c_uint32),
                ("nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64), ("hits", C.c_uint64)]


class WireShapeBuffer(C.Structure):  # one per-type shape buffer of gpu_state.rs:53-77
    _fields_ = [("object_type", C.c_char_p), ("shape_bytes", C.c_void_p), ("shape_count", C.c_uint64)]


class Shard(C.Structure):
    _fields_ = [("index", C.c_uint32), ("count", C.c_uint32)]


assert C.sizeof(Vec3) == 24 and C.sizeof(Mat3) == 72 and C.sizeof(CameraC) == 200
assert C.sizeof(ConfigC) == 40 and C.sizeof(MaterialC) == 56
assert C.sizeof(SphereC) == 32 and C.sizeof(PlaneC) == 48 and C.sizeof(TriangleC) == 72

RTX_OK = 0
RTX_ERR_INVALID_ARGUMENT = 1
RTX_ERR_CUDA = 2
RTX_ERR_OUT_OF_MEMORY = 3
RTX_ERR_UNSUPPORTED_SHAPE = 4
RTX_ERR_INTERNAL = 5
RTX_ERR_REFERENCE_PANIC = 6
RTX_ACCEL_AUTO, RTX_ACCEL_BRUTE, RTX_ACCEL_LBVH = 0, 1, 2
RTX_FLAG_NO_RADIANCE = 4
RTX_FLAG_LANE_KERNEL = 8
RTX_FLAG_SHARD_IN_PLACE = 16
RTX_FLAG_PIXEL_QUEUE = 32
RTX_FLAG_SAMPLE_QUEUE = 64
RTX_FLAG_SHARD_ROWS = 128
RTX_FLAG_RNG_ROW = 256
RTX_PEER_HANDLE_BYTES = 64
RTX_TILE = 16
RTX_NOALLOC_SCRATCH_BYTES = 32 << 20
MISS_ID = 0xFFFFFFFF

_vp = C.c_void_p
_P = C.POINTER

# name -> (restype, argtypes); every symbol include/rtx_cuda.h declares
SIGNATURES = {
    "rtx_camera_new": (C.c_int32, [_P(Vec3), _P(Vec3), C.c_double, _P(CameraC)]),
    "rtx_camera_set_direction": (C.c_int32, [_P(CameraC), _P(Vec3)]),
    "rtx_camera_to_cam_space": (C.c_int32, [_P(CameraC), _P(Vec3), _P(Vec3)]),
    "rtx_camera_to_world_space": (C.c_int32, [_P(CameraC), _P(Vec3), _P(Vec3)]),
    "rtx_camera_rotate_to_world_space": (C.c_int32, [_P(CameraC), _P(Vec3), _P(Vec3)]),
    "rtx_config_default": (C.c_int32, [_P(ConfigC)]),
    "rtx_scene_create": (C.c_int32, [_P(CameraC), _P(ConfigC), C.c_int32, _P(_vp)]),
    "rtx_scene_destroy": (C.c_int32, [_vp]),
    "rtx_scene_set_camera": (C.c_int32, [_vp, _P(CameraC)]),
    "rtx_scene_set_config": (C.c_int32, [_vp, _P(ConfigC)]),
    "rtx_scene_get_camera": (C.c_int32, [_vp, _P(CameraC)]),
    "rtx_scene_get_config": (C.c_int32, [_vp, _P(ConfigC)]),
    "rtx_scene_add_spheres": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_add_planes": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_add_triangles": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_clear_objects": (C.c_int32, [_vp]),
    "rtx_scene_update_spheres": (C.c_int32, [_vp, _vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_update_planes": (C.c_int32, [_vp, _vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_update_triangles": (C.c_int32, [_vp, _vp, _vp, _vp, C.c_uint64]),
    "rtx_scene_rebuild_accel": (C.c_int32, [_vp]),
    "rtx_scene_accel_refits": (C.c_uint32, [_vp]),
    "rtx_scene_accel_builds": (C.c_uint32, [_vp]),
    "rtx_scene_last_launches": (C.c_uint32, [_vp]),
    "rtx_scene_object_count": (C.c_uint64, [_vp]),
    "rtx_scene_render": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _vp, _vp, _P(Stats)]),
    "rtx_scene_render_device": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _P(Shard),
                                            _vp, _vp, _vp, _vp]),
    "rtx_scene_render_device_band": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _P(Shard),
                                                 C.c_uint32, C.c_uint32, _vp, _vp, _vp, _vp]),
    "rtx_scene_render_device_noalloc": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _vp, _vp, _vp]),
    "rtx_scene_render_rows_to_host": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _P(Shard), _vp, _vp,
                                                  _vp, _vp]),
    "rtx_host_frame_create": (C.c_int32, [C.c_char_p, C.c_uint64, _P(_vp)]),
    "rtx_host_frame_open": (C.c_int32, [C.c_char_p, C.c_uint64, _P(_vp)]),
    "rtx_host_frame_close": (C.c_int32, [_vp, C.c_uint64, C.c_char_p]),
    "rtx_multi_create": (C.c_int32, [_P(CameraC), _P(ConfigC), _P(C.c_int32), C.c_uint32, _P(_vp)]),
    "rtx_multi_destroy": (C.c_int32, [_vp]),
    "rtx_multi_device_count": (C.c_uint32, [_vp]),
    "rtx_multi_scene": (_vp, [_vp, C.c_uint32]),
    "rtx_multi_set_camera": (C.c_int32, [_vp, _P(CameraC)]),
    "rtx_multi_set_config": (C.c_int32, [_vp, _P(ConfigC)]),
    "rtx_multi_add_spheres": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_multi_add_planes": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_multi_add_triangles": (C.c_int32, [_vp, _vp, _vp, C.c_uint64]),
    "rtx_multi_clear_objects": (C.c_int32, [_vp]),
    "rtx_multi_render": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _vp, _vp, _P(Stats)]),
    "rtx_shard_band_tiles": (None, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                    _P(C.c_uint32), _P(C.c_uint32)]),
    "rtx_shard_strip_count": (C.c_uint64, [C.c_uint32, C.c_uint32, C.c_uint32]),
    "rtx_shard_band_strips": (None, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P(C.c_uint32), _P(C.c_uint32)]),
    "rtx_shard_tile_count": (C.c_uint64, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]),
    "rtx_deinterleave_tiles": (C.c_int32, [C.c_uint32, C.c_uint32, C.c_uint32, _vp, _vp, _vp, _vp, _vp]),
    "rtx_scene_render_to_rgb8": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint32, _vp, _P(Stats)]),
    "rtx_scene_render_preview_f32": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, _vp, _vp, _P(Stats)]),
    "rtx_scene_render_preview_f32_device": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, _vp, _vp, _vp, _vp]),
    "rtx_quantize_rgb8_device": (C.c_int32, [_vp, C.c_uint32, C.c_uint32, _vp, _vp]),
    "rtx_last_error": (C.c_char_p, [_vp]),
    "rtx_abi_version": (C.c_uint32, []),
    "rtx_device_count": (C.c_int32, []),
    "rtx_host_alloc": (C.c_int32, [C.c_size_t, _P(_vp)]),
    "rtx_host_free": (C.c_int32, [_vp]),
    "rtx_wire_decode_camera": (C.c_int32, [_vp, _P(CameraC)]),
    "rtx_wire_decode_config": (C.c_int32, [_vp, _P(ConfigC)]),
    "rtx_wire_decode_objects": (C.c_int32, [_vp, C.c_uint64, _P(WireShapeBuffer), C.c_uint32, _vp, _vp, _vp]),
    "rtx_scene_add_wire_objects": (C.c_int32, [_vp, _vp, C.c_uint64, _P(WireShapeBuffer), C.c_uint32]),
    "rtx_scene_set_wire_camera": (C.c_int32, [_vp, _vp]),
    "rtx_scene_set_wire_config": (C.c_int32, [_vp, _vp]),
    "rtx_peer_alloc": (C.c_int32, [C.c_int32, C.c_uint64, _P(_vp)]),
    "rtx_peer_free": (C.c_int32, [C.c_int32, _vp]),
    "rtx_peer_export": (C.c_int32, [C.c_int32, _vp, _vp]),
    "rtx_peer_open": (C.c_int32, [C.c_int32, _vp, _P(_vp)]),
    "rtx_peer_close": (C.c_int32, [C.c_int32, _vp]),
    "rtx_probe_fp64_peak": (C.c_int32, [C.c_int32, _P(C.c_double), _P(C.c_double)]),
    "rtx_debug_closest_object": (C.c_int32, [_vp, C.c_uint32, _vp, C.c_uint64, _vp, _vp]),
    "rtx_debug_normal_at": (C.c_int32, [_vp, _vp, _vp, C.c_uint64, _vp]),
    "rtx_debug_random_bounce_dir": (C.c_int32, [C.c_int32, _vp, _vp, C.c_uint64, _vp]),
    "rtx_debug_wyrand_f64": (C.c_int32, [C.c_int32, _P(C.c_uint64), C.c_uint64, _vp]),
    "rtx_debug_vdiv3": (C.c_int32, [C.c_int32, _vp, _vp, C.c_uint64, _vp, _vp]),
    "rtx_debug_contraction_probe": (C.c_int32, [C.c_int32, C.c_double, C.c_double, C.c_double, _P(C.c_double)]),
    "rtx_stream_state": (C.c_uint64, [C.c_uint64, C.c_uint64]),
}

_lib = None


class RtxError(RuntimeError):
    """Non-zero rtx_status.  The Rust shim maps this to panic!, keeping the reference's contract."""

    def __init__(self, status, message):
        super().__init__(f"rtx status {status}: {message}")
        self.status = status


def load():
    """Loads librtx_cuda.so and declares every signature.  Raises if the library is not built."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.path.normpath(LIB_PATH)
    if not os.path.exists(path):
        raise ImportError(
            f"{path} not found: build the CUDA extension first (python -c 'import __graft_entry__ as g; g.build()'). "
            "The rtx CUDA path has no CPU fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status, scene=None):
    if status != RTX_OK:
        msg = load().rtx_last_error(scene)
        raise RtxError(status, msg.decode() if msg else "")
