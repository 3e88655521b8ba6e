"""The reference's f32 wire format (trait GpuSerialize) — serializers that mirror the reference's own, and the
decoders of the C ABI (rtx_wire_*; rust-raytracing_b200/csrc/rtx_wire.cu).

Reference layouts (all little-endian f32 unless noted):
  f64.serialize()         `(x as f32).to_le_bytes()`                           raytracing/gpu/serialize.rs:8-12
  Vector3.serialize()     x, y, z                                              (12 B)
  Camera                  position, pad(4), direction, fov                     raytracing/camera.rs:70-78   (32 B)
  Config                  rays_per_pixel u32, max_bounces u32, 3 x f32         raytracing/scene.rs:67-76    (20 B)
  Object record           base_color, roughness, emission, shape_id u32,
                          index u32, pad(12)                                   raytracing/gpu/object.rs:34-42 (48 B)
  Sphere / Plane / Triangle   object/sphere.rs:36-42 (16 B), plane.rs:38-46 (32 B), triangle.rs:130-136 (48 B)
gpu_state.rs:53-77 numbers shape types in order of first appearance and keeps one byte buffer per type.
"""
import ctypes as C
import struct

import numpy as np

from . import SHAPE_PLANE, SHAPE_SPHERE, SHAPE_TRIANGLE, _abi

TAGS = {SHAPE_SPHERE: "rtx::raytracing::object::sphere::sphere",   # format!("{}::sphere", module_path!())
        SHAPE_PLANE: "rtx::raytracing::object::plane::plane",
        SHAPE_TRIANGLE: "rtx::raytracing::object::triangle::triangle"}


def _f32(*vals):
    return struct.pack("<%df" % len(vals), *[float(np.float32(v)) for v in vals])


def serialize_camera(position, direction, fov):
    return _f32(*position) + b"\0" * 4 + _f32(*direction) + _f32(fov)


def serialize_config(rays_per_pixel, max_bounces, focal_length, focal_offset, non_focal_offset):
    return struct.pack("<II", rays_per_pixel & 0xFFFFFFFF, max_bounces & 0xFFFFFFFF) + \
        _f32(focal_length, focal_offset, non_focal_offset)


def serialize_shape(kind, row):
    """row: the f64 shape row of rtx_b200 (sphere: c.xyz r; plane: pos.xyz n.xyz; triangle: 9 doubles)."""
    if kind == SHAPE_SPHERE:
        return _f32(*row[:3]) + _f32(row[3])
    if kind == SHAPE_PLANE:
        return _f32(*row[:3]) + b"\0" * 4 + _f32(*row[3:6]) + b"\0" * 4
    return b"".join(_f32(*row[3 * v:3 * v + 3]) + b"\0" * 4 for v in range(3))


def serialize_object_record(material_row, shape_id, index):
    """material_row: base.xyz, emission.xyz, roughness (the rtx_b200 material row)."""
    return _f32(*material_row[:3]) + _f32(material_row[6]) + _f32(*material_row[3:6]) + \
        struct.pack("<II", shape_id, index) + b"\0" * 12


def serialize_scene(kinds, shapes, materials):
    """What gpu::Scene::add_object accumulates for objects given in insertion order:
    (object_records bytes, [(object_type tag, shape buffer bytes, count)] indexed by shape_id)."""
    order, buffers, counts, records = {}, [], [], []
    for kind, row, mat in zip(kinds, shapes, materials):
        kind = int(kind)
        if kind not in order:
            order[kind] = len(buffers)
            buffers.append(bytearray())
            counts.append(0)
        sid = order[kind]
        records.append(serialize_object_record(mat, sid, counts[sid]))
        buffers[sid] += serialize_shape(kind, row)
        counts[sid] += 1
    by_id = sorted(order.items(), key=lambda kv: kv[1])
    return b"".join(records), [(TAGS[k], bytes(buffers[i]), counts[i]) for k, i in by_id]


def _shape_buffers(buffers):
    keep = []
    arr = (_abi.WireShapeBuffer * max(len(buffers), 1))()
    for i, (tag, data, count) in enumerate(buffers):
        raw = C.create_string_buffer(bytes(data), len(data))
        keep.append(raw)
        arr[i].object_type = tag.encode() if isinstance(tag, str) else tag
        arr[i].shape_bytes = C.cast(raw, C.c_void_p)
        arr[i].shape_count = count
    return arr, keep


def decode_camera(data):
    """-> _abi.CameraC (Camera::new on the widened values, camera.rs:19-28).  Pure host code."""
    out = _abi.CameraC()
    _abi.check(_abi.load().rtx_wire_decode_camera(bytes(data), C.byref(out)))
    return out


def decode_config(data):
    out = _abi.ConfigC()
    _abi.check(_abi.load().rtx_wire_decode_config(bytes(data), C.byref(out)))
    return out


def decode_objects(records, buffers):
    """-> (kinds[n] u32, shapes[n][9] f64, materials[n][7] f64 as base, emission, roughness).  Pure host code."""
    n = len(records) // 48
    kinds = np.zeros(n, dtype=np.uint32)
    shapes = np.zeros((n, 9), dtype=np.float64)
    mats = np.zeros((n, 7), dtype=np.float64)
    arr, keep = _shape_buffers(buffers)
    _abi.check(_abi.load().rtx_wire_decode_objects(bytes(records), n, arr, len(buffers), kinds.ctypes.data,
                                                   shapes.ctypes.data, mats.ctypes.data))
    del keep
    return kinds, shapes, mats
