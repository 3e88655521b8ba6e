"""rtx_b200 — host-side mirror of the `rtx` crate's scene/camera/render API over the CUDA C ABI.

Same names, argument meaning and error behaviour as the reference (paths relative to its src/):
  Vector3            math/vector.rs:12-20
  Camera             raytracing/camera.rs:7-68
  Config             raytracing/scene.rs:16-65
  Material           raytracing/object.rs:79-135
  Sphere/Plane/Triangle   raytracing/object/{sphere,plane,triangle}.rs
  Object             raytracing/object.rs:10-28
  Scene              raytracing/scene.rs:78-190  (new / add_object / render / render_to_image)

Everything that computes a pixel runs in librtx_cuda.so on the GPU; this file only marshals.  Failures
raise RtxError (the reference panics).
"""
import ctypes as C
import math

import numpy as np

from . import _abi
from ._abi import (MISS_ID, RTX_ACCEL_AUTO, RTX_ACCEL_BRUTE, RTX_ACCEL_LBVH, RTX_FLAG_LANE_KERNEL, RTX_FLAG_NO_RADIANCE,
                   RTX_FLAG_PIXEL_QUEUE, RTX_FLAG_SAMPLE_QUEUE, RTX_FLAG_SHARD_IN_PLACE, RTX_FLAG_SHARD_ROWS, RTX_FLAG_RNG_ROW, RTX_TILE, RtxError,
                   Shard, Stats)

__all__ = ["Vector3", "Camera", "Config", "Material", "Sphere", "Plane", "Triangle", "Object", "Scene", "RtxError",
           "HostFrame", "MultiScene",
           "MISS_ID", "RTX_ACCEL_AUTO", "RTX_ACCEL_BRUTE", "RTX_ACCEL_LBVH", "RTX_FLAG_LANE_KERNEL", "RTX_FLAG_NO_RADIANCE",
           "RTX_FLAG_PIXEL_QUEUE", "RTX_FLAG_SAMPLE_QUEUE", "RTX_FLAG_SHARD_IN_PLACE", "RTX_FLAG_SHARD_ROWS", "RTX_FLAG_RNG_ROW", "RTX_TILE",
           "device_count", "probe_fp64_peak", "shard_tile_count"]

SHAPE_SPHERE, SHAPE_PLANE, SHAPE_TRIANGLE = 0, 1, 2


def _v3(v):
    if isinstance(v, Vector3):
        return _abi.Vec3(v.x, v.y, v.z)
    x, y, z = v
    return _abi.Vec3(float(x), float(y), float(z))


class Vector3:
    """math/vector.rs:12-20 — three f64."""
    __slots__ = ("x", "y", "z")

    def __init__(self, x=0.0, y=0.0, z=0.0):
        self.x, self.y, self.z = float(x), float(y), float(z)

    new = classmethod(lambda cls, x, y, z: cls(x, y, z))
    zeros = classmethod(lambda cls: cls(0.0, 0.0, 0.0))
    ones = classmethod(lambda cls: cls(1.0, 1.0, 1.0))
    x_axis = classmethod(lambda cls: cls(1.0, 0.0, 0.0))
    y_axis = classmethod(lambda cls: cls(0.0, 1.0, 0.0))
    z_axis = classmethod(lambda cls: cls(0.0, 0.0, 1.0))

    @classmethod
    def of(cls, v):
        return v if isinstance(v, Vector3) else cls(*v)

    def __iter__(self):
        return iter((self.x, self.y, self.z))

    def __eq__(self, o):  # derive(PartialEq): -0.0 == 0.0
        o = Vector3.of(o)
        return self.x == o.x and self.y == o.y and self.z == o.z

    def __neg__(self):
        return Vector3(-self.x, -self.y, -self.z)

    def __repr__(self):
        return f"({self.x}, {self.y}, {self.z})"


class Camera:
    """raytracing/camera.rs:7-68.  `fov` is horizontal, radians.  The matrices are computed by the
    library (rtx_camera_new) with the reference's evaluation order."""

    def __init__(self, position, direction, fov):
        self._c = _abi.CameraC()
        p, d = _v3(position), _v3(direction)
        _abi.check(_abi.load().rtx_camera_new(C.byref(p), C.byref(d), float(fov), C.byref(self._c)))

    new = classmethod(lambda cls, position, direction, fov: cls(position, direction, fov))

    @property
    def fov(self):
        return self._c.fov

    @fov.setter
    def fov(self, v):
        self._c.fov = float(v)

    @property
    def position(self):
        p = self._c.position
        return Vector3(p.x, p.y, p.z)

    @position.setter
    def position(self, v):
        self._c.position = _v3(v)

    def get_direction(self):
        d = self._c.direction
        return Vector3(d.x, d.y, d.z)

    def set_direction(self, direction):
        """camera.rs:35-40 — keeps the upstream quirk: matrices are derived from the OLD direction."""
        d = _v3(direction)
        _abi.check(_abi.load().rtx_camera_set_direction(C.byref(self._c), C.byref(d)))

    def _xf(self, fn, vec):
        v, out = _v3(vec), _abi.Vec3()
        _abi.check(fn(C.byref(self._c), C.byref(v), C.byref(out)))
        return Vector3(out.x, out.y, out.z)

    def to_cam_space(self, vec):
        return self._xf(_abi.load().rtx_camera_to_cam_space, vec)

    def to_world_space(self, vec):
        return self._xf(_abi.load().rtx_camera_to_world_space, vec)

    def rotate_to_world_space(self, vec):
        return self._xf(_abi.load().rtx_camera_rotate_to_world_space, vec)

    def to_world_matrix(self):
        m = self._c.to_world_space
        return np.array([[r.x, r.y, r.z] for r in (m.x, m.y, m.z)], dtype=np.float64)


class Config:
    """raytracing/scene.rs:16-65."""

    def __init__(self, rays_per_pixel=16, max_bounces=10, focal_length=10.0, focal_offset=1e-4,
                 non_focal_offset=1e-1):
        self.rays_per_pixel = int(rays_per_pixel)
        self.max_bounces = int(max_bounces)
        self.focal_length = float(focal_length)
        self.focal_offset = float(focal_offset)
        self.non_focal_offset = float(non_focal_offset)

    default = classmethod(lambda cls: cls())

    def _with(self, **kw):
        c = Config(self.rays_per_pixel, self.max_bounces, self.focal_length, self.focal_offset,
                   self.non_focal_offset)
        for k, v in kw.items():
            setattr(c, k, v)
        return c

    def with_rays_per_pixel(self, v):
        return self._with(rays_per_pixel=int(v))

    def with_max_bounces(self, v):
        return self._with(max_bounces=int(v))

    def with_focal_length(self, v):
        return self._with(focal_length=float(v))

    def with_focal_offset(self, v):
        return self._with(focal_offset=float(v))

    def with_non_focal_offset(self, v):
        return self._with(non_focal_offset=float(v))

    def _c(self):
        return _abi.ConfigC(self.rays_per_pixel, self.max_bounces, self.focal_length, self.focal_offset,
                            self.non_focal_offset)


class Material:
    """raytracing/object.rs:79-135."""

    def __init__(self, base_color, emission_color, roughness):
        self.base_color = Vector3.of(base_color)
        self.emission_color = Vector3.of(emission_color)
        self.roughness = float(roughness)

    new = classmethod(lambda cls, b, e, r: cls(b, e, r))
    colored = classmethod(lambda cls, color: cls(color, (0, 0, 0), 1.0))       # object.rs:111-113
    light = classmethod(lambda cls, light_color: cls((0, 0, 0), light_color, 1.0))  # object.rs:130-132
    mirror = classmethod(lambda cls: cls((1, 1, 1), (0, 0, 0), 1.0))          # object.rs:133-135 (roughness 1.0 on the CPU path)

    def row(self):
        return [*self.base_color, *self.emission_color, self.roughness]


class Sphere:  # object/sphere.rs:9-17
    kind = SHAPE_SPHERE

    def __init__(self, position, radius):
        self.position, self.radius = Vector3.of(position), float(radius)

    new = classmethod(lambda cls, position, radius: cls(position, radius))

    def row(self):
        return [*self.position, self.radius]


class Plane:  # object/plane.rs:9-17
    kind = SHAPE_PLANE

    def __init__(self, position, normal):
        self.position, self.normal = Vector3.of(position), Vector3.of(normal)

    new = classmethod(lambda cls, position, normal: cls(position, normal))

    def row(self):
        return [*self.position, *self.normal]


class Triangle:  # object/triangle.rs:9-18
    kind = SHAPE_TRIANGLE

    def __init__(self, vertices):
        self.vertices = [Vector3.of(v) for v in vertices]
        assert len(self.vertices) == 3

    new = classmethod(lambda cls, vertices: cls(vertices))

    def row(self):
        return [c for v in self.vertices for c in v]


class Object:
    """raytracing/object.rs:10-28.  Only the three built-in shapes can cross the C ABI; a user-defined
    CustomShape (object.rs:53-76) raises RTX_ERR_UNSUPPORTED_SHAPE."""

    def __init__(self, shape, material):
        if not isinstance(shape, (Sphere, Plane, Triangle)):
            raise RtxError(_abi.RTX_ERR_UNSUPPORTED_SHAPE,
                           "user-defined CustomShape cannot cross the C ABI (only Sphere, Plane, Triangle)")
        self.shape, self.material = shape, material

    new = classmethod(lambda cls, shape, material: cls(shape, material))


_SHAPE_WIDTH = {SHAPE_SPHERE: 4, SHAPE_PLANE: 6, SHAPE_TRIANGLE: 9}


class Scene:
    """raytracing/scene.rs:78-190 over librtx_cuda.so.

    `objects` order is the object index (nearest-hit tie-break, scene.rs:250).  `seed` is the one added
    parameter: the reference's RNG is time-seeded and unreproducible.
    """

    def __init__(self, config=None, camera=None, device=0):
        self.config = config if config is not None else Config()
        # Scene::default (scene.rs:86-94): fov literal 90f64 is 90 RADIANS upstream; kept.
        self.camera = camera if camera is not None else Camera((0, 0, 0), (1, 0, 0), 90.0)
        self.device = int(device)
        self._batches = []  # (kind, shapes[n,k] f64, materials[n,7] f64) in insertion order
        self._handle = None
        self._synced_batches = 0
        self.last_stats = None

    new = classmethod(lambda cls, config, camera, device=0: cls(config, camera, device))

    # -- construction ---------------------------------------------------------------------------
    def add_object(self, obj):
        """scene.rs:126-128"""
        if not isinstance(obj, Object):
            raise RtxError(_abi.RTX_ERR_UNSUPPORTED_SHAPE, "add_object expects an rtx_b200.Object")
        self._append(obj.shape.kind, np.array([obj.shape.row()], dtype=np.float64),
                     np.array([obj.material.row()], dtype=np.float64))

    def add_spheres(self, centers, radii, materials):
        """Bulk add_object: centers[n,3], radii[n], materials[n,7] = base.xyz, emission.xyz, roughness."""
        centers = np.asarray(centers, dtype=np.float64).reshape(-1, 3)
        radii = np.asarray(radii, dtype=np.float64).reshape(-1, 1)
        self._append(SHAPE_SPHERE, np.hstack([centers, radii]), materials)

    def add_planes(self, positions, normals, materials):
        positions = np.asarray(positions, dtype=np.float64).reshape(-1, 3)
        normals = np.asarray(normals, dtype=np.float64).reshape(-1, 3)
        self._append(SHAPE_PLANE, np.hstack([positions, normals]), materials)

    def add_triangles(self, vertices, materials):
        """vertices[n,3,3] (or [n,9])."""
        self._append(SHAPE_TRIANGLE, np.asarray(vertices, dtype=np.float64).reshape(-1, 9), materials)

    def _append(self, kind, shapes, materials):
        shapes = np.ascontiguousarray(shapes, dtype=np.float64)
        materials = np.ascontiguousarray(np.asarray(materials, dtype=np.float64).reshape(-1, 7))
        if materials.shape[0] == 1 and shapes.shape[0] > 1:
            materials = np.ascontiguousarray(np.repeat(materials, shapes.shape[0], axis=0))
        assert shapes.shape == (materials.shape[0], _SHAPE_WIDTH[kind])
        if self._batches and self._batches[-1][0] == kind and len(self._batches) > self._synced_batches:
            k, s, m = self._batches[-1]
            self._batches[-1] = (k, np.vstack([s, shapes]), np.vstack([m, materials]))
        else:
            self._batches.append((kind, shapes, materials))

    def update_objects(self, indices, shapes, materials=None):
        """In-place replacement of existing objects (rtx_scene_update_*): `indices` = global object indices, all
        of one shape kind; `shapes` = rows in this module's layout (sphere: c.xyz r; plane: pos.xyz n.xyz;
        triangle: 9 doubles); `materials` = rows (base.xyz, emission.xyz, roughness) or None to keep them.
        Only the touched records are re-uploaded and the LBVH is refitted instead of rebuilt."""
        indices = np.ascontiguousarray(indices, dtype=np.uint64).reshape(-1)
        shapes = np.ascontiguousarray(shapes, dtype=np.float64).reshape(indices.shape[0], -1)
        if materials is not None:
            materials = np.ascontiguousarray(materials, dtype=np.float64).reshape(indices.shape[0], 7)
        # locate every object in the host-side batches
        starts = np.cumsum([0] + [b[1].shape[0] for b in self._batches])
        kinds = set()
        where = []
        for i in indices.tolist():
            if not 0 <= i < starts[-1]:
                raise RtxError(_abi.RTX_ERR_INVALID_ARGUMENT, "object index out of range")
            b = int(np.searchsorted(starts, i, side="right") - 1)
            kinds.add(self._batches[b][0])
            where.append((b, i - int(starts[b])))
        if len(kinds) > 1:
            raise RtxError(_abi.RTX_ERR_INVALID_ARGUMENT, "objects of different shape kinds in one update")
        if not where:
            return
        kind = kinds.pop()
        ncol = {SHAPE_SPHERE: 4, SHAPE_PLANE: 6, SHAPE_TRIANGLE: 9}[kind]
        if shapes.shape[1] != ncol:
            raise RtxError(_abi.RTX_ERR_INVALID_ARGUMENT, "shape rows have the wrong width for this kind")
        for (b, r), row, k in zip(where, shapes, range(len(where))):
            self._batches[b][1][r] = row
            if materials is not None:
                self._batches[b][2][r] = materials[k]
        if self._handle is not None and all(b < self._synced_batches for b, _ in where):
            fn = {SHAPE_SPHERE: "rtx_scene_update_spheres", SHAPE_PLANE: "rtx_scene_update_planes",
                  SHAPE_TRIANGLE: "rtx_scene_update_triangles"}[kind]
            _abi.check(getattr(_abi.load(), fn)(self._handle, indices.ctypes.data, shapes.ctypes.data,
                                                materials.ctypes.data if materials is not None else None,
                                                indices.shape[0]), self._handle)
        elif self._handle is not None:  # some of them were never uploaded: start over
            self.close()

    def rebuild_accel(self):
        if self._handle is not None:
            _abi.check(_abi.load().rtx_scene_rebuild_accel(self._handle), self._handle)

    @property
    def accel_refits(self):
        return int(_abi.load().rtx_scene_accel_refits(self._handle)) if self._handle is not None else 0

    @property
    def accel_builds(self):
        return int(_abi.load().rtx_scene_accel_builds(self._handle)) if self._handle is not None else 0

    def add_wire_objects(self, records, buffers):
        """≙ gpu::Scene::add_object's serialized state (gpu/gpu_state.rs:53-77): 48-byte object records plus one
        (object_type tag, bytes, count) shape buffer per shape_id, in the reference's f32 wire format (wire.py)."""
        from . import wire
        kinds, shapes, mats = wire.decode_objects(records, buffers)
        ncol = {SHAPE_SPHERE: 4, SHAPE_PLANE: 6, SHAPE_TRIANGLE: 9}
        for k, s, m in zip(kinds.tolist(), shapes, mats):
            self._append(int(k), s[None, :ncol[int(k)]].copy(), m[None, :].copy())

    @classmethod
    def from_wire(cls, camera_bytes, config_bytes, records, buffers, device=0):
        """A scene from the byte buffers a reference `gpu::Scene` holds (camera.rs:70-78, scene.rs:67-76, ...)."""
        from . import wire
        cc, cf = wire.decode_camera(camera_bytes), wire.decode_config(config_bytes)
        cam = Camera((cc.position.x, cc.position.y, cc.position.z), (cc.direction.x, cc.direction.y, cc.direction.z), cc.fov)
        cfg = Config(cf.rays_per_pixel, cf.max_bounces, cf.focal_length, cf.focal_offset, cf.non_focal_offset)
        sc = cls(cfg, cam, device)
        sc.add_wire_objects(records, buffers)
        return sc

    @property
    def object_count(self):
        return sum(b[1].shape[0] for b in self._batches)

    def objects_flat(self):
        """(types[n] int32, shapes[n,9] f64 zero-padded, materials[n,7]) in object-index order."""
        n = self.object_count
        types = np.zeros(n, dtype=np.int32)
        shapes = np.zeros((n, 9), dtype=np.float64)
        mats = np.zeros((n, 7), dtype=np.float64)
        o = 0
        for kind, s, m in self._batches:
            k = s.shape[0]
            types[o:o + k] = kind
            shapes[o:o + k, :s.shape[1]] = s
            mats[o:o + k] = m
            o += k
        return types, shapes, mats

    # -- native handle ----------------------------------------------------------------------------
    def _sync(self):
        lib = _abi.load()
        if self._handle is None:
            h = C.c_void_p()
            cfg = self.config._c()
            _abi.check(lib.rtx_scene_create(C.byref(self.camera._c), C.byref(cfg), self.device, C.byref(h)))
            self._handle = h
            self._synced_batches = 0
        else:
            cfg = self.config._c()
            _abi.check(lib.rtx_scene_set_camera(self._handle, C.byref(self.camera._c)), self._handle)
            _abi.check(lib.rtx_scene_set_config(self._handle, C.byref(cfg)), self._handle)
        add = {SHAPE_SPHERE: lib.rtx_scene_add_spheres, SHAPE_PLANE: lib.rtx_scene_add_planes,
               SHAPE_TRIANGLE: lib.rtx_scene_add_triangles}
        for kind, s, m in self._batches[self._synced_batches:]:
            _abi.check(add[kind](self._handle, s.ctypes.data, m.ctypes.data, s.shape[0]), self._handle)
        self._synced_batches = len(self._batches)
        return self._handle

    def close(self):
        if self._handle is not None:
            _abi.load().rtx_scene_destroy(self._handle)
            self._handle = None
            self._synced_batches = 0

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- render -----------------------------------------------------------------------------------
    def render(self, width, height, seed=0, flags=RTX_ACCEL_AUTO, want_ids=False, out=None, out_ids=None):
        """scene.rs:144-170 → float64 array [height][width][3] (row 0 = bottom), indexed img[y][x]."""
        h = self._sync()
        rgb = out if out is not None else np.empty((height, width, 3), dtype=np.float64)
        ids = out_ids
        if ids is None and want_ids:
            ids = np.empty((height, width), dtype=np.uint32)
        st = Stats()
        _abi.check(_abi.load().rtx_scene_render(h, width, height, seed, flags, rgb.ctypes.data,
                                                ids.ctypes.data if ids is not None else None, C.byref(st)), h)
        self.last_stats = st
        return (rgb, ids) if want_ids or out_ids is not None else rgb

    def render_ids(self, width, height, seed=0, flags=RTX_ACCEL_AUTO):
        """Hit-ID-only render: uint32 [height][width], MISS_ID where the primary ray hits nothing."""
        h = self._sync()
        ids = np.empty((height, width), dtype=np.uint32)
        st = Stats()
        _abi.check(_abi.load().rtx_scene_render(h, width, height, seed, flags | RTX_FLAG_NO_RADIANCE, None,
                                                ids.ctypes.data, C.byref(st)), h)
        self.last_stats = st
        return ids

    def render_to_image(self, width, height, seed=0, flags=RTX_ACCEL_AUTO):
        """scene.rs:173-178 → uint8 [height][width][3], row 0 = top (flipped), ×256 saturating, no gamma."""
        h = self._sync()
        img = np.empty((height, width, 3), dtype=np.uint8)
        st = Stats()
        _abi.check(_abi.load().rtx_scene_render_to_rgb8(h, width, height, seed, flags, img.ctypes.data, C.byref(st)), h)
        self.last_stats = st
        return img

    def render_preview(self, width, height, want_ids=False):
        """The f32 preview (the wgpu backend's arithmetic, base_shader.wgsl): float32 [height][width][3], row 0 = bottom;
        with want_ids also the object index sample 0's first ray hits."""
        h = self._sync()
        rgb = np.empty((height, width, 3), dtype=np.float32)
        ids = np.empty((height, width), dtype=np.uint32) if want_ids else None
        st = Stats()
        _abi.check(_abi.load().rtx_scene_render_preview_f32(h, width, height, rgb.ctypes.data,
                                                            ids.ctypes.data if ids is not None else None, C.byref(st)), h)
        self.last_stats = st
        return (rgb, ids) if want_ids else rgb

    def render_preview_device(self, width, height, d_rgb, d_ids, d_counters, stream):
        h = self._sync()
        _abi.check(_abi.load().rtx_scene_render_preview_f32_device(h, width, height, d_rgb, d_ids, d_counters, stream), h)

    def render_device(self, width, height, seed, flags, shard, d_rgb, d_ids, d_counters, stream):
        """Asynchronous device-buffer render (raw device pointers / cudaStream_t as ints)."""
        h = self._sync()
        sh = Shard(*shard) if shard is not None else None
        _abi.check(_abi.load().rtx_scene_render_device(h, width, height, seed, flags,
                                                       C.byref(sh) if sh is not None else None,
                                                       d_rgb, d_ids, d_counters, stream), h)

    def render_device_noalloc(self, width, height, seed, flags, d_rgb, d_ids, stream):
        """rtx_scene_render_device_noalloc: whole-image device render whose scratch is the library's static device
        array (no cudaMalloc/cudaFree in the call).  The scene must be resident and have been rendered once at this
        size through render_device; raises RtxError otherwise, or when the frame's scratch exceeds the array."""
        h = self._sync()
        _abi.check(_abi.load().rtx_scene_render_device_noalloc(h, width, height, seed, flags, d_rgb, d_ids, stream), h)

    def render_rows_to_host(self, width, height, seed, flags, shard, host_rgb_ptr, host_ids_ptr, d_counters, stream):
        """rtx_scene_render_rows_to_host: this shard's rows (strips of 4, RTX_FLAG_SHARD_ROWS) rendered in bands and
        delivered by this GPU's copy engine to their place in the whole-frame host buffer.  Asynchronous on `stream`."""
        h = self._sync()
        sh = Shard(*shard) if shard is not None else None
        _abi.check(_abi.load().rtx_scene_render_rows_to_host(h, width, height, seed, flags,
                                                             C.byref(sh) if sh is not None else None, host_rgb_ptr,
                                                             host_ids_ptr, d_counters, stream), h)

    def render_device_band(self, width, height, seed, flags, shard, tile_row_first, tile_row_count, d_rgb, d_ids,
                           d_counters, stream):
        """render_device for the band of 16-pixel tile rows [tile_row_first, tile_row_first + tile_row_count) only."""
        h = self._sync()
        sh = Shard(*shard) if shard is not None else None
        _abi.check(_abi.load().rtx_scene_render_device_band(h, width, height, seed, flags,
                                                            C.byref(sh) if sh is not None else None,
                                                            tile_row_first, tile_row_count, d_rgb, d_ids, d_counters,
                                                            stream), h)

    @property
    def last_device_launches(self):
        """Kernels launched by the most recent render_device call."""
        return int(_abi.load().rtx_scene_last_launches(self._handle)) if self._handle is not None else 0

    # -- per-function device evaluation (parity tests) ---------------------------------------------
    def debug_closest_object(self, rays, flags=RTX_ACCEL_BRUTE):
        h = self._sync()
        rays = np.ascontiguousarray(rays, dtype=np.float64).reshape(-1, 6)
        t = np.empty(rays.shape[0], dtype=np.float64)
        ids = np.empty(rays.shape[0], dtype=np.uint32)
        _abi.check(_abi.load().rtx_debug_closest_object(h, flags, rays.ctypes.data, rays.shape[0], t.ctypes.data,
                                                        ids.ctypes.data), h)
        return t, ids

    def debug_normal_at(self, ids, positions):
        h = self._sync()
        ids = np.ascontiguousarray(ids, dtype=np.uint32)
        positions = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
        out = np.empty_like(positions)
        _abi.check(_abi.load().rtx_debug_normal_at(h, ids.ctypes.data, positions.ctypes.data, ids.shape[0],
                                                   out.ctypes.data), h)
        return out


def device_count():
    return _abi.load().rtx_device_count()


def shard_tile_count(width, height, index, count):
    return _abi.load().rtx_shard_tile_count(width, height, index, count)


def probe_fp64_peak(device=0):
    """(non-fused DADD+DMUL Tflop/s, DFMA Tflop/s) measured on the device — the roofline denominators."""
    a, b = C.c_double(), C.c_double()
    _abi.check(_abi.load().rtx_probe_fp64_peak(device, C.byref(a), C.byref(b)))
    return a.value, b.value


def debug_wyrand_f64(state, n, device=0):
    st = C.c_uint64(state)
    out = np.empty(n, dtype=np.float64)
    _abi.check(_abi.load().rtx_debug_wyrand_f64(device, C.byref(st), n, out.ctypes.data))
    return out, st.value


def debug_random_bounce_dir(states, inputs, device=0):
    states = np.ascontiguousarray(states, dtype=np.uint64).copy()
    inputs = np.ascontiguousarray(inputs, dtype=np.float64).reshape(-1, 7)
    out = np.empty((inputs.shape[0], 3), dtype=np.float64)
    _abi.check(_abi.load().rtx_debug_random_bounce_dir(device, states.ctypes.data, inputs.ctypes.data,
                                                       inputs.shape[0], out.ctypes.data))
    return out, states


def debug_vdiv3(a, s, device=0):
    """(shared-reciprocal quotient, three plain IEEE divisions) of a[n][3] / s[n] on the device."""
    a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1, 3)
    s = np.ascontiguousarray(s, dtype=np.float64).reshape(-1)
    out, ref = np.empty_like(a), np.empty_like(a)
    _abi.check(_abi.load().rtx_debug_vdiv3(device, a.ctypes.data, s.ctypes.data, a.shape[0], out.ctypes.data,
                                           ref.ctypes.data))
    return out, ref


def debug_contraction_probe(a, b, c, device=0):
    out = C.c_double()
    _abi.check(_abi.load().rtx_debug_contraction_probe(device, a, b, c, C.byref(out)))
    return out.value


def stream_state(seed, key):
    return _abi.load().rtx_stream_state(seed, key)


class HostFrame:
    """One pinned host frame [height][width][3] f64 shared by the processes of a node (rtx_host_frame_*: POSIX shared
    memory registered with CUDA in every process).  create=True on one rank, False on the others (after the creator)."""

    def __init__(self, name, width, height, create):
        self.name, self.width, self.height, self.creator = name.encode(), width, height, bool(create)
        self.bytes = width * height * 24
        p = C.c_void_p()
        fn = _abi.load().rtx_host_frame_create if create else _abi.load().rtx_host_frame_open
        _abi.check(fn(self.name, self.bytes, C.byref(p)))
        self.ptr = int(p.value)
        self.array = np.ctypeslib.as_array((C.c_double * (width * height * 3)).from_address(self.ptr)).reshape(height, width, 3)

    def close(self):
        if self.ptr:
            self.array = None
            _abi.check(_abi.load().rtx_host_frame_close(C.c_void_p(self.ptr), self.bytes, self.name if self.creator else None))
            self.ptr = 0


class MultiScene:
    """rtx_multi_*: N GPUs of this process behind one handle; render() blocks and returns the whole image on the host
    (≙ Scene::render, scene.rs:144-170).  Objects are taken from a Scene built the usual way."""

    def __init__(self, scene, devices):
        lib = _abi.load()
        self.devices = list(devices)
        dev = (C.c_int32 * len(self.devices))(*self.devices)
        cfg = scene.config._c()
        h = C.c_void_p()
        _abi.check(lib.rtx_multi_create(C.byref(scene.camera._c), C.byref(cfg), dev, len(self.devices), C.byref(h)))
        self._handle = h
        add = {SHAPE_SPHERE: lib.rtx_multi_add_spheres, SHAPE_PLANE: lib.rtx_multi_add_planes,
               SHAPE_TRIANGLE: lib.rtx_multi_add_triangles}
        for kind, s, m in scene._batches:
            _abi.check(add[kind](h, s.ctypes.data, m.ctypes.data, s.shape[0]))
        self.last_stats = None

    def render(self, width, height, seed=0, flags=RTX_ACCEL_AUTO, out=None, want_ids=False):
        rgb = out if out is not None else np.empty((height, width, 3), dtype=np.float64)
        ids = np.empty((height, width), dtype=np.uint32) if want_ids else None
        st = Stats()
        _abi.check(_abi.load().rtx_multi_render(self._handle, width, height, seed, flags, rgb.ctypes.data,
                                                ids.ctypes.data if ids is not None else None, C.byref(st)))
        self.last_stats = st
        return (rgb, ids) if want_ids else rgb

    def close(self):
        if self._handle is not None:
            _abi.check(_abi.load().rtx_multi_destroy(self._handle))
            self._handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
