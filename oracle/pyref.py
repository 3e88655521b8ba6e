"""pyref.py — a second, independent restatement of the reference CPU path in pure Python (small cases only).

TEST INFRASTRUCTURE ONLY.  Python floats are IEEE-754 binary64 and CPython never fuses a*b+c; math.sqrt /
math.sin / math.cos call the same libm as the C++ oracle.  It exists so that two independent
transcriptions of the reference (this file, written from the Rust source in expression form, and
rtx_oracle.cpp) must agree bit for bit — the closest thing to a pin the oracle can have while the Rust
crate cannot be compiled here.  Citations are into the reference's src/.
"""
import math
import struct

M64 = (1 << 64) - 1
INF = float("inf")


# ---- math/vector.rs ---------------------------------------------------------------------------
def add(a, b): return (a[0] + b[0], a[1] + b[1], a[2] + b[2])
def sub(a, b): return (a[0] - b[0], a[1] - b[1], a[2] - b[2])
def muls(a, s): return (a[0] * s, a[1] * s, a[2] * s)
def mulv(a, b): return (a[0] * b[0], a[1] * b[1], a[2] * b[2])
def divs(a, s): return (fdiv(a[0], s), fdiv(a[1], s), fdiv(a[2], s))
def neg(a): return (-a[0], -a[1], -a[2])
def dot(a, b): return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]                      # vector.rs:85-87
def cross(a, b):                                                                  # vector.rs:89-95
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])
def length(a):                                                                    # vector.rs:101-103
    m = mulv(a, a)
    return fsqrt(m[0] + m[1] + m[2])
def norm(a): return divs(a, length(a))                                            # vector.rs:105-107


def fdiv(a, b):
    """IEEE division incl. the x/0 and 0/0 cases Python raises on."""
    try:
        return a / b
    except ZeroDivisionError:
        if a != a or a == 0.0:
            return float("nan")
        neg_res = (math.copysign(1.0, a) < 0) != (math.copysign(1.0, b) < 0)
        return -INF if neg_res else INF


def fsqrt(a):
    if a != a or a < 0.0:
        return float("nan")
    return math.sqrt(a)


# ---- fastrand 2.3.0 (WyRand) + the added seeding convention ------------------------------------
class WyRand:
    def __init__(self, s): self.s = s & M64

    def u64(self):
        self.s = (self.s + 0x2d358dccaa6c78a5) & M64
        t = self.s * (self.s ^ 0x8bb84b93962eacc9)
        return (t & M64) ^ (t >> 64)

    def f64(self):
        bits = 0x3FF0000000000000 + (self.u64() >> 12)
        return struct.unpack("<d", struct.pack("<Q", bits))[0] - 1.0


def mix64(z):
    z &= M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def stream_state(seed, key):
    return mix64(seed + 0x9E3779B97F4A7C15 * (key + 1))


# ---- camera.rs ------------------------------------------------------------------------------------
def mat_transpose(m): return tuple(tuple(m[r][c] for r in range(3)) for c in range(3))


def mat_det(m):                                                                    # specific_math.rs:23-43
    (a, b, c), (d, e, f), (g, h, i) = m
    return (a * e * i + b * f * g + c * d * h) - (g * e * c + h * f * a + i * d * b)


def mat_adj(m):                                                                    # specific_math.rs:45-71
    (a, b, c), (d, e, f), (g, h, i) = m
    return ((e * i - f * h, c * h - b * i, b * f - c * e),
            (f * g - d * i, a * i - c * g, c * d - a * f),
            (d * h - e * g, b * g - a * h, a * e - b * d))


def mat_inv(m):
    adj, det = mat_adj(m), mat_det(m)
    return tuple(divs(r, det) for r in adj)


def matvec(m, v): return (dot(v, m[0]), dot(v, m[1]), dot(v, m[2]))                # mat/mul.rs:42-51


def to_world_mat(direction):                                                       # camera.rs:42-49
    fwd = norm(direction)
    right = cross(fwd, (0., 0., -1.))
    up = cross(fwd, right)
    return mat_transpose((right, up, fwd))


# ---- shapes ---------------------------------------------------------------------------------------
def sphere_distance(sh, pos, d):                                                   # sphere.rs:19-30
    offset = sub(pos, sh[0:3])
    rd = norm(d)
    a = dot(rd, rd)
    b = 2.0 * dot(offset, rd)
    c = dot(offset, offset) - sh[3] * sh[3]
    disc = b * b - 4.0 * a * c
    if disc <= 1e-100:
        return None
    return fdiv(-b - fsqrt(disc), 2.0 * a)


def plane_distance(sh, pos, d):                                                    # plane.rs:20-31
    offset = sub(pos, sh[0:3])
    nrm_raw = tuple(sh[3:6])
    nrm = norm(nrm_raw)
    rd = norm(d)
    if dot(rd, nrm_raw) >= 0. or dot(offset, nrm_raw) <= 0.:
        return None
    t = fdiv(dot(offset, nrm), dot(rd, nrm))
    ip = add(offset, muls(rd, t))
    return length(sub(offset, ip))


def tri_normal(sh):                                                                # triangle.rs:104-107
    v0 = tuple(sh[0:3])
    return norm(cross(sub(sh[3:6], v0), sub(sh[6:9], v0)))


def tri_contains(sh, point):                                                       # triangle.rs:37-101
    v0 = tuple(sh[0:3])
    r, s, p = sub(sh[3:6], v0), sub(sh[6:9], v0), sub(point, v0)
    l1, l2, l3 = (r[0], s[0], p[0]), (r[1], s[1], p[1]), (r[2], s[2], p[2])
    if l1[0] == 0.0:
        if l2[0] == 0.0:
            if l3[0] == 0.0:
                return False
            l3, l1 = l1, l3
        else:
            l1, l2 = l2, l1
    l1 = divs(l1, l1[0])
    l2 = sub(l2, muls(l1, fdiv(l2[0], l1[0])))
    l3 = sub(l3, muls(l1, fdiv(l3[0], l1[0])))
    if not (l1[0] == 1.0 and l2[0] == 0.0 and l3[0] == 0.0):
        return False  # reference would panic
    if l2[1] == 0.0:
        if l3[1] == 0.0:
            return False
        l2, l3 = l3, l2
    l2 = divs(l2, l2[1])
    l1 = sub(l1, muls(l2, fdiv(l1[1], l2[1])))
    l3 = sub(l3, muls(l2, fdiv(l3[1], l2[1])))
    if not (l1[1] == 0.0 and l2[1] == 1.0 and l3[1] == 0.0):
        return False
    a, b = l1[2], l2[2]
    return 0. <= a <= 1. and 0. <= b <= 1. and (a + b) <= 1.


def tri_distance(sh, pos, d):                                                      # triangle.rs:108-127
    n = tri_normal(sh)
    v0 = tuple(sh[0:3])
    if dot(n, sub(v0, d)) < 0.0:
        return None
    rd = norm(d)
    if dot(rd, n) == 0.0:
        dist = INF
    else:
        dist = abs(fdiv(dot(n, sub(v0, pos)), dot(rd, n)))
    if dist == INF:
        return None
    if not tri_contains(sh, add(pos, muls(d, dist))):
        return None
    return dist


def obj_distance(o, pos, d):
    return (sphere_distance, plane_distance, tri_distance)[o["type"]](o["shape"], pos, d)


def obj_normal_at(o, p):                                                           # object.rs:37-39
    sh = o["shape"]
    if o["type"] == 0:
        n = norm(sub(p, sh[0:3]))
    elif o["type"] == 1:
        n = tuple(sh[3:6])
    else:
        n = tri_normal(sh)
    return norm(n)


def is_normal_positive(d):
    return d == d and d != INF and d >= 2.2250738585072014e-308


# ---- scene.rs -------------------------------------------------------------------------------------
def closest_object(objs, pos, d):                                                  # scene.rs:243-251
    best = None
    for i, o in enumerate(objs):
        t = obj_distance(o, pos, d)
        if t is None or not is_normal_positive(t):
            continue
        if best is None or t < best[0]:
            best = (t, i)
    return best


def random_direction(g):                                                           # vector.rs:36-45
    z = g.f64() * 2.0 - 1.0
    theta = g.f64() * 2.0 * math.pi
    r = fsqrt(1.0 - z * z)
    return norm((r * math.cos(theta), r * math.sin(theta), z))


def random_bounce_dir(g, rd, n, rough):                                            # scene.rs:279-291
    rnd = random_direction(g)
    refl = sub(rd, muls(muls(n, 2.0), dot(rd, n)))
    fin = norm(add(rnd, muls(sub(refl, rnd), 1.0 - rough)))
    return fin if dot(fin, n) > 0.0 else neg(fin)


def render(cam_pos, to_world, fov, cfg, objs, width, height, seed):
    """Returns (rgb[h][w] of 3-tuples, ids[h][w], total_rays).  objs: list of dict(type, shape, base, emission,
    roughness).  RNG mode = one stream per (pixel, sample)."""
    rpp, maxb = cfg["rays_per_pixel"], cfg["max_bounces"]
    vfov = height / width * fov                                                    # scene.rs:145
    img, ids, total = [], [], 0
    for y in range(height):
        row, idrow = [], []
        v = y / height
        for x in range(width):
            u = x / width
            ax, ay = fov * (u - 0.5), vfov * (v - 0.5)                             # scene.rs:214-215
            base_dir = matvec(to_world, (math.sin(ax), math.sin(ay), math.cos(ax) * math.cos(ay)))
            acc = (0., 0., 0.)
            first = 0xFFFFFFFF
            for s in range(rpp):
                g = WyRand(stream_state(seed, (y * width + x) * rpp + s))
                j0 = (g.f64(), g.f64(), g.f64())
                pos = add(cam_pos, muls(j0, cfg["non_focal_offset"]))
                focal = add(cam_pos, muls(base_dir, cfg["focal_length"]))
                j1 = (g.f64(), g.f64(), g.f64())
                target = add(focal, muls(j1, cfg["focal_offset"]))
                d = norm(sub(target, pos))
                light, res = (1., 1., 1.), (0., 0., 0.)
                if objs:
                    for seg in range(maxb + 1):                                    # scene.rs:227
                        if light[0] == 0.0 and light[1] == 0.0 and light[2] == 0.0:
                            break
                        hit = closest_object(objs, pos, d)
                        total += 1
                        if seg == 0 and s == 0:
                            first = hit[1] if hit else 0xFFFFFFFF
                        if hit is None:
                            break
                        t, i = hit
                        o = objs[i]
                        pos = add(pos, muls(d, t))
                        d = random_bounce_dir(g, d, obj_normal_at(o, pos), o["roughness"])
                        res = add(res, mulv(light, o["emission"]))
                        light = mulv(light, o["base"])
                acc = add(acc, res)
            row.append(divs(acc, float(rpp)))
            idrow.append(first)
        img.append(row)
        ids.append(idrow)
    return img, ids, total
