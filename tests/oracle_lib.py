"""ctypes bindings of the CPU oracle (oracle/libvxo.so) and, when present, of the reference itself
built for the host (oracle/_ref/*.so).  Test infrastructure: imported only by tests/, bench.py's
cpu_baseline / --impl reference legs and __graft_entry__.smoke()."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

NORM_HOST, NORM_NONE = 1, 2
F2U_CUDA, F2U_X86 = 0, 1

_u32p = np.ctypeslib.ndpointer(np.uint32, flags="C")
_f32p = np.ctypeslib.ndpointer(np.float32, flags="C")


class Hits(C.Structure):
    _fields_ = [("hit", C.c_void_p), ("pos", C.c_void_p), ("normal", C.c_void_p), ("steps", C.c_void_p),
                ("voxel", C.c_void_p), ("brick", C.c_void_p), ("coarse_steps", C.c_void_p),
                ("fine_steps", C.c_void_p), ("counters", C.c_void_p)]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("fwd", C.c_float * 3), ("up", C.c_float * 3),
                ("right", C.c_float * 3), ("fov_deg", C.c_float), ("tan_half", C.c_float),
                ("ortho_w", C.c_float), ("ortho_h", C.c_float)]


class Env(C.Structure):
    _fields_ = [("light_dir", C.c_float * 3), ("light_color", C.c_float * 3), ("ambient", C.c_float * 3)]


class RenderOpts(C.Structure):
    _fields_ = [("checkerboard", C.c_int), ("frame_number", C.c_uint32), ("debug_view", C.c_int),
                ("shadows", C.c_int), ("crosshair", C.c_int), ("norm_mode", C.c_int)]


def build_oracle():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "lib"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "libvxo.so")
        if not os.path.exists(path):
            build_oracle()
        L = C.CDLL(path)
        L.vxo_sample_index.restype = C.c_uint64
        L.vxo_sample_index.argtypes = [C.c_uint32] * 5
        L.vxo_hash.restype = C.c_uint32
        L.vxo_hash.argtypes = [C.c_uint32]
        L.vxo_perlin.restype = C.c_float
        L.vxo_perlin.argtypes = [C.c_float] * 4 + [C.c_int, C.c_int]
        L.vxo_terrain_noise.restype = C.c_float
        L.vxo_terrain_noise.argtypes = [C.c_float] * 3 + [C.c_int]
        L.vxo_voxel_solid.restype = C.c_int
        L.vxo_voxel_solid.argtypes = [C.c_uint32] * 3 + [C.c_int]
        L.vxo_populate_dense.restype = None
        L.vxo_populate_dense.argtypes = [_u32p] + [C.c_uint32] * 6 + [C.c_int, C.c_int]
        L.vxo_world_from_dense.restype = C.c_void_p
        L.vxo_world_from_dense.argtypes = [_u32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.vxo_world_from_noise.restype = C.c_void_p
        L.vxo_world_from_noise.argtypes = [C.c_uint32] * 3 + [C.c_int] * 3
        L.vxo_world_free.argtypes = [C.c_void_p]
        for name in ("dense_bits", "top_bits", "brick_bits", "bounds"):
            f = getattr(L, "vxo_world_" + name)
            f.restype = C.c_void_p
            f.argtypes = [C.c_void_p]
        L.vxo_ray_aabb.restype = C.c_int
        L.vxo_ray_aabb.argtypes = [_f32p] * 6
        L.vxo_trace.restype = None
        L.vxo_trace.argtypes = [C.c_void_p, _f32p, _f32p, C.c_size_t, C.c_int, C.c_int, C.POINTER(Hits), C.c_int]
        L.vxo_trace_dense.restype = None
        L.vxo_trace_dense.argtypes = [C.c_void_p, _f32p, _f32p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.vxo_get_directions.restype = None
        L.vxo_get_directions.argtypes = [_f32p] * 4
        L.vxo_render.restype = None
        L.vxo_render.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts),
                                 C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
        L.vxo_render_sample.restype = None
        L.vxo_render_sample.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), C.c_uint32,
                                        C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
        L.vxo_render_sample_planes.restype = None
        L.vxo_render_sample_planes.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), C.c_uint32,
                                               C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_int]
        L.vxo_frame_rays.restype = C.c_size_t
        L.vxo_frame_rays.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), C.c_uint32,
                                     C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_size_t]
        L.vxo_world_from_brickmap_view.restype = C.c_void_p
        L.vxo_world_from_brickmap_view.argtypes = [_u32p, _u32p, _f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        _lib = L
    return _lib


def _view(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype)


class OracleWorld:
    """Reference-form brickmap held by the oracle (dense bits, top bits, per-brick bits, bounds)."""

    def __init__(self, handle, X, Y, Z, factor):
        if not handle:
            raise ValueError("oracle rejected world %dx%dx%d factor %d" % (X, Y, Z, factor))
        self.h, self.X, self.Y, self.Z, self.factor = handle, X, Y, Z, factor
        self.cx, self.cy, self.cz = X // factor, Y // factor, Z // factor
        self.nbricks = self.cx * self.cy * self.cz
        self.words_per_brick = factor ** 3 // 32

    @classmethod
    def from_noise(cls, X, Y, Z, factor=8, f2u=F2U_CUDA, nthreads=0):
        return cls(lib().vxo_world_from_noise(X, Y, Z, factor, f2u, nthreads), X, Y, Z, factor)

    @classmethod
    def from_dense(cls, bits, X, Y, Z, factor=8):
        bits = np.ascontiguousarray(bits, np.uint32)
        return cls(lib().vxo_world_from_dense(bits, X, Y, Z, factor), X, Y, Z, factor)

    @classmethod
    def from_brickmap(cls, top, brick_bits, bounds, X, Y, Z, factor=8):
        """Non-owning view of a reference-form brickmap (e.g. vx.World.download()); no dense grid."""
        top = np.ascontiguousarray(top, np.uint32)
        brick_bits = np.ascontiguousarray(brick_bits, np.uint32)
        bounds = np.ascontiguousarray(bounds, np.float32).reshape(-1)
        self = cls(lib().vxo_world_from_brickmap_view(top, brick_bits, bounds, X, Y, Z, factor), X, Y, Z, factor)
        self._keep = (top, brick_bits, bounds)
        return self

    def render_sample(self, cam, env, opts, W, H, ty_begin, ty_end, ty_stride, nthreads=0):
        stats = np.zeros(8, np.uint64)
        lib().vxo_render_sample(self.h, C.byref(cam), C.byref(env), C.byref(opts), W, H, ty_begin, ty_end, ty_stride,
                                None, stats.ctypes.data, nthreads)
        return stats

    def render_sample_planes(self, cam, env, opts, W, H, ty_begin, ty_end, ty_stride, bgra, steps, nthreads=0):
        """The sampled thread rows' pixels into bgra (H x W x 4 uint8) and step counts into steps (H x W int32)."""
        assert bgra.dtype == np.uint8 and bgra.size == W * H * 4 and steps.dtype == np.int32 and steps.size == W * H
        lib().vxo_render_sample_planes(self.h, C.byref(cam), C.byref(env), C.byref(opts), W, H, ty_begin, ty_end, ty_stride,
                                       bgra.ctypes.data, steps.ctypes.data, nthreads)

    def frame_rays(self, cam, env, opts, W, H, ty_begin, ty_end, ty_stride):
        """(origins, dirs) screenDispatch hands to Raytrace on the sampled rows: primaries + one sun ray per primary hit."""
        rows = len(range(ty_begin, min(ty_end, H), max(ty_stride, 1)))
        cap = rows * ((W + 31) // 32 * 32) * 2
        o = np.zeros((cap, 3), np.float32)
        d = np.zeros((cap, 3), np.float32)
        n = lib().vxo_frame_rays(self.h, C.byref(cam), C.byref(env), C.byref(opts), W, H, ty_begin, ty_end, ty_stride,
                                 o.ctypes.data, d.ctypes.data, cap)
        assert n <= cap
        return o[:n], d[:n]

    def __del__(self):
        if getattr(self, "h", None):
            lib().vxo_world_free(self.h)
            self.h = None

    @property
    def dense(self):
        return _view(lib().vxo_world_dense_bits(self.h), (self.X * self.Y * self.Z + 31) // 32, np.uint32)

    @property
    def top(self):
        return _view(lib().vxo_world_top_bits(self.h), (self.nbricks + 31) // 32, np.uint32)

    @property
    def brick_bits(self):
        return _view(lib().vxo_world_brick_bits(self.h), self.nbricks * self.words_per_brick, np.uint32)

    @property
    def bounds(self):
        return _view(lib().vxo_world_bounds(self.h), self.nbricks * 6, np.float32).reshape(-1, 6)

    def trace(self, origins, dirs, norm_mode=NORM_HOST, max_steps=2048, nthreads=0):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        out = dict(hit=np.zeros(n, np.uint8), pos=np.zeros((n, 3), np.float32), normal=np.zeros((n, 3), np.float32),
                   steps=np.zeros(n, np.int32), voxel=np.zeros((n, 3), np.int32), brick=np.zeros(n, np.uint32),
                   coarse_steps=np.zeros(n, np.int32), fine_steps=np.zeros(n, np.int32),
                   counters=np.zeros((n, 3), np.uint32))
        h = Hits(**{k: v.ctypes.data for k, v in out.items()})
        lib().vxo_trace(self.h, o, d, n, norm_mode, max_steps, C.byref(h), nthreads)
        return out

    def trace_dense(self, origins, dirs, norm_mode=NORM_HOST, nthreads=0):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        hit, steps = np.zeros(n, np.uint8), np.zeros(n, np.int32)
        lib().vxo_trace_dense(self.h, o, d, n, norm_mode, hit.ctypes.data, steps.ctypes.data, nthreads)
        return hit, steps

    def render(self, cam, env, opts, W, H, bgra=None, nthreads=0):
        if bgra is None:
            bgra = np.zeros((H, W, 4), np.uint8)
        steps = np.zeros((H, W), np.int32)
        lib().vxo_render(self.h, C.byref(cam), C.byref(env), C.byref(opts), W, H, bgra.ctypes.data,
                         steps.ctypes.data, nthreads)
        return bgra, steps


def populate_dense(X, Y, Z, x0=0, y0=0, z0=0, f2u=F2U_CUDA, nthreads=0):
    bits = np.zeros((X * Y * Z + 31) // 32, np.uint32)
    lib().vxo_populate_dense(bits, X, Y, Z, x0, y0, z0, f2u, nthreads)
    return bits


def get_directions(euler):
    e = np.asarray(euler, np.float32)
    f, u, r = (np.zeros(3, np.float32) for _ in range(3))
    lib().vxo_get_directions(e, f, u, r)
    return f, u, r


def make_camera(origin, euler, fov_deg=90.0, tan_half=0.0, ortho=None):
    f, u, r = get_directions(euler)
    cam = Camera()
    cam.origin[:] = [float(v) for v in origin]
    cam.fwd[:] = f.tolist()
    cam.up[:] = u.tolist()
    cam.right[:] = r.tolist()
    cam.fov_deg = fov_deg
    cam.tan_half = tan_half
    cam.ortho_w, cam.ortho_h = (float(ortho[0]), float(ortho[1])) if ortho else (0.0, 0.0)  # SetOrthoWindowSize, ORTHO build
    return cam


def default_env():
    """VoxelApp/main.cu:58-62: L = normalize(1,1,1) (host), LightColor 2, Ambient 0.5."""
    e = Env()
    inv = np.float32(1.0 / np.sqrt(np.float64(np.float32(3.0))))
    e.light_dir[:] = [float(np.float32(1.0) * inv)] * 3
    e.light_color[:] = [2.0, 2.0, 2.0]
    e.ambient[:] = [0.5, 0.5, 0.5]
    return e


# ---------------------------------------------------------------------------------------------
# the reference itself, built for the host by oracle/Makefile (present only where it was built)
# ---------------------------------------------------------------------------------------------
def ref_host():
    path = os.path.join(ORACLE_DIR, "_ref", "libvxref_host.so")
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    L.vxref_build.restype = C.c_void_p
    L.vxref_build.argtypes = [_u32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
    L.vxref_export.argtypes = [C.c_void_p, _u32p, _u32p, _f32p]
    L.vxref_trace.argtypes = [C.c_void_p, _f32p, _f32p, C.c_size_t, C.c_int, C.c_void_p, _f32p, _f32p,
                              C.c_void_p, C.c_int]
    L.vxref_dda_dense.argtypes = [C.c_void_p, _f32p, _f32p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    L.vxref_ray_aabb.restype = C.c_int
    L.vxref_ray_aabb.argtypes = [_f32p] * 6
    L.vxref_sample_index.restype = C.c_uint32
    L.vxref_sample_index.argtypes = [C.c_uint32] * 5
    if hasattr(L, "vxref_from_brickmap"):
        L.vxref_from_brickmap.restype = C.c_void_p
        L.vxref_from_brickmap.argtypes = [_u32p, _u32p, _f32p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int]
        L.vxref_free_view.argtypes = [C.c_void_p]
        L.vxref_normalize.argtypes = [_f32p, _f32p]
        L.vxref_reflect.argtypes = [_f32p, _f32p, _f32p]
        L.vxref_dot.restype = C.c_float
        L.vxref_dot.argtypes = [_f32p, _f32p]
    return L


class RefWorld:
    """The reference's own march (GPUDDA::Raytrace compiled for the host, oracle/_ref/libvxref_host.so) on a brickmap in
    the reference's form that the caller keeps alive (e.g. vx.World.download())."""

    def __init__(self, top, brick_bits, bounds, X, Y, Z, factor=8):
        self.L = ref_host()
        if self.L is None or not hasattr(self.L, "vxref_from_brickmap"):
            raise RuntimeError("oracle/_ref/libvxref_host.so not built")
        self._keep = (np.ascontiguousarray(top, np.uint32), np.ascontiguousarray(brick_bits, np.uint32),
                      np.ascontiguousarray(bounds, np.float32).reshape(-1))
        self.h = self.L.vxref_from_brickmap(*self._keep, X, Y, Z, factor)

    def trace(self, origins, dirs, max_steps=2048, nthreads=0):
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        hit, steps = np.zeros(n, np.uint8), np.zeros(n, np.int32)
        pos, nrm = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        self.L.vxref_trace(self.h, o, d, n, max_steps, hit.ctypes.data, pos, nrm, steps.ctypes.data, nthreads)
        return dict(hit=hit, pos=pos, normal=nrm, steps=steps)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.vxref_free_view(self.h)
            self.h = None


def ref_noise():
    path = os.path.join(ORACLE_DIR, "_ref", "libvxref_noise.so")
    if not os.path.exists(path):
        return None
    L = C.CDLL(path)
    L.vxref_hash.restype = C.c_uint32
    L.vxref_hash.argtypes = [C.c_uint32]
    L.vxref_random_float.restype = C.c_float
    L.vxref_random_float.argtypes = [C.c_uint32]
    L.vxref_perlin.restype = C.c_float
    L.vxref_perlin.argtypes = [C.c_float] * 4 + [C.c_int]
    L.vxref_terrain_noise.restype = C.c_float
    L.vxref_terrain_noise.argtypes = [C.c_float] * 3
    L.vxref_fbm.restype = C.c_float
    L.vxref_fbm.argtypes = [C.c_float] * 3 + [C.c_int]
    return L
