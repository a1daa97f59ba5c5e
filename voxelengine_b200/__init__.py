"""voxelengine_b200 — B200-native brickmap voxel ray tracer (host-side Python binding).

Thin ctypes layer over the C ABI in include/vxrt.h (libvxrt.so, hand-written CUDA for sm_100a).
The classes mirror the reference's objects for this path (JoshuaLim007/VoxelEngine, VoxelRT/):

    World                <- CreateVoxels + GenerateLowresVoxelBuffer + Upload*
                            (VoxelWorldBuilder.cuh:12-32, VolumeRaytracer.cuh:379-516, VolumeRaytracer.cu:527-572)
    VoxelRaytracer3D     <- GPUDDA::VoxelRaytracer3D                 (VolumeRaytracer.cuh:291-377)
    render_screen        <- GPUDDA::Graphics::RenderScreen           (Renderer.cu:305-328)
    get_directions       <- GPUDDA::Graphics::GetDirections          (Renderer.cu:27-42)

There is no CPU path here: if libvxrt.so is missing or no CUDA device is present, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VXRT_LIB") or os.path.join(_HERE, "lib", "libvxrt.so")  # VXRT_LIB: tuning builds (tools/)

NORM_DEVICE, NORM_HOST, NORM_NONE = 0, 1, 2
STATIC_MAX_RAY_IDS = 262144  # VX_STATIC_MAX_RAY_IDS of include/vxrt.h (World.render_static)
MAX_STEPS = 2048


class VxError(RuntimeError):
    pass


class WorldInfo(C.Structure):
    _fields_ = [("dims", C.c_uint32 * 3), ("coarse", C.c_uint32 * 3), ("factor", C.c_int32), ("device", C.c_int32),
                ("bricks", C.c_uint64), ("occupied_bricks", C.c_uint64), ("solid_voxels", C.c_uint64),
                ("bytes_top", C.c_uint64), ("bytes_records", C.c_uint64), ("bytes_pool", C.c_uint64),
                ("build_ms", C.c_float), ("tiles_ms", C.c_float)]


class SlabInfo(C.Structure):
    _fields_ = [("brick_begin", C.c_uint64), ("brick_count", C.c_uint64), ("occupied_bricks", C.c_uint64), ("solid_voxels", C.c_uint64),
                ("words_per_brick", C.c_uint32), ("build_ms", C.c_float)]


class Hits(C.Structure):
    _fields_ = [("hit", C.c_void_p), ("pos", C.c_void_p), ("normal", C.c_void_p), ("steps", C.c_void_p),
                ("voxel", C.c_void_p), ("brick", C.c_void_p), ("coarse_steps", C.c_void_p),
                ("fine_steps", C.c_void_p), ("counters", C.c_void_p)]


class TraceOpts(C.Structure):
    _fields_ = [("norm_mode", C.c_int32), ("max_steps", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("fwd", C.c_float * 3), ("up", C.c_float * 3), ("right", C.c_float * 3),
                ("fov_deg", C.c_float), ("tan_half", C.c_float),
                ("ortho_w", C.c_float), ("ortho_h", C.c_float)]


class Env(C.Structure):
    _fields_ = [("light_dir", C.c_float * 3), ("light_color", C.c_float * 3), ("ambient", C.c_float * 3)]


class RenderOpts(C.Structure):
    _fields_ = [("checkerboard", C.c_int32), ("frame_number", C.c_uint32), ("debug_view", C.c_int32),
                ("shadows", C.c_int32), ("crosshair", C.c_int32), ("norm_mode", C.c_int32),
                ("tile_x0", C.c_uint32), ("tile_y0", C.c_uint32), ("tile_w", C.c_uint32), ("tile_h", C.c_uint32),
                ("d_stats", C.c_void_p), ("band_count", C.c_uint32), ("band_index", C.c_uint32)]


# Every symbol include/vxrt.h declares (tests check that the library exports all of them).
ABI_SYMBOLS = [
    "vx_last_error", "vx_version", "vx_device_count", "vx_world_build_noise", "vx_world_build_noise_full",
    "vx_world_from_dense_bits", "vx_world_destroy", "vx_world_get_info", "vx_world_download", "vx_world_download_bricks", "vx_tracer_create",
    "vx_tracer_destroy", "vx_trace", "vx_trace_device", "vx_hits_postprocess", "vx_get_directions", "vx_render",
    "vx_render_host", "vx_render_static", "vx_bounce_rays", "vx_tracer_set_timing", "vx_tracer_set_fused_shade", "vx_tracer_frame_times", "vx_device_tanf", "vx_device_normalize", "vx_measure_read_bandwidth",
    "vx_shared_alloc", "vx_shared_open", "vx_shared_close", "vx_shared_free", "vx_device_to_host", "vx_device_fill",
    "vx_host_register", "vx_host_unregister", "vx_bands_to_host",
    "vx_world_edit", "vx_world_save", "vx_world_load",
    "vx_slab_build_noise", "vx_slab_get_info", "vx_slab_export", "vx_slab_destroy", "vx_world_from_parts",
]

_lib = None


def lib():
    """Load libvxrt.so (built by `make` / __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VxError("libvxrt.so not built (%s): run `make` at the repo root; there is no CPU fallback" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, u32, i32, sz = C.c_void_p, C.c_uint32, C.c_int, C.c_size_t
    L.vx_last_error.restype = C.c_char_p
    L.vx_version.restype = i32
    L.vx_device_count.argtypes = [C.POINTER(i32)]
    for name in ("vx_world_build_noise", "vx_world_build_noise_full"):
        getattr(L, name).argtypes = [C.POINTER(vp), u32, u32, u32, i32, u32, u32, u32, i32]
    L.vx_world_from_dense_bits.argtypes = [C.POINTER(vp), vp, u32, u32, u32, i32, i32]
    L.vx_world_destroy.argtypes = [vp]
    L.vx_world_get_info.argtypes = [vp, C.POINTER(WorldInfo)]
    L.vx_world_download.argtypes = [vp, vp, vp, vp]
    L.vx_world_download_bricks.argtypes = [vp, vp, sz, vp, vp, vp]
    L.vx_tracer_create.argtypes = [C.POINTER(vp), vp, sz]
    L.vx_tracer_destroy.argtypes = [vp]
    L.vx_trace.argtypes = [vp, vp, vp, sz, C.POINTER(TraceOpts), C.POINTER(Hits)]
    L.vx_trace_device.argtypes = [vp, vp, vp, sz, C.POINTER(TraceOpts), C.POINTER(Hits), vp]
    L.vx_hits_postprocess.argtypes = [vp, vp, sz, vp, vp]
    L.vx_get_directions.argtypes = [vp, vp, vp, vp]
    L.vx_get_directions.restype = None
    L.vx_render.argtypes = [vp, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), vp, vp, u32, u32, vp]
    L.vx_render_static.argtypes = [vp, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), vp, vp, u32, u32, vp]
    L.vx_render_host.argtypes = [vp, C.POINTER(Camera), C.POINTER(Env), C.POINTER(RenderOpts), vp, u32, u32]
    L.vx_bounce_rays.argtypes = [vp, C.POINTER(Camera), C.POINTER(RenderOpts), u32, u32, vp, vp, vp, vp, vp]
    L.vx_tracer_set_timing.argtypes = [vp, i32]
    L.vx_tracer_frame_times.argtypes = [vp, C.POINTER(C.c_float * 4)]
    L.vx_device_tanf.argtypes = [C.c_float, C.POINTER(C.c_float), i32]
    L.vx_device_normalize.argtypes = [vp, vp, sz, i32]
    L.vx_measure_read_bandwidth.argtypes = [i32, sz, i32, C.POINTER(C.c_float)]
    L.vx_tracer_set_fused_shade.argtypes = [vp, i32]
    L.vx_shared_alloc.argtypes = [sz, i32, C.POINTER(vp), C.c_char_p]
    L.vx_shared_open.argtypes = [C.c_char_p, i32, C.POINTER(vp)]
    L.vx_shared_close.argtypes = [vp, i32]
    L.vx_shared_free.argtypes = [vp, i32]
    L.vx_device_to_host.argtypes = [vp, vp, sz, i32, vp]
    L.vx_device_fill.argtypes = [vp, i32, sz, i32, vp]
    L.vx_world_edit.argtypes = [vp, vp, sz, i32]
    L.vx_world_save.argtypes = [vp, C.c_char_p]
    L.vx_world_load.argtypes = [C.POINTER(vp), C.c_char_p, i32]
    L.vx_slab_build_noise.argtypes = [C.POINTER(vp), u32, u32, u32, i32, u32, u32, u32, u32, u32, i32]
    L.vx_slab_get_info.argtypes = [vp, C.POINTER(SlabInfo)]
    L.vx_slab_export.argtypes = [vp, vp, vp, vp, vp]
    L.vx_slab_destroy.argtypes = [vp]
    L.vx_world_from_parts.argtypes = [C.POINTER(vp), u32, u32, u32, i32, vp, vp, vp, C.c_uint64, C.c_uint64, i32]
    L.vx_host_register.argtypes = [vp, sz]
    L.vx_host_unregister.argtypes = [vp]
    L.vx_bands_to_host.argtypes = [vp, vp, u32, u32, u32, u32, i32, vp]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise VxError("vxrt error %d: %s" % (rc, lib().vx_last_error().decode()))


def device_count():
    n = C.c_int(0)
    _check(lib().vx_device_count(C.byref(n)))
    return n.value


def get_directions(euler):
    """Graphics::GetDirections (Renderer.cu:27-42): returns (forward, up, right) float32 vectors."""
    e = np.ascontiguousarray(euler, np.float32)
    f, u, r = (np.zeros(3, np.float32) for _ in range(3))
    lib().vx_get_directions(e.ctypes.data, f.ctypes.data, u.ctypes.data, r.ctypes.data)
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
    """The app's environment, VoxelApp/main.cu:58-62."""
    e = Env()
    inv = np.float32(1.0 / np.sqrt(np.float64(np.float32(3.0))))
    e.light_dir[:] = [float(np.float32(1.0) * inv)] * 3
    e.light_color[:] = [2.0, 2.0, 2.0]
    e.ambient[:] = [0.5, 0.5, 0.5]
    return e


def device_tanf(x, device=0):
    out = C.c_float(0)
    _check(lib().vx_device_tanf(float(x), C.byref(out), device))
    return out.value


def measure_read_bandwidth(nbytes, reps=20, device=0):
    """GB/s of an all-SM streaming read of an nbytes buffer (L2-resident when nbytes << 126 MB)."""
    out = C.c_float(0)
    _check(lib().vx_measure_read_bandwidth(device, int(nbytes), int(reps), C.byref(out)))
    return out.value


SHARED_HANDLE_BYTES = 64


def shared_alloc(nbytes, device=0):
    """(device pointer, 64-byte handle) of a zeroed buffer other processes of this node can map (vx_shared_alloc)."""
    ptr, handle = C.c_void_p(), C.create_string_buffer(SHARED_HANDLE_BYTES)
    _check(lib().vx_shared_alloc(int(nbytes), device, C.byref(ptr), handle))
    return ptr.value, handle.raw


def shared_open(handle, device=0):
    """Map another process's vx_shared_alloc buffer; the pointer is valid in this process's kernels (peer memory)."""
    if len(handle) != SHARED_HANDLE_BYTES:
        raise ValueError("handle must be %d bytes" % SHARED_HANDLE_BYTES)
    ptr = C.c_void_p()
    _check(lib().vx_shared_open(bytes(handle), device, C.byref(ptr)))
    return ptr.value


def shared_close(ptr, device=0):
    _check(lib().vx_shared_close(ptr, device))


def shared_free(ptr, device=0):
    _check(lib().vx_shared_free(ptr, device))


def device_to_host(host_ptr, dev_ptr, nbytes, device=0, stream=0):
    _check(lib().vx_device_to_host(host_ptr, dev_ptr, int(nbytes), device, stream))


def device_fill(dev_ptr, byte, nbytes, device=0, stream=0):
    _check(lib().vx_device_fill(dev_ptr, int(byte), int(nbytes), device, stream))


def host_register(ptr, nbytes):
    _check(lib().vx_host_register(ptr, int(nbytes)))


def host_unregister(ptr):
    _check(lib().vx_host_unregister(ptr))


def bands_to_host(host_ptr, dev_ptr, W, H, band_count=0, band_index=0, device=0, stream=0):
    """Enqueue the copy of this rank's 8-row bands of a W x H BGRA frame from its device frame to the host frame."""
    _check(lib().vx_bands_to_host(host_ptr, dev_ptr, W, H, band_count, band_index, device, stream))


def device_normalize(dirs, device=0):
    d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    out = np.empty_like(d)
    _check(lib().vx_device_normalize(d.ctypes.data, out.ctypes.data, d.shape[0], device))
    return out


class World:
    """Brickmap resident on one GPU."""

    def __init__(self, handle):
        self.h = handle

    @classmethod
    def from_noise(cls, X, Y, Z, factor=8, origin=(0, 0, 0), device=0, full_octaves=False):
        h = C.c_void_p()
        fn = lib().vx_world_build_noise_full if full_octaves else lib().vx_world_build_noise
        _check(fn(C.byref(h), X, Y, Z, factor, origin[0], origin[1], origin[2], device))
        return cls(h)

    @classmethod
    def from_dense_bits(cls, bits, X, Y, Z, factor=8, device=0):
        b = np.ascontiguousarray(bits, np.uint32)
        if b.size != (X * Y * Z + 31) // 32:
            raise VxError("dense bit grid has %d words, expected %d" % (b.size, (X * Y * Z + 31) // 32))
        h = C.c_void_p()
        _check(lib().vx_world_from_dense_bits(C.byref(h), b.ctypes.data, X, Y, Z, factor, device))
        return cls(h)

    @classmethod
    def from_parts(cls, X, Y, Z, factor, d_top_words, d_bounds, d_pool_words, occupied_bricks, solid_voxels, device=0):
        """vx_world_from_parts: the world whose top-level bits, packed bounds and compact mask pool (device pointers, whole
        world, brick order) were built slab by slab (Slab) and brought together by the caller."""
        h = C.c_void_p()
        _check(lib().vx_world_from_parts(C.byref(h), X, Y, Z, factor, int(d_top_words), int(d_bounds), int(d_pool_words) or None,
                                         int(occupied_bricks), int(solid_voxels), device))
        return cls(h)

    @classmethod
    def load(cls, path, device=0):
        """A world written by save()."""
        h = C.c_void_p()
        _check(lib().vx_world_load(C.byref(h), os.fsencode(path), device))
        return cls(h)

    def save(self, path):
        _check(lib().vx_world_save(self.h, os.fsencode(path)))

    def render_static(self, cam, env, opts, d_bgra, W, H, d_steps=0, stream=0):
        """vx_render_static: VoxelRaytracer3D.render_device without a tracer and without any allocation (the frame's scratch is
        a fixed device array of the library; frames up to STATIC_MAX_RAY_IDS ray ids, e.g. 640x360; cam.tan_half must be set;
        calls that overlap on the GPU race).  No counterpart in the reference."""
        _check(lib().vx_render_static(self.h, C.byref(cam), C.byref(env), C.byref(opts), int(d_bgra), int(d_steps) or None, W, H,
                                      int(stream) or None))

    def edit(self, voxels, solid=True):
        """Set (or clear) the voxels (n x 3 integer coordinates) in place: masks, bounds and top-level bits of the touched
        bricks are refreshed on the device; the result equals a rebuild from the edited dense grid."""
        v = np.ascontiguousarray(voxels, np.uint32).reshape(-1, 3)
        _check(lib().vx_world_edit(self.h, v.ctypes.data, v.shape[0], 1 if solid else 0))

    def close(self):
        if self.h:
            lib().vx_world_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def info(self):
        i = WorldInfo()
        _check(lib().vx_world_get_info(self.h, C.byref(i)))
        return i

    def download(self):
        """(top_words, brick_words, bounds[n,6]) in the reference's GenerateLowresVoxelBuffer form."""
        i = self.info
        nb = int(i.bricks)
        wpb = i.factor ** 3 // 32
        top = np.zeros((nb + 31) // 32, np.uint32)
        bw = np.zeros(nb * wpb, np.uint32)
        b6 = np.zeros((nb, 6), np.float32)
        _check(lib().vx_world_download(self.h, top.ctypes.data, bw.ctypes.data, b6.ctypes.data))
        return top, bw, b6

    def download_top(self):
        """The top-level occupancy bits alone (reference tiled index)."""
        top = np.zeros((int(self.info.bricks) + 31) // 32, np.uint32)
        _check(lib().vx_world_download(self.h, top.ctypes.data, None, None))
        return top

    def download_bricks(self, coarse_xyz):
        """(occupied[n], brick_words[n, factor^3/32], bounds[n, 6]) of the bricks at the given coarse coordinates, in the
        reference's form; occupied bit 0 = traversal grid's top-level bit, bit 1 = the reference-indexed one."""
        c = np.ascontiguousarray(coarse_xyz, np.uint32).reshape(-1, 3)
        n = c.shape[0]
        wpb = self.info.factor ** 3 // 32
        occ = np.zeros(n, np.uint8)
        bw = np.zeros((n, wpb), np.uint32)
        b6 = np.zeros((n, 6), np.float32)
        _check(lib().vx_world_download_bricks(self.h, c.ctypes.data, n, occ.ctypes.data, bw.ctypes.data, b6.ctypes.data))
        return occ, bw, b6


class Slab:
    """One builder's z-slab of a partitioned world build (vx_slab_build_noise)."""

    def __init__(self, X, Y, Z, factor, slab_index, slab_count, origin=(0, 0, 0), device=0):
        self.h = C.c_void_p()
        _check(lib().vx_slab_build_noise(C.byref(self.h), X, Y, Z, factor, origin[0], origin[1], origin[2], slab_index, slab_count, device))

    @property
    def info(self):
        i = SlabInfo()
        _check(lib().vx_slab_get_info(self.h, C.byref(i)))
        return i

    def export(self, d_top_words, d_bounds, d_pool_words, stream=0):
        _check(lib().vx_slab_export(self.h, int(d_top_words), int(d_bounds), int(d_pool_words) or None, int(stream) or None))

    def close(self):
        if self.h:
            lib().vx_slab_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class VoxelRaytracer3D:
    """GPUDDA::VoxelRaytracer3D (VolumeRaytracer.cuh:291-377): ray buffers for `count` rays bound to a world."""

    def __init__(self, world, count):
        self.world = world
        self.count = int(count)
        self.h = C.c_void_p()
        _check(lib().vx_tracer_create(C.byref(self.h), world.h, self.count))

    def close(self):
        if self.h:
            lib().vx_tracer_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    PLANES = dict(hit=(np.uint8, 1), pos=(np.float32, 3), normal=(np.float32, 3), steps=(np.int32, 1), voxel=(np.int32, 3),
                  brick=(np.uint32, 1), coarse_steps=(np.int32, 1), fine_steps=(np.int32, 1), counters=(np.uint32, 3))
    REFERENCE_PLANES = ("hit", "pos", "normal", "steps")  # what the reference's own Raytrace returns (VolumeRaytracer.cu:590-599)

    def raytrace(self, origins, dirs, norm_mode=NORM_DEVICE, max_steps=0, counters=False, planes=None, out=None):
        """VoxelRaytracer3D::Raytrace (VolumeRaytracer.cu:574-618) with host arrays.
        Returns a dict of numpy planes (hit, pos, normal, steps, voxel, brick, coarse_steps, fine_steps
        [, counters]) plus the reference's host-derived valid / distance.  `planes` restricts the result to the named
        planes (only those are computed, copied down and returned; valid / distance need "pos"); `out` reuses the arrays
        of an earlier call instead of allocating."""
        o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        n = o.shape[0]
        if d.shape[0] != n:
            raise VxError("origins and dirs differ in length")
        names = [k for k in self.PLANES if k != "counters"] + (["counters"] if counters else []) if planes is None else list(planes)
        res = {}
        for k in names:
            dt, w = self.PLANES[k]
            a = out.get(k) if out else None
            if a is None or a.dtype != dt or a.shape != ((n, w) if w > 1 else (n,)):
                a = np.zeros((n, w) if w > 1 else n, dt)
            res[k] = a
        h = Hits(**{k: v.ctypes.data for k, v in res.items()})
        opts = TraceOpts(norm_mode, max_steps)
        _check(lib().vx_trace(self.h, o.ctypes.data, d.ctypes.data, n, C.byref(opts), C.byref(h)))
        if "pos" in res and planes is None:
            res["valid"] = np.zeros(n, np.uint8)
            res["distance"] = np.zeros(n, np.float32)
            _check(lib().vx_hits_postprocess(o.ctypes.data, res["pos"].ctypes.data, n, res["valid"].ctypes.data,
                                             res["distance"].ctypes.data))
        return res

    def raytrace_device(self, d_origins, d_dirs, n, planes, norm_mode=NORM_DEVICE, max_steps=0, stream=0):
        """Device-pointer variant: `planes` maps plane name -> device pointer (int)."""
        h = Hits(**{k: int(v) for k, v in planes.items()})
        opts = TraceOpts(norm_mode, max_steps)
        _check(lib().vx_trace_device(self.h, int(d_origins), int(d_dirs), int(n), C.byref(opts), C.byref(h), int(stream)))

    def render_device(self, cam, env, opts, d_bgra, W, H, d_steps=0, stream=0):
        _check(lib().vx_render(self.h, C.byref(cam), C.byref(env), C.byref(opts), int(d_bgra), int(d_steps) or None, W, H,
                               int(stream) or None))

    def bounce_rays(self, cam, opts, W, H, d_origins, d_dirs, d_pixels, d_count, stream=0):
        """vx_bounce_rays: the AO sample ray of every primary hit of the frame, into device arrays (pointers as ints)."""
        _check(lib().vx_bounce_rays(self.h, C.byref(cam), C.byref(opts), W, H, int(d_origins), int(d_dirs), int(d_pixels), int(d_count),
                                    int(stream) or None))

    KERNELS = ("primary_rays", "shadow_setup", "shadow_rays", "shade_pixels")

    def set_timing(self, enable=True):
        """Record CUDA events around the launches of every following frame (vx_tracer_set_timing)."""
        _check(lib().vx_tracer_set_timing(self.h, 1 if enable else 0))

    def set_fused_shade(self, mode):
        """1: shade inside the sun-ray march (for a frame in another GPU's memory); 0: dense shade_pixels pass (default)."""
        _check(lib().vx_tracer_set_fused_shade(self.h, int(mode)))

    def frame_times(self):
        """Device ms of the last frame's launches, in KERNELS order (waits for the frame)."""
        ms = (C.c_float * 4)()
        _check(lib().vx_tracer_frame_times(self.h, C.byref(ms)))
        return [float(v) for v in ms]

    def render_host(self, cam, env, opts, W, H, bgra=None):
        if bgra is None:
            bgra = np.zeros((H, W, 4), np.uint8)
        _check(lib().vx_render_host(self.h, C.byref(cam), C.byref(env), C.byref(opts), bgra.ctypes.data, W, H))
        return bgra


def render_screen(rt, W, H, d_bgra, origin, fwd, up, right, env=None, fov_deg=90.0, frame_number=0, checkerboard=True,
                  debug_view=True, shadows=False, stream=0):
    """Graphics::RenderScreen (Renderer.cu:305-328) with the reference's shipped defaults
    (checkerboard on, DEBUG_VIEW on, shadow call disabled); globals are explicit arguments."""
    cam = Camera()
    cam.origin[:] = [float(v) for v in origin]
    cam.fwd[:] = [float(v) for v in fwd]
    cam.up[:] = [float(v) for v in up]
    cam.right[:] = [float(v) for v in right]
    cam.fov_deg = fov_deg
    opts = RenderOpts(int(checkerboard), frame_number, int(debug_view), int(shadows), 1, NORM_DEVICE, 0, 0, 0, 0, None)
    rt.render_device(cam, env or default_env(), opts, d_bgra, W, H, stream=stream)
