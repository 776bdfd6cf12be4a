"""ctypes binding of include/lajolla_gpu.h (the drop-in boundary for
`Image3 render(const Scene&)`, /root/reference/src/render.h:9)."""
import ctypes as C
import os
import numpy as np

RNG_TILE_STREAMS = 0     # mode A: render.cpp:82 per-tile pcg32 streams (tile-serial)
RNG_SAMPLE_STREAMS = 1   # mode B: per-(pixel,sample) streams
ENGINE_WAVEFRONT = 0
ENGINE_MEGAKERNEL = 1
PARTITION_AUTO, PARTITION_TILES, PARTITION_SPP = 0, 1, 2   # lj_render_params.partition (multi-device contexts)

_HERE = os.path.dirname(os.path.abspath(__file__))


class LjError(RuntimeError):
    pass


def lib_path():
    # LJ_LIB: an alternative build of the same library (kernel-variant experiments, tools/)
    return os.environ.get("LJ_LIB") or os.path.join(_HERE, "liblajolla_b200.so")


class RenderParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("rng_mode", C.c_int32), ("engine", C.c_int32),
                ("sample_begin", C.c_int32), ("sample_end", C.c_int32),
                ("tile_rank", C.c_int32), ("tile_world", C.c_int32),
                ("max_depth_override", C.c_int32), ("paths_in_flight", C.c_int32), ("flags", C.c_int32),
                ("partition", C.c_int32)]

    def __init__(self, **kw):
        super().__init__()
        self.max_depth_override = -100
        self.rng_mode = RNG_SAMPLE_STREAMS
        for k, v in kw.items():
            setattr(self, k, v)


class Stats(C.Structure):
    _fields_ = [("render_ms", C.c_double), ("upload_ms", C.c_double), ("bvh_build_ms", C.c_double),
                ("samples", C.c_uint64), ("rays_primary", C.c_uint64), ("rays_extension", C.c_uint64),
                ("rays_shadow", C.c_uint64), ("nodes_visited", C.c_uint64), ("prims_tested", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("extend_ms", C.c_double), ("shadow_ms", C.c_double),
                ("shade_ms", C.c_double), ("generate_ms", C.c_double), ("bvh_nodes", C.c_uint64),
                ("bvh_prims", C.c_uint64), ("shadow_nodes_visited", C.c_uint64), ("shadow_prims_tested", C.c_uint64),
                ("iterations", C.c_uint64), ("tables_build_ms", C.c_double),
                ("n_devices", C.c_int32), ("partition", C.c_int32), ("reduce_ms", C.c_double),
                ("device_ms_min", C.c_double), ("device_ms_max", C.c_double)]

    @property
    def rays(self):
        return self.rays_primary + self.rays_extension + self.rays_shadow

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class Ray(C.Structure):
    _fields_ = [("org", C.c_float * 3), ("tnear", C.c_float), ("dir", C.c_float * 3), ("tfar", C.c_float)]


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("u", C.c_float), ("v", C.c_float), ("geom_id", C.c_int32),
                ("prim_id", C.c_int32), ("ng", C.c_float * 3)]


RAY_DTYPE = np.dtype([("org", "<f4", 3), ("tnear", "<f4"), ("dir", "<f4", 3), ("tfar", "<f4")])
HIT_DTYPE = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("geom_id", "<i4"), ("prim_id", "<i4"), ("ng", "<f4", 3)])

# every symbol include/lajolla_gpu.h declares
ABI_SYMBOLS = ["lj_init", "lj_init_devices", "lj_ctx_device_count", "lj_shutdown", "lj_scene_create", "lj_scene_destroy", "lj_render", "lj_render_device",
               "lj_intersect_batch", "lj_scene_info", "lj_last_error", "lj_scene_desc_save", "lj_scene_desc_load",
               "lj_scene_desc_free", "lj_scene_desc_pin", "lj_image_encoded_size", "lj_image_encode", "lj_image_write",
               "lj_accum_finalize", "lj_accum_finalize_static", "lj_accum_encode", "lj_accum_write", "lj_accum_save", "lj_accum_load",
               "lj_scene_create_ex", "lj_scene_read_derived", "lj_mesh_load_obj", "lj_mesh_load_serialized",
               "lj_mesh_compute_normals", "lj_mesh_free", "lj_volume_load", "lj_volume_free"]

SCENE_BUILD_TABLES = 1      # lj_scene_create_ex flags
SCENE_POISON_DERIVED = 2

# numpy mirrors of lj_shape / lj_light (include/lajolla_gpu.h) for reading record arrays
SHAPE_DTYPE = np.dtype([("kind", "<i4"), ("material_id", "<i4"), ("area_light_id", "<i4"), ("interior_medium_id", "<i4"),
                        ("exterior_medium_id", "<i4"), ("has_normals", "<i4"), ("has_uvs", "<i4"), ("_pad", "<i4"),
                        ("center", "<f8", 3), ("radius", "<f8"), ("vertex_offset", "<u4"), ("num_vertices", "<u4"),
                        ("triangle_offset", "<u4"), ("num_triangles", "<u4"), ("total_area", "<f8"),
                        ("tri_cdf_offset", "<u8"), ("tri_pmf_offset", "<u8")])
LIGHT_DTYPE = np.dtype([("kind", "<i4"), ("shape_id", "<i4"), ("intensity", "<f8", 3), ("values_texture", "<i4"), ("_pad", "<i4"),
                        ("to_world", "<f8", 16), ("to_local", "<f8", 16), ("scale", "<f8"), ("dist_width", "<i4"),
                        ("dist_height", "<i4"), ("cdf_rows_offset", "<u8"), ("pdf_rows_offset", "<u8"),
                        ("cdf_marginals_offset", "<u8"), ("pdf_marginals_offset", "<u8"), ("total_values", "<f8")])
IMAGE_DTYPE = np.dtype([("channels", "<i4"), ("num_levels", "<i4"), ("width", "<i4", 8), ("height", "<i4", 8), ("offset", "<u8", 8)])



class MeshData(C.Structure):       # lj_mesh_data
    _fields_ = [("num_vertices", C.c_uint64), ("num_triangles", C.c_uint64), ("has_normals", C.c_int32), ("has_uvs", C.c_int32),
                ("positions", C.POINTER(C.c_double)), ("normals", C.POINTER(C.c_double)), ("uvs", C.POINTER(C.c_double)),
                ("indices", C.POINTER(C.c_int32)), ("decode_ms", C.c_double)]


class VolumeData(C.Structure):     # lj_volume_data
    _fields_ = [("resolution", C.c_int32 * 3), ("channels", C.c_int32), ("p_min", C.c_double * 3), ("p_max", C.c_double * 3),
                ("max_data", C.c_double * 3), ("data", C.POINTER(C.c_double)), ("num_values", C.c_uint64), ("decode_ms", C.c_double)]


_IDENTITY = np.eye(4, dtype=np.float64)
_lib = None


def load_library():
    """dlopen liblajolla_b200.so; fails loudly if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    p = lib_path()
    if not os.path.exists(p):
        raise LjError(f"{p} is missing: build it with `make -C lajolla_public_b200` "
                      "(or __graft_entry__.build()); there is no CPU fallback")
    lib = C.CDLL(p)
    lib.lj_last_error.restype = C.c_char_p
    lib.lj_init.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    if hasattr(lib, "lj_init_devices"):      # (absent from older builds loaded through LJ_LIB for before/after runs)
        lib.lj_init_devices.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_void_p)]
        lib.lj_ctx_device_count.argtypes = [C.c_void_p]
    lib.lj_shutdown.argtypes = [C.c_void_p]
    lib.lj_shutdown.restype = None
    lib.lj_scene_create.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]
    lib.lj_scene_destroy.argtypes = [C.c_void_p]
    lib.lj_scene_destroy.restype = None
    lib.lj_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.POINTER(Stats)]
    lib.lj_render_device.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.POINTER(Stats)]
    lib.lj_intersect_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int]
    lib.lj_scene_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(Stats)]
    lib.lj_scene_desc_save.argtypes = [C.c_void_p, C.c_char_p]
    lib.lj_scene_desc_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p)]
    lib.lj_scene_desc_free.argtypes = [C.c_void_p]
    lib.lj_scene_desc_free.restype = None
    lib.lj_scene_desc_pin.argtypes = [C.c_void_p, C.c_int]
    lib.lj_scene_create_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.POINTER(C.c_void_p)]
    lib.lj_scene_read_derived.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64]
    lib.lj_image_encoded_size.argtypes = [C.c_int32, C.c_int32, C.c_int32]
    lib.lj_image_encoded_size.restype = C.c_uint64
    lib.lj_image_encode.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.lj_image_write.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32, C.c_int32]
    lib.lj_accum_finalize.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    if hasattr(lib, "lj_accum_finalize_static"):
        lib.lj_accum_finalize_static.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    lib.lj_accum_encode.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]
    lib.lj_accum_write.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    lib.lj_accum_save.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32]
    lib.lj_accum_load.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    lib.lj_mesh_load_obj.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.POINTER(MeshData)]
    lib.lj_mesh_load_serialized.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_int32, C.c_void_p, C.c_void_p, C.POINTER(MeshData)]
    lib.lj_mesh_compute_normals.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]
    lib.lj_mesh_free.argtypes = [C.POINTER(MeshData)]
    lib.lj_mesh_free.restype = None
    lib.lj_volume_load.argtypes = [C.c_void_p, C.c_char_p, C.c_uint64, C.c_int32, C.POINTER(VolumeData)]
    lib.lj_volume_free.argtypes = [C.POINTER(VolumeData)]
    lib.lj_volume_free.restype = None
    _lib = lib
    return lib


def _check(lib, rc, what):
    if rc != 0:
        raise LjError(f"{what} failed ({rc}): {lib.lj_last_error().decode(errors='replace')}")


class LajollaGPU:
    """One context per CUDA device (lj_init / lj_shutdown), or one over several devices of the box
    (lj_init_devices: `device` a list of ordinals, or "all")."""

    def __init__(self, device=0):
        self.lib = load_library()
        self.handle = C.c_void_p()
        if isinstance(device, (list, tuple)) or device == "all":
            ids = [] if device == "all" else [int(d) for d in device]
            arr = (C.c_int * max(len(ids), 1))(*ids)
            _check(self.lib, self.lib.lj_init_devices(arr if ids else None, len(ids), C.byref(self.handle)), "lj_init_devices")
        else:
            _check(self.lib, self.lib.lj_init(int(device), C.byref(self.handle)), "lj_init")
        self.device = device
        self.n_devices = int(self.lib.lj_ctx_device_count(self.handle)) if hasattr(self.lib, "lj_ctx_device_count") else 1

    def load_scene(self, blob_path, res=None, spp=None, integrator=None):
        return Scene(self, blob_path, res=res, spp=spp, integrator=integrator)

    def scene_from_desc(self, desc, flags=0):
        return Scene(self, desc=desc, flags=flags)

    # ---- output step (imwrite, image.cpp:135-173) ----
    def encode_image(self, img, fmt):
        """f32 (h, w, 3) image -> bytes of the .pfm / .exr file the reference's imwrite would write."""
        img = np.ascontiguousarray(img, dtype=np.float32)
        h, w = img.shape[:2]
        cap = self.lib.lj_image_encoded_size(int(fmt), w, h)
        buf = (C.c_ubyte * max(int(cap), 1))()
        n = C.c_uint64()
        _check(self.lib, self.lib.lj_image_encode(self.handle, int(fmt), img.ctypes.data, w, h, buf, cap, C.byref(n)), "lj_image_encode")
        return bytes(buf[:n.value])

    def write_image(self, path, img):
        img = np.ascontiguousarray(img, dtype=np.float32)
        _check(self.lib, self.lib.lj_image_write(self.handle, os.fsencode(path), img.ctypes.data, img.shape[1], img.shape[0]), "lj_image_write")

    def accum_finalize(self, dev_ptr, w, h, spp):
        out = np.empty((h, w, 3), dtype=np.float32)
        _check(self.lib, self.lib.lj_accum_finalize(self.handle, C.c_void_p(dev_ptr), w, h, int(spp), out.ctypes.data), "lj_accum_finalize")
        return out

    def accum_finalize_static(self, dev_ptr, w, h, spp, out=None):
        """lj_accum_finalize_static: the same without any allocation (static device scratch, <= 2^20 pixels)"""
        if out is None:
            out = np.empty((h, w, 3), dtype=np.float32)
        _check(self.lib, self.lib.lj_accum_finalize_static(self.handle, C.c_void_p(dev_ptr), w, h, int(spp), out.ctypes.data), "lj_accum_finalize_static")
        return out

    def accum_encode(self, dev_ptr, w, h, spp, fmt):
        cap = self.lib.lj_image_encoded_size(int(fmt), w, h)
        buf = (C.c_ubyte * max(int(cap), 1))()
        n = C.c_uint64()
        _check(self.lib, self.lib.lj_accum_encode(self.handle, int(fmt), C.c_void_p(dev_ptr), w, h, int(spp), buf, cap, C.byref(n)), "lj_accum_encode")
        return bytes(buf[:n.value])

    def accum_write(self, path, dev_ptr, w, h, spp):
        _check(self.lib, self.lib.lj_accum_write(self.handle, os.fsencode(path), C.c_void_p(dev_ptr), w, h, int(spp)), "lj_accum_write")

    def accum_save(self, path, dev_ptr, w, h, sample_begin, sample_end):
        _check(self.lib, self.lib.lj_accum_save(self.handle, os.fsencode(path), C.c_void_p(dev_ptr), w, h, int(sample_begin), int(sample_end)), "lj_accum_save")

    def accum_load(self, path, dev_ptr, w, h):
        b, e = C.c_int32(), C.c_int32()
        _check(self.lib, self.lib.lj_accum_load(self.handle, os.fsencode(path), C.c_void_p(dev_ptr), w, h, C.byref(b), C.byref(e)), "lj_accum_load")
        return b.value, e.value

    # ---- on-disk asset formats decoded on the device (parse_obj / load_serialized / load_volume_from_file) ----
    def _mesh_out(self, m):
        nv, nt = int(m.num_vertices), int(m.num_triangles)

        def arr(ptr, count, dtype, shape):
            if not ptr or count == 0:
                return np.zeros(shape, dtype=dtype)
            return np.ctypeslib.as_array(ptr, shape=(count,)).astype(dtype, copy=True).reshape(shape)
        out = {"positions": arr(m.positions, nv * 3, np.float64, (nv, 3)),
               "normals": arr(m.normals, nv * 3, np.float64, (nv, 3)) if m.has_normals else np.zeros((0, 3)),
               "uvs": arr(m.uvs, nv * 2, np.float64, (nv, 2)) if m.has_uvs else np.zeros((0, 2)),
               "indices": arr(m.indices, nt * 3, np.int32, (nt, 3)), "decode_ms": m.decode_ms}
        self.lib.lj_mesh_free(C.byref(m))
        return out

    @staticmethod
    def _xforms(to_world, to_local):
        tw = np.ascontiguousarray(_IDENTITY if to_world is None else to_world, dtype=np.float64).reshape(4, 4)
        tl = np.ascontiguousarray(np.linalg.inv(tw) if to_local is None else to_local, dtype=np.float64).reshape(4, 4)
        return tw, tl

    def load_obj(self, data, to_world=None, to_local=None):
        """parse_obj(file, to_world) on the device; `data` = the file's bytes.  Pass the reference's inverse(to_world)
        as `to_local` for bit-identical normals (numpy's inverse is used otherwise)."""
        tw, tl = self._xforms(to_world, to_local)
        m = MeshData()
        _check(self.lib, self.lib.lj_mesh_load_obj(self.handle, data, len(data), tw.ctypes.data, tl.ctypes.data, C.byref(m)), "lj_mesh_load_obj")
        return self._mesh_out(m)

    def load_serialized(self, data, shape_index=0, to_world=None, to_local=None):
        tw, tl = self._xforms(to_world, to_local)
        m = MeshData()
        _check(self.lib, self.lib.lj_mesh_load_serialized(self.handle, data, len(data), int(shape_index), tw.ctypes.data, tl.ctypes.data, C.byref(m)),
               "lj_mesh_load_serialized")
        return self._mesh_out(m)

    def compute_normals(self, positions, indices):
        pos = np.ascontiguousarray(positions, dtype=np.float64).reshape(-1, 3)
        idx = np.ascontiguousarray(indices, dtype=np.int32).reshape(-1, 3)
        out = np.empty_like(pos)
        _check(self.lib, self.lib.lj_mesh_compute_normals(self.handle, pos.ctypes.data, pos.shape[0], idx.ctypes.data, idx.shape[0], out.ctypes.data),
               "lj_mesh_compute_normals")
        return out

    def load_volume(self, data, channels=3):
        v = VolumeData()
        _check(self.lib, self.lib.lj_volume_load(self.handle, data, len(data), int(channels), C.byref(v)), "lj_volume_load")
        res = tuple(v.resolution)
        grid = np.ctypeslib.as_array(v.data, shape=(int(v.num_values),)).copy().reshape(res[2], res[1], res[0], channels)
        out = {"resolution": res, "p_min": np.array(v.p_min), "p_max": np.array(v.p_max), "max_data": np.array(v.max_data)[:channels],
               "data": grid, "decode_ms": v.decode_ms}
        self.lib.lj_volume_free(C.byref(v))
        return out

    def close(self):
        if self.handle:
            self.lib.lj_shutdown(self.handle)
            self.handle = C.c_void_p()


IMAGE_PFM, IMAGE_EXR = 0, 1


class HostSceneDesc:
    """A flattened scene held in HOST memory (lj_scene_desc_load); optionally page-locked."""

    def __init__(self, blob_path, pin=False, res=None, spp=None, integrator=None):
        self.lib = load_library()
        self.handle = C.c_void_p()
        _check(self.lib, self.lib.lj_scene_desc_load(os.fsencode(blob_path), C.byref(self.handle)), "lj_scene_desc_load")
        self.nbytes = os.path.getsize(blob_path)
        if res is not None:   # film-size override (lj_camera.width/height follow the two 4x4 f64 matrices)
            wh = (C.c_int32 * 2).from_address(self.handle.value + 256)
            wh[0], wh[1] = int(res[0]), int(res[1])
        if spp is not None:   # lj_options.samples_per_pixel
            opt = (C.c_int32 * 6).from_address(self.handle.value + 280)
            opt[1] = int(spp)
        if integrator is not None:   # lj_options.integrator (LJ_INTEGRATOR_*): what parse_scene.cpp:256-304 selects
            opt = (C.c_int32 * 6).from_address(self.handle.value + 280)
            opt[0] = int(integrator)
        self.pinned = bool(pin) and self.lib.lj_scene_desc_pin(self.handle, 1) == 0

    def _u64(self, off):
        return C.c_uint64.from_address(self.handle.value + off).value

    def _array(self, ptr_off, count, dtype):
        ptr = C.c_void_p.from_address(self.handle.value + ptr_off).value
        if not ptr or count == 0:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype)

    def arrays(self):
        """numpy views of the host arrays the scene-build step derives (offsets: see lj_scene_desc)."""
        return {"shapes": self._array(400, self._u64(344), SHAPE_DTYPE), "lights": self._array(408, self._u64(352), LIGHT_DTYPE),
                "images1": self._array(440, self._u64(384), IMAGE_DTYPE), "images3": self._array(448, self._u64(392), IMAGE_DTYPE),
                "texels": self._array(528, self._u64(472), np.float64), "tables": self._array(536, self._u64(480), np.float64),
                "light_pmf_offset": self._u64(552), "light_cdf_offset": self._u64(560)}

    def close(self):
        if self.handle:
            if self.pinned:
                self.lib.lj_scene_desc_pin(self.handle, 0)
            self.lib.lj_scene_desc_free(self.handle)
            self.handle = C.c_void_p()


class Scene:
    """A flattened scene blob (.ljs, written by bin/lj_flatten or LJ_SAVE_SCENE) resident on the GPU."""

    def __init__(self, gpu, blob_path=None, res=None, spp=None, desc=None, flags=0, integrator=None):
        self.gpu = gpu
        lib = gpu.lib
        own = desc is None
        if own:
            desc = HostSceneDesc(blob_path, res=res, spp=spp, integrator=integrator)
        self.handle = C.c_void_p()
        try:
            if flags:
                _check(lib, lib.lj_scene_create_ex(gpu.handle, desc.handle, int(flags), C.byref(self.handle)), "lj_scene_create_ex")
            else:
                _check(lib, lib.lj_scene_create(gpu.handle, desc.handle, C.byref(self.handle)), "lj_scene_create")
        finally:
            if own:
                desc.close()
        w, h, spp = C.c_int32(), C.c_int32(), C.c_int32()
        self.build_stats = Stats()
        _check(lib, lib.lj_scene_info(self.handle, C.byref(w), C.byref(h), C.byref(spp), C.byref(self.build_stats)), "lj_scene_info")
        self.width, self.height, self.spp = w.value, h.value, spp.value

    def read_derived(self, num_tables, num_texels, num_shapes, num_lights):
        """device copies of tables / texels / shape and light records (lj_scene_read_derived)"""
        tables = np.empty(num_tables, dtype=np.float64); texels = np.empty(num_texels, dtype=np.float64)
        shapes = np.empty(num_shapes, dtype=SHAPE_DTYPE); lights = np.empty(num_lights, dtype=LIGHT_DTYPE)
        _check(self.gpu.lib, self.gpu.lib.lj_scene_read_derived(self.handle, tables.ctypes.data, num_tables, texels.ctypes.data, num_texels,
                                                               shapes.ctypes.data, num_shapes, lights.ctypes.data, num_lights),
               "lj_scene_read_derived")
        return tables, texels, shapes, lights

    def render(self, params=None, out=None):
        """render(): host f32 image (h, w, 3), y=0 top.  `out` may be a pinned/host f32 array to fill."""
        lib = self.gpu.lib
        params = params or RenderParams()
        if out is None:
            out = np.empty((self.height, self.width, 3), dtype=np.float32)
        assert out.dtype == np.float32 and out.size == self.height * self.width * 3 and out.flags["C_CONTIGUOUS"]
        st = Stats()
        _check(lib, lib.lj_render(self.handle, C.byref(params), out.ctypes.data_as(C.c_void_p), C.byref(st)), "lj_render")
        return out, st

    def render_into_device(self, dev_ptr, params=None):
        """render radiance SUMS (f64, h*w*3) into caller-owned device memory (e.g. a torch tensor's data_ptr())."""
        lib = self.gpu.lib
        params = params or RenderParams()
        st = Stats()
        _check(lib, lib.lj_render_device(self.handle, C.byref(params), C.c_void_p(int(dev_ptr)), C.byref(st)), "lj_render_device")
        return st

    def intersect(self, rays, any_hit=False):
        """rtcIntersect1 / rtcOccluded1 for a batch: rays is a RAY_DTYPE array; returns a HIT_DTYPE array."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        lib = self.gpu.lib
        _check(lib, lib.lj_intersect_batch(self.handle, rays.ctypes.data_as(C.c_void_p), rays.shape[0],
                                           hits.ctypes.data_as(C.c_void_p), 1 if any_hit else 0), "lj_intersect_batch")
        return hits

    def close(self):
        if self.handle:
            self.gpu.lib.lj_scene_destroy(self.handle)
            self.handle = C.c_void_p()


def read_pfm(path):
    """PFM as written by the reference's imwrite (image.cpp:139-149): rows top to bottom, scale -1."""
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), dtype="<f4" if scale < 0 else ">f4")
    return data.reshape(h, w, 3).astype(np.float32)


def write_pfm(path, img):
    img = np.ascontiguousarray(img, dtype="<f4")
    with open(path, "wb") as f:
        f.write(b"PF\n%d %d\n-1\n" % (img.shape[1], img.shape[0]))
        f.write(img.tobytes())

