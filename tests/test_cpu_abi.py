"""CPU-side checks of the drop-in boundary (no GPU compute): the C-ABI library loads, exports
every entry point include/lajolla_gpu.h declares, fails LOUDLY without a CUDA device (no CPU
fallback), and the host-side blob reader round-trips a flattened scene."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

HEADER = os.path.join(ROOT, "include", "lajolla_gpu.h")


def header_entry_points():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lj_[a-z0-9_]+)\s*\(", src)))


def test_header_lists_entry_points():
    names = header_entry_points()
    for must in ("lj_init", "lj_shutdown", "lj_scene_create", "lj_scene_destroy", "lj_render", "lj_render_device",
                 "lj_intersect_batch", "lj_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from lajolla_public_b200 import load_library, lib_path
    from lajolla_public_b200.api import ABI_SYMBOLS
    lib = load_library()
    declared = header_entry_points()
    assert sorted(ABI_SYMBOLS) == declared, "api.ABI_SYMBOLS and the header disagree"
    for name in declared:
        assert hasattr(lib, name), f"{lib_path()} does not export {name}"
    # the exported dynamic symbols carry no C++ mangling for the ABI names
    out = subprocess.run(["nm", "-D", "--defined-only", lib_path()], stdout=subprocess.PIPE, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if l.strip()}
    for name in declared:
        assert name in exported


def test_library_is_sm100a_cuda_code():
    """the shared object embeds sm_100a SASS for the hot-path kernels (not a host-only stub)"""
    from lajolla_public_b200 import lib_path
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not present")
    out = subprocess.run([cuobjdump, "-lelf", lib_path()], stdout=subprocess.PIPE, text=True).stdout
    assert "sm_100a" in out
    syms = subprocess.run([cuobjdump, "-elf", lib_path()], stdout=subprocess.PIPE, text=True).stdout
    for kernel in ("k_wf_extend", "k_wf_shadow", "k_wf_shade", "k_wf_generate", "k_tiles_mode_a"):
        assert kernel in syms, f"kernel {kernel} missing from the fatbinary"


def test_no_device_fails_loudly():
    """without a CUDA device lj_init must fail with an error code and a message: there is no CPU path"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the no-device path cannot be exercised")
    from lajolla_public_b200 import LajollaGPU, LjError
    with pytest.raises(LjError) as ei:
        LajollaGPU(0)
    assert "lj_init failed" in str(ei.value)
    # the multi-device entry point fails the same way (and loads neither NCCL nor anything else on the way)
    with pytest.raises(LjError) as ei:
        LajollaGPU("all")
    assert "lj_init_devices failed" in str(ei.value) and "no CUDA device" in str(ei.value)
    from lajolla_public_b200 import load_library
    assert load_library().lj_ctx_device_count(None) == 0
    # and so does the allocation-free finalize: argument checks come first, no device is touched
    assert load_library().lj_accum_finalize_static(None, None, 1, 1, 1, None) < 0


def test_null_arguments_are_rejected():
    from lajolla_public_b200 import load_library
    lib = load_library()
    assert lib.lj_scene_create(None, None, None) < 0
    assert lib.lj_render(None, None, None, None) < 0
    assert lib.lj_render_device(None, None, None, None) < 0
    assert lib.lj_intersect_batch(None, None, 0, None, 0) < 0
    assert lib.lj_last_error() not in (None, b"")
    h = C.c_void_p()
    assert lib.lj_scene_desc_load(b"/nonexistent/file.ljs", C.byref(h)) < 0


def test_struct_sizes_match_header():
    """ctypes mirrors of lj_render_params / lj_stats / lj_ray / lj_hit have the C layout"""
    from lajolla_public_b200.api import RenderParams, Stats, Ray, Hit, RAY_DTYPE, HIT_DTYPE
    src = r'''
    #include <stdio.h>
    #include "lajolla_gpu.h"
    int main(void) { printf("%zu %zu %zu %zu %zu %zu\n", sizeof(lj_render_params), sizeof(lj_stats), sizeof(lj_ray),
                            sizeof(lj_hit), sizeof(lj_camera), sizeof(lj_scene_desc)); return 0; }
    '''
    exe = "/tmp/lj_sizes"
    r = subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src, text=True,
                       stderr=subprocess.PIPE)
    assert r.returncode == 0, r.stderr   # the header is plain C
    sizes = list(map(int, subprocess.run([exe], stdout=subprocess.PIPE, text=True).stdout.split()))
    assert C.sizeof(RenderParams) == sizes[0]
    assert C.sizeof(Stats) == sizes[1]
    assert C.sizeof(Ray) == sizes[2] == RAY_DTYPE.itemsize
    assert C.sizeof(Hit) == sizes[3] == HIT_DTYPE.itemsize
    assert sizes[4] == 280      # api.HostSceneDesc pokes width/height at +256 and options at +280
    assert sizes[5] == 568      # blob header size used by test_blob_roundtrip


def test_blob_roundtrip(tmp_path):
    """lj_scene_desc_load -> lj_scene_desc_save keeps every array byte for byte (the header's host
    pointers are written as null), and a second round trip is the identity"""
    from lajolla_public_b200 import load_library
    from lajolla_public_b200.api import HostSceneDesc
    lib = load_library()
    for name in ("cbox", "veach_mi"):
        src = os.path.join(GOLDEN, name + ".ljs")
        d = HostSceneDesc(src)
        out = str(tmp_path / (name + ".ljs"))
        assert lib.lj_scene_desc_save(d.handle, os.fsencode(out)) == 0
        d.close()
        a, b = open(src, "rb").read(), open(out, "rb").read()
        assert len(a) == len(b)
        hdr = 4 + 568                       # magic + sizeof(lj_scene_desc)
        assert a[hdr:] == b[hdr:]           # all arrays
        assert a[:4 + 400] == b[:4 + 400]   # camera, options, bounds, ids, counts (before the first pointer)
        d2 = HostSceneDesc(out)
        out2 = str(tmp_path / (name + "_2.ljs"))
        assert lib.lj_scene_desc_save(d2.handle, os.fsencode(out2)) == 0
        d2.close()
        assert open(out2, "rb").read() == b


def test_blob_film_override():
    from lajolla_public_b200.api import HostSceneDesc
    d = HostSceneDesc(os.path.join(GOLDEN, "cbox.ljs"), res=(128, 96), spp=7)
    wh = (C.c_int32 * 2).from_address(d.handle.value + 256)
    opt = (C.c_int32 * 6).from_address(d.handle.value + 280)
    assert (wh[0], wh[1]) == (128, 96)
    assert opt[1] == 7 and opt[0] == 5      # LJ_INTEGRATOR_PATH
    d.close()


def test_asset_loaders_reject_null_arguments_without_a_device():
    """the asset decoders validate their arguments before touching CUDA and have no host decode path:
    without a context they fail with LJ_EINVAL and a message; the free functions accept an empty record"""
    from lajolla_public_b200 import load_library
    from lajolla_public_b200.api import MeshData, VolumeData
    lib = load_library()
    m, v = MeshData(), VolumeData()
    eye = np.eye(4)
    assert lib.lj_mesh_load_obj(None, b"v 0 0 0\n", 8, eye.ctypes.data, eye.ctypes.data, C.byref(m)) == -1
    assert b"lj_mesh_load_obj" in lib.lj_last_error()
    assert lib.lj_mesh_load_serialized(None, b"\x1c\x04\x04\x00", 4, 0, eye.ctypes.data, eye.ctypes.data, C.byref(m)) == -1
    assert lib.lj_mesh_compute_normals(None, None, 0, None, 0, None) == -1
    assert lib.lj_volume_load(None, b"VOL\x03", 4, 3, C.byref(v)) == -1
    assert b"lj_volume_load" in lib.lj_last_error()
    lib.lj_mesh_free(C.byref(m)); lib.lj_volume_free(C.byref(v))      # zeroed records: no-ops
    assert not m.positions and not v.data
