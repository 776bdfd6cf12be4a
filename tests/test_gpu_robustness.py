"""Loud failures instead of silent damage: a description with out-of-range ids / kinds / offsets is
refused by lj_scene_create (the kernels index with those values unchecked), and a BVH deeper than the traversal
stack covers is refused instead of dropping subtrees."""
import ctypes as C
import os

import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


def desc_field(desc, ptr_off, index, stride, field_off, ctype):
    base = C.c_void_p.from_address(desc.handle.value + ptr_off).value
    return ctype.from_address(base + index * stride + field_off)


# (what, pointer offset in lj_scene_desc, record stride, field offset, bad value, error text)
# offsets: lj_scene_desc pointer block starts at 400 (shapes, lights, materials, media, textures, images1, images3)
CORRUPTIONS = [
    ("material kind", 416, 64, 0, 99, "unknown kind"),
    ("material texture slot", 416, 64, 16, 12345, "texture slot"),
    ("shape material id", 400, 104, 4, 77, "material id"),
    ("shape area light id", 400, 104, 8, 50, "area light id"),
    ("light shape id", 408, 352, 4, 1000, "shape id"),
]


@pytest.mark.parametrize("what,ptr_off,stride,field_off,bad,text", CORRUPTIONS)
def test_malformed_description_is_refused(gpu, what, ptr_off, stride, field_off, bad, text):
    from lajolla_public_b200 import LjError
    from lajolla_public_b200.api import HostSceneDesc
    desc = HostSceneDesc(os.path.join(GOLDEN, "cbox.ljs"))
    try:
        f = desc_field(desc, ptr_off, 0, stride, field_off, C.c_int32)
        good = f.value
        f.value = bad
        with pytest.raises(LjError, match=text):
            gpu.scene_from_desc(desc)
        f.value = good
        gpu.scene_from_desc(desc).close()       # and the untouched description still loads
    finally:
        desc.close()


def test_light_table_offset_is_checked(gpu):
    from lajolla_public_b200 import LjError
    from lajolla_public_b200.api import HostSceneDesc
    desc = HostSceneDesc(os.path.join(GOLDEN, "cbox.ljs"))
    try:
        off = C.c_uint64.from_address(desc.handle.value + 560)      # light_cdf_offset
        good = off.value
        off.value = 1 << 40
        with pytest.raises(LjError, match="light distribution"):
            gpu.scene_from_desc(desc)
        off.value = good
    finally:
        desc.close()


def test_deep_tree_is_refused_not_truncated(gpu):
    """the traversal stack drops work when full (lj_traverse8.cuh); the build refuses trees it cannot cover"""
    from lajolla_public_b200 import LjError
    os.environ["LJ_BVH8_MAX_LEVELS"] = "1"
    try:
        with pytest.raises(LjError, match="levels deep"):
            gpu.load_scene(os.path.join(GOLDEN, "cbox.ljs"))
    finally:
        os.environ.pop("LJ_BVH8_MAX_LEVELS")
    gpu.load_scene(os.path.join(GOLDEN, "cbox.ljs")).close()
