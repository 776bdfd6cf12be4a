// Scene -> lj_scene_desc.  Host side, runs once per render() call.
// Field sources (all under /root/reference/src): scene.h:58-81, shape.h:26-50,
// light.h:15-30, material.h:11-100, texture.h:75-92, mipmap.h:8-48,
// medium.h:10-22, volume.h:13-27, camera.h:10-25, table_dist.h:9-37.
#include "scene_flatten.h"

#include "scene.h"   // reference header (found via -I<reference>/src)

#include <cstring>
#include <type_traits>

namespace {

template <class T> struct always_false : std::false_type {};

uint64_t push_table(FlatScene &fs, const std::vector<Real> &v) {
    uint64_t off = fs.tables.size();
    fs.tables.insert(fs.tables.end(), v.begin(), v.end());
    return off;
}

void set3(double *dst, const Vector3 &v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }
void set3(double *dst, Real v) { dst[0] = v; dst[1] = 0; dst[2] = 0; }

void set_mat(double *dst, const Matrix4x4 &m) {
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) dst[i * 4 + j] = m(i, j);
}

template <typename T>
int add_texture(FlatScene &fs, const Texture<T> &tex) {
    lj_texture t; memset(&t, 0, sizeof t);
    t.channels = std::is_same<T, Real>::value ? 1 : 3;
    t.image_id = -1;
    t.uscale = t.vscale = 1;
    if (auto *c = std::get_if<ConstantTexture<T>>(&tex)) {
        t.kind = LJ_TEX_CONSTANT; set3(t.value, c->value);
    } else if (auto *im = std::get_if<ImageTexture<T>>(&tex)) {
        t.kind = LJ_TEX_IMAGE; t.image_id = im->texture_id;
        t.uscale = im->uscale; t.vscale = im->vscale; t.uoffset = im->uoffset; t.voffset = im->voffset;
    } else if (auto *cb = std::get_if<CheckerboardTexture<T>>(&tex)) {
        t.kind = LJ_TEX_CHECKERBOARD; set3(t.value, cb->color0); set3(t.color1, cb->color1);
        t.uscale = cb->uscale; t.vscale = cb->vscale; t.uoffset = cb->uoffset; t.voffset = cb->voffset;
    }
    fs.textures.push_back(t);
    return (int)fs.textures.size() - 1;
}

template <typename T>
lj_image add_image(FlatScene &fs, const Mipmap<T> &mip) {
    lj_image im; memset(&im, 0, sizeof im);
    im.channels = std::is_same<T, Real>::value ? 1 : 3;
    im.num_levels = (int)mip.images.size();
    for (int l = 0; l < im.num_levels && l < 8; l++) {
        const Image<T> &img = mip.images[l];
        im.width[l] = img.width; im.height[l] = img.height;
        im.offset[l] = fs.texels.size();
        const double *p = (const double *)img.data.data();
        fs.texels.insert(fs.texels.end(), p, p + (size_t)img.width * img.height * im.channels);
    }
    return im;
}

lj_volume flat_volume(FlatScene &fs, const VolumeSpectrum &vol) {
    lj_volume v; memset(&v, 0, sizeof v);
    if (auto *c = std::get_if<ConstantVolume<Spectrum>>(&vol)) {
        v.kind = LJ_VOLUME_CONSTANT; set3(v.value, c->value); v.scale = 1;
    } else if (auto *g = std::get_if<GridVolume<Spectrum>>(&vol)) {
        v.kind = LJ_VOLUME_GRID;
        v.resolution[0] = g->resolution.x; v.resolution[1] = g->resolution.y; v.resolution[2] = g->resolution.z;
        set3(v.p_min, g->p_min); set3(v.p_max, g->p_max); set3(v.max_data, g->max_data);
        v.scale = g->scale;
        v.data_offset = fs.grid_data.size();
        const double *p = (const double *)g->data.data();
        fs.grid_data.insert(fs.grid_data.end(), p, p + g->data.size() * 3);
    }
    return v;
}

lj_material flat_material(FlatScene &fs, const Material &mat) {
    lj_material m; memset(&m, 0, sizeof m);
    for (int &t : m.tex) t = -1;
    m.kind = (int)mat.index();
    std::visit([&](const auto &b) {
        using B = std::decay_t<decltype(b)>;
        if constexpr (std::is_same_v<B, Lambertian>) {
            m.tex[LJ_SLOT_REFLECTANCE] = add_texture(fs, b.reflectance);
        } else if constexpr (std::is_same_v<B, RoughPlastic>) {
            m.tex[LJ_SLOT_DIFFUSE] = add_texture(fs, b.diffuse_reflectance);
            m.tex[LJ_SLOT_SPECULAR] = add_texture(fs, b.specular_reflectance);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.eta = b.eta;
        } else if constexpr (std::is_same_v<B, RoughDielectric>) {
            m.tex[LJ_SLOT_SPEC_REFL] = add_texture(fs, b.specular_reflectance);
            m.tex[LJ_SLOT_SPEC_TRANS] = add_texture(fs, b.specular_transmittance);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.eta = b.eta;
        } else if constexpr (std::is_same_v<B, DisneyDiffuse>) {
            m.tex[LJ_SLOT_BASE_COLOR] = add_texture(fs, b.base_color);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.tex[LJ_SLOT_SUBSURFACE] = add_texture(fs, b.subsurface);
        } else if constexpr (std::is_same_v<B, DisneyMetal>) {
            m.tex[LJ_SLOT_BASE_COLOR] = add_texture(fs, b.base_color);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.tex[LJ_SLOT_ANISOTROPIC] = add_texture(fs, b.anisotropic);
        } else if constexpr (std::is_same_v<B, DisneyGlass>) {
            m.tex[LJ_SLOT_BASE_COLOR] = add_texture(fs, b.base_color);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.tex[LJ_SLOT_ANISOTROPIC] = add_texture(fs, b.anisotropic);
            m.eta = b.eta;
        } else if constexpr (std::is_same_v<B, DisneyClearcoat>) {
            m.tex[LJ_SLOT_CLEARCOAT_GLOSS] = add_texture(fs, b.clearcoat_gloss);
        } else if constexpr (std::is_same_v<B, DisneySheen>) {
            m.tex[LJ_SLOT_BASE_COLOR] = add_texture(fs, b.base_color);
            m.tex[LJ_SLOT_SHEEN_TINT] = add_texture(fs, b.sheen_tint);
        } else if constexpr (std::is_same_v<B, DisneyBSDF>) {
            m.tex[LJ_SLOT_BASE_COLOR] = add_texture(fs, b.base_color);
            m.tex[LJ_SLOT_SPEC_TRANSMISSION] = add_texture(fs, b.specular_transmission);
            m.tex[LJ_SLOT_METALLIC] = add_texture(fs, b.metallic);
            m.tex[LJ_SLOT_SUBSURFACE] = add_texture(fs, b.subsurface);
            m.tex[LJ_SLOT_SPECULAR_D] = add_texture(fs, b.specular);
            m.tex[LJ_SLOT_ROUGHNESS] = add_texture(fs, b.roughness);
            m.tex[LJ_SLOT_SPECULAR_TINT] = add_texture(fs, b.specular_tint);
            m.tex[LJ_SLOT_ANISOTROPIC] = add_texture(fs, b.anisotropic);
            m.tex[LJ_SLOT_SHEEN] = add_texture(fs, b.sheen);
            m.tex[LJ_SLOT_SHEEN_TINT] = add_texture(fs, b.sheen_tint);
            m.tex[LJ_SLOT_CLEARCOAT] = add_texture(fs, b.clearcoat);
            m.tex[LJ_SLOT_CLEARCOAT_GLOSS] = add_texture(fs, b.clearcoat_gloss);
            m.eta = b.eta;
        } else {
            static_assert(always_false<B>::value, "unhandled material");
        }
    }, mat);
    return m;
}

} // namespace

void FlatScene::finalize() {
    desc.num_shapes = shapes.size(); desc.shapes = shapes.data();
    desc.num_lights = lights.size(); desc.lights = lights.data();
    desc.num_materials = materials.size(); desc.materials = materials.data();
    desc.num_media = media.size(); desc.media = media.data();
    desc.num_textures = textures.size(); desc.textures = textures.data();
    desc.num_images1 = images1.size(); desc.images1 = images1.data();
    desc.num_images3 = images3.size(); desc.images3 = images3.data();
    desc.num_positions = positions.size() / 3; desc.positions = positions.data();
    desc.normals = normals.data(); desc.uvs = uvs.data();
    desc.num_indices = indices.size() / 3; desc.indices = indices.data();
    desc.num_texels = texels.size(); desc.texels = texels.data();
    desc.num_tables = tables.size(); desc.tables = tables.data();
    desc.num_grid_data = grid_data.size(); desc.grid_data = grid_data.data();
}

void flatten_scene(const Scene &scene, FlatScene &fs) {
    memset(&fs.desc, 0, sizeof fs.desc);
    lj_scene_desc &d = fs.desc;

    // camera (camera.h:10-25) and pixel filter (filter.h:30-45)
    set_mat(d.camera.sample_to_cam, scene.camera.sample_to_cam);
    set_mat(d.camera.cam_to_world, scene.camera.cam_to_world);
    d.camera.width = scene.camera.width; d.camera.height = scene.camera.height;
    d.camera.medium_id = scene.camera.medium_id;
    d.camera.filter_kind = (int)scene.camera.filter.index();
    if (auto *f = std::get_if<Box>(&scene.camera.filter)) d.camera.filter_param = f->width;
    if (auto *f = std::get_if<Tent>(&scene.camera.filter)) d.camera.filter_param = f->width;
    if (auto *f = std::get_if<Gaussian>(&scene.camera.filter)) d.camera.filter_param = f->stddev;

    d.options.integrator = (int)scene.options.integrator;
    d.options.samples_per_pixel = scene.options.samples_per_pixel;
    d.options.max_depth = scene.options.max_depth;
    d.options.rr_depth = scene.options.rr_depth;
    d.options.vol_path_version = scene.options.vol_path_version;
    d.options.max_null_collisions = scene.options.max_null_collisions;
    d.bounds_radius = scene.bounds.radius;
    set3(d.bounds_center, scene.bounds.center);
    d.envmap_light_id = scene.envmap_light_id;

    // image pools first, so that texture.image_id indexes them directly
    for (const Mipmap1 &m : scene.texture_pool.image1s) fs.images1.push_back(add_image(fs, m));
    for (const Mipmap3 &m : scene.texture_pool.image3s) fs.images3.push_back(add_image(fs, m));

    for (const Material &m : scene.materials) fs.materials.push_back(flat_material(fs, m));

    for (const Shape &shape : scene.shapes) {
        lj_shape s; memset(&s, 0, sizeof s);
        s.kind = (int)shape.index();
        s.material_id = get_material_id(shape); s.area_light_id = get_area_light_id(shape);
        s.interior_medium_id = get_interior_medium_id(shape); s.exterior_medium_id = get_exterior_medium_id(shape);
        if (auto *sp = std::get_if<Sphere>(&shape)) {
            set3(s.center, sp->position); s.radius = sp->radius;
        } else if (auto *mesh = std::get_if<TriangleMesh>(&shape)) {
            s.vertex_offset = (uint32_t)(fs.positions.size() / 3);
            s.num_vertices = (uint32_t)mesh->positions.size();
            s.triangle_offset = (uint32_t)(fs.indices.size() / 3);
            s.num_triangles = (uint32_t)mesh->indices.size();
            s.has_normals = mesh->normals.size() > 0; s.has_uvs = mesh->uvs.size() > 0;
            for (const Vector3 &p : mesh->positions) { fs.positions.push_back(p.x); fs.positions.push_back(p.y); fs.positions.push_back(p.z); }
            for (size_t i = 0; i < mesh->positions.size(); i++) {
                Vector3 n = s.has_normals ? mesh->normals[i] : Vector3{0, 0, 0};
                Vector2 uv = s.has_uvs ? mesh->uvs[i] : Vector2{0, 0};
                fs.normals.push_back(n.x); fs.normals.push_back(n.y); fs.normals.push_back(n.z);
                fs.uvs.push_back(uv.x); fs.uvs.push_back(uv.y);
            }
            for (const Vector3i &t : mesh->indices) { fs.indices.push_back(t.x); fs.indices.push_back(t.y); fs.indices.push_back(t.z); }
            s.total_area = mesh->total_area;
            s.tri_pmf_offset = push_table(fs, mesh->triangle_sampler.pmf);
            s.tri_cdf_offset = push_table(fs, mesh->triangle_sampler.cdf);
        }
        fs.shapes.push_back(s);
    }

    for (const Light &light : scene.lights) {
        lj_light l; memset(&l, 0, sizeof l);
        l.kind = (int)light.index(); l.shape_id = -1; l.values_texture = -1;
        if (auto *a = std::get_if<DiffuseAreaLight>(&light)) {
            l.shape_id = a->shape_id; set3(l.intensity, a->intensity);
        } else if (auto *e = std::get_if<Envmap>(&light)) {
            l.values_texture = add_texture(fs, e->values);
            set_mat(l.to_world, e->to_world); set_mat(l.to_local, e->to_local);
            l.scale = e->scale;
            const TableDist2D &t = e->sampling_dist;
            l.dist_width = t.width; l.dist_height = t.height; l.total_values = t.total_values;
            l.cdf_rows_offset = push_table(fs, t.cdf_rows); l.pdf_rows_offset = push_table(fs, t.pdf_rows);
            l.cdf_marginals_offset = push_table(fs, t.cdf_marginals); l.pdf_marginals_offset = push_table(fs, t.pdf_marginals);
        }
        fs.lights.push_back(l);
    }
    d.light_pmf_offset = push_table(fs, scene.light_dist.pmf);
    d.light_cdf_offset = push_table(fs, scene.light_dist.cdf);

    for (const Medium &medium : scene.media) {
        lj_medium m; memset(&m, 0, sizeof m);
        m.kind = (int)medium.index();
        PhaseFunction pf = get_phase_function(medium);
        m.phase_kind = (int)pf.index();
        if (auto *hg = std::get_if<HenyeyGreenstein>(&pf)) m.g = hg->g;
        if (auto *h = std::get_if<HomogeneousMedium>(&medium)) {
            set3(m.sigma_a, h->sigma_a); set3(m.sigma_s, h->sigma_s);
        } else if (auto *het = std::get_if<HeterogeneousMedium>(&medium)) {
            m.albedo = flat_volume(fs, het->albedo); m.density = flat_volume(fs, het->density);
        }
        fs.media.push_back(m);
    }
    fs.finalize();
}
