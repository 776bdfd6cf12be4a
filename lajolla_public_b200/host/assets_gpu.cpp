// GPU drop-ins for the reference's asset loaders, same signatures, linked INSTEAD of
// src/parse_obj.cpp, src/load_serialized.cpp and src/volume.cpp (bin/lajolla_ga):
//     TriangleMesh parse_obj(const fs::path &filename, const Matrix4x4 &to_world);                       // parse_obj.h:9
//     TriangleMesh load_serialized(const fs::path &filename, int shape_index, const Matrix4x4 &to_world); // load_serialized.h:8
//     template <> GridVolume<Real|Spectrum> load_volume_from_file(const fs::path &filename);              // volume.h:171-175
// Each reads the file's bytes and hands them to the C ABI (lj_mesh_load_obj / lj_mesh_load_serialized /
// lj_volume_load), which decodes on the device; failures become the reference's Error() exception.
#include "parse_obj.h"
#include "load_serialized.h"
#include "volume.h"
#include "flexception.h"
#include "host_ctx.h"

#include <fstream>
#include <iterator>
#include <string>
#include <vector>

static std::vector<char> slurp(const fs::path &filename, const char *what) {
    std::ifstream f(filename.c_str(), std::ios::binary);
    if (!f.is_open()) Error(std::string(what) + ": unable to open " + filename.string());
    return std::vector<char>((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
}

static void matrices(const Matrix4x4 &to_world, double tw[16], double tl[16]) {
    const Matrix4x4 inv = inverse(to_world);                       // matrix.h:80, as parse_obj.cpp:132 / load_serialized.cpp:222
    for (int i = 0; i < 16; i++) { tw[i] = to_world(i / 4, i % 4); tl[i] = inv(i / 4, i % 4); }
}

static TriangleMesh to_mesh(lj_mesh_data &m) {
    TriangleMesh mesh;
    mesh.positions.resize(m.num_vertices);
    for (size_t i = 0; i < m.num_vertices; i++) mesh.positions[i] = Vector3{m.positions[3 * i], m.positions[3 * i + 1], m.positions[3 * i + 2]};
    if (m.has_normals) {
        mesh.normals.resize(m.num_vertices);
        for (size_t i = 0; i < m.num_vertices; i++) mesh.normals[i] = Vector3{m.normals[3 * i], m.normals[3 * i + 1], m.normals[3 * i + 2]};
    }
    if (m.has_uvs) {
        mesh.uvs.resize(m.num_vertices);
        for (size_t i = 0; i < m.num_vertices; i++) mesh.uvs[i] = Vector2{m.uvs[2 * i], m.uvs[2 * i + 1]};
    }
    mesh.indices.resize(m.num_triangles);
    for (size_t i = 0; i < m.num_triangles; i++) mesh.indices[i] = Vector3i{m.indices[3 * i], m.indices[3 * i + 1], m.indices[3 * i + 2]};
    lj_mesh_free(&m);
    return mesh;
}

TriangleMesh parse_obj(const fs::path &filename, const Matrix4x4 &to_world) {
    std::vector<char> text = slurp(filename, "Unable to open the obj file");
    double tw[16], tl[16]; matrices(to_world, tw, tl);
    lj_mesh_data m;
    if (lj_mesh_load_obj(lj_host_ctx(), text.data(), text.size(), tw, tl, &m) != LJ_OK)
        Error(filename.string() + ": " + lj_last_error());
    return to_mesh(m);
}

TriangleMesh load_serialized(const fs::path &filename, int shape_index, const Matrix4x4 &to_world) {
    std::vector<char> bytes = slurp(filename, "load_serialized");
    double tw[16], tl[16]; matrices(to_world, tw, tl);
    lj_mesh_data m;
    if (lj_mesh_load_serialized(lj_host_ctx(), bytes.data(), bytes.size(), shape_index, tw, tl, &m) != LJ_OK)
        Error(filename.string() + ": " + lj_last_error());
    return to_mesh(m);
}

template <>
GridVolume<Real> load_volume_from_file(const fs::path &filename) {
    std::vector<char> bytes = slurp(filename, "load_volume");
    lj_volume_data v;
    if (lj_volume_load(lj_host_ctx(), bytes.data(), bytes.size(), 1, &v) != LJ_OK) Error(lj_last_error() + std::string(" Filename:") + filename.string());
    GridVolume<Real> g{Vector3i{v.resolution[0], v.resolution[1], v.resolution[2]}, Vector3{v.p_min[0], v.p_min[1], v.p_min[2]},
                       Vector3{v.p_max[0], v.p_max[1], v.p_max[2]}, std::vector<Real>(v.data, v.data + v.num_values), v.max_data[0], Real(1)};
    lj_volume_free(&v);
    return g;
}

template <>
GridVolume<Spectrum> load_volume_from_file(const fs::path &filename) {
    std::vector<char> bytes = slurp(filename, "load_volume");
    lj_volume_data v;
    if (lj_volume_load(lj_host_ctx(), bytes.data(), bytes.size(), 3, &v) != LJ_OK) Error(lj_last_error() + std::string(" Filename:") + filename.string());
    std::vector<Spectrum> data(v.num_values / 3);
    for (size_t i = 0; i < data.size(); i++) data[i] = Spectrum{v.data[3 * i], v.data[3 * i + 1], v.data[3 * i + 2]};
    GridVolume<Spectrum> g{Vector3i{v.resolution[0], v.resolution[1], v.resolution[2]}, Vector3{v.p_min[0], v.p_min[1], v.p_min[2]},
                           Vector3{v.p_max[0], v.p_max[1], v.p_max[2]}, data, Spectrum{v.max_data[0], v.max_data[1], v.max_data[2]}, Real(1)};
    lj_volume_free(&v);
    return g;
}
