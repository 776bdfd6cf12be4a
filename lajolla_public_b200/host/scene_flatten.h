// Flattens the reference's `Scene` (std::variant / std::vector based,
// src/scene.h:58-81) into the POD `lj_scene_desc` of include/lajolla_gpu.h.
#pragma once
#include "lajolla_gpu.h"
#include <cstdint>
#include <vector>

struct Scene;

struct FlatScene {
    lj_scene_desc desc;
    std::vector<lj_shape> shapes;
    std::vector<lj_light> lights;
    std::vector<lj_material> materials;
    std::vector<lj_medium> media;
    std::vector<lj_texture> textures;
    std::vector<lj_image> images1, images3;
    std::vector<double> positions, normals, uvs, texels, tables, grid_data;
    std::vector<int32_t> indices;
    void finalize(); // point desc at the vectors
};

void flatten_scene(const Scene &scene, FlatScene &out);
