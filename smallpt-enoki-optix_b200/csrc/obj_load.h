// obj_load.h — OBJ/MTL ingestion (host only).  See obj_load.cpp.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

namespace wpt {

// the eight results of the reference's load_meshes (src/main.cpp:133-141), AoS-flattened
struct ObjMesh {
    std::vector<float>   positions, normals, texcoords;                 // 3V, 3Vn, 2Vt
    std::vector<int32_t> position_triplets, normal_triplets, texcoord_triplets;   // 3T each, -1 = absent
    std::vector<int32_t> material_id;                                    // T, tinyobj id + 1 (0 = default)
    std::vector<std::string> material_names;                             // M (without the default)
    std::vector<float>   kd;                                             // 3M
};

bool load_obj(const char *path, ObjMesh &mesh, std::string &err);

}  // namespace wpt
