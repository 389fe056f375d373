// tinyobj_ref_wrap.cpp — TEST INFRASTRUCTURE ONLY.  Compiles the reference tree's own tinyobjloader
// (dependencies/tinyobjloader/tiny_obj_loader.h, included from where it lies under /root/reference;
// nothing is copied) and flattens its result exactly like the reference's load_meshes does
// (src/main.cpp:143-223: triangulate = true, vertex_color = false, shapes in order, material id + 1).
// Used to pin smallpt-enoki-optix_b200/csrc/obj_load.cpp and to generate tests/golden/obj/*.npz.
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"

#include <cstdint>
#include <cstring>
#include <vector>

struct RefObj {
    std::vector<float> pos, nrm, uv, kd;
    std::vector<int32_t> pidx, nidx, tidx, mat;
};

extern "C" {
// returns NULL on parse failure
void *tor_load(const char *path)
{
    tinyobj::ObjReader reader;
    tinyobj::ObjReaderConfig cfg;
    cfg.triangulate = true;       // src/main.cpp:146
    cfg.vertex_color = false;     // src/main.cpp:147
    if (!reader.ParseFromFile(path, cfg)) return nullptr;
    RefObj *r = new RefObj();
    for (const tinyobj::shape_t &s : reader.GetShapes()) {                  // src/main.cpp:165-188
        const size_t nf = s.mesh.num_face_vertices.size();
        for (size_t f = 0; f < nf; f++) {
            for (int k = 0; k < 3; k++) {
                const tinyobj::index_t &i = s.mesh.indices[3 * f + k];
                r->pidx.push_back(i.vertex_index);
                r->nidx.push_back(i.normal_index);
                r->tidx.push_back(i.texcoord_index);
            }
            r->mat.push_back(s.mesh.material_ids[f] + 1);
        }
    }
    const tinyobj::attrib_t &a = reader.GetAttrib();
    r->pos.assign(a.vertices.begin(), a.vertices.end());
    r->nrm.assign(a.normals.begin(), a.normals.end());
    r->uv.assign(a.texcoords.begin(), a.texcoords.end());
    for (const tinyobj::material_t &m : reader.GetMaterials())              // src/main.cpp:237-241
        for (int k = 0; k < 3; k++) r->kd.push_back(m.diffuse[k]);
    return r;
}
void tor_counts(const void *h, uint32_t out[5])
{
    const RefObj *r = (const RefObj *)h;
    out[0] = (uint32_t)(r->pos.size() / 3); out[1] = (uint32_t)(r->nrm.size() / 3);
    out[2] = (uint32_t)(r->uv.size() / 2);  out[3] = (uint32_t)r->mat.size();
    out[4] = (uint32_t)(r->kd.size() / 3);
}
void tor_copy(const void *h, float *pos, float *nrm, float *uv, float *kd, int32_t *pidx, int32_t *nidx,
              int32_t *tidx, int32_t *mat)
{
    const RefObj *r = (const RefObj *)h;
    auto cp = [](auto *dst, const auto &v) { if (dst && !v.empty()) std::memcpy(dst, v.data(), v.size() * sizeof(v[0])); };
    cp(pos, r->pos); cp(nrm, r->nrm); cp(uv, r->uv); cp(kd, r->kd);
    cp(pidx, r->pidx); cp(nidx, r->nidx); cp(tidx, r->tidx); cp(mat, r->mat);
}
void tor_free(void *h) { delete (RefObj *)h; }
}
