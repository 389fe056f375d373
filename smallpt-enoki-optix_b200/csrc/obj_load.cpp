// obj_load.cpp — Wavefront OBJ / MTL ingestion into the flat arrays wpt_scene_create takes.
//
// Replaces load_meshes (src/main.cpp:133-251), which drives tinyobj::ObjReader with
// triangulate = true, vertex_color = false (src/main.cpp:144-148) and then flattens all shapes:
//   * tri_id = global face index, shapes in order of appearance, faces in file order
//     (src/main.cpp:165-188);
//   * per-face material id = tinyobj id + 1, id 0 = the default material (src/main.cpp:185,229-234);
//   * every material is forced to white (src/main.cpp:244); the commented-out intent, the Kd colour
//     (src/main.cpp:239-242), is available with use_kd != 0.
// The parser is written from scratch but follows tinyobjloader's published rules where they decide
// the OUTPUT (dependencies/tinyobjloader/tiny_obj_loader.h): 1-based and negative-relative indices
// (fixIndex, :741-760), the i, i/j, i//k, i/j/k corner forms (:1071-1121), its decimal-to-double
// routine (:807-938: digit accumulation in double, 5^e * 2^e exponent assembly) so that positions
// come out bit-identical, its polygon triangulation (:1364-1560: projection axes from the first
// non-degenerate corner, signed area, ear clipping with a rotating guess vertex), `usemtl` lookup
// with first-definition-wins names (:2361-2384), and `mtllib` with several candidate files (:2389-2437).
// Checked against the real tinyobjloader by tests/test_obj_loader.py (fixtures in tests/golden/obj/).
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <map>
#include <string>
#include <vector>

#include "obj_load.h"

namespace wpt {
namespace {

inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline bool is_digit(char c) { return (unsigned)(c - '0') < 10u; }
inline bool is_eol(char c) { return c == '\r' || c == '\n' || c == '\0'; }

// decimal text [s, e) -> double the way tiny_obj_loader.h:807-938 does it (NOT strtod: the digit
// accumulation rounds differently in rare cases, and parity of the vertex data is the point)
bool parse_double(const char *s, const char *e, double *out)
{
    if (s >= e) return false;
    double m = 0.0;
    int exp10 = 0, nread = 0;
    bool neg = false, eneg = false, lead_dot = false;
    const char *c = s;
    if (*c == '+' || *c == '-') {
        neg = *c == '-';
        c++;
        if (c != e && *c == '.') lead_dot = true;
    } else if (is_digit(*c)) {
    } else if (*c == '.') {
        lead_dot = true;
    } else {
        return false;
    }
    if (!lead_dot) {
        while (c != e && is_digit(*c)) { m = m * 10 + (*c - '0'); c++; nread++; }
        if (nread == 0) return false;
    }
    if (c != e && *c == '.') {
        c++;
        int k = 1;
        static const double frac[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
        while (c != e && is_digit(*c)) {
            m += (*c - '0') * (k < 8 ? frac[k] : std::pow(10.0, -k));
            k++; c++;
        }
    } else if (c != e && (*c == 'e' || *c == 'E')) {
    } else {
        c = e;   // anything else ends the number
    }
    if (c != e && (*c == 'e' || *c == 'E')) {
        c++;
        if (c != e && (*c == '+' || *c == '-')) { eneg = *c == '-'; c++; }
        else if (c != e && is_digit(*c)) {}
        else return false;
        int k = 0;
        while (c != e && is_digit(*c)) { exp10 = exp10 * 10 + (*c - '0'); c++; k++; }
        if (eneg) exp10 = -exp10;
        if (k == 0) return false;
    }
    const double v = exp10 ? std::ldexp(m * std::pow(5.0, exp10), exp10) : m;
    *out = neg ? -v : v;
    return true;
}

float parse_real(const char *&p, double dflt = 0.0)
{
    p += std::strspn(p, " \t");
    const char *end = p + std::strcspn(p, " \t\r");
    double v = dflt;
    parse_double(p, end, &v);
    p = end;
    return (float)v;
}

std::string parse_word(const char *&p)
{
    p += std::strspn(p, " \t");
    const size_t n = std::strcspn(p, " \t\r");
    std::string s(p, n);
    p += n;
    return s;
}

struct Corner { int v = -1, vt = -1, vn = -1; };

bool fix_index(int idx, int n, int *out)
{
    if (idx > 0) { *out = idx - 1; return true; }
    if (idx == 0) return false;
    *out = n + idx;
    return true;
}

bool parse_corner(const char *&p, int nv, int nvn, int nvt, Corner *out)
{
    Corner c;
    if (!fix_index(std::atoi(p), nv, &c.v)) return false;
    p += std::strcspn(p, "/ \t\r");
    if (*p != '/') { *out = c; return true; }
    p++;
    if (*p == '/') {   // i//k
        p++;
        if (!fix_index(std::atoi(p), nvn, &c.vn)) return false;
        p += std::strcspn(p, "/ \t\r");
        *out = c;
        return true;
    }
    if (!fix_index(std::atoi(p), nvt, &c.vt)) return false;
    p += std::strcspn(p, "/ \t\r");
    if (*p != '/') { *out = c; return true; }
    p++;
    if (!fix_index(std::atoi(p), nvn, &c.vn)) return false;
    p += std::strcspn(p, "/ \t\r");
    *out = c;
    return true;
}

// crossing-number point-in-polygon test on a triangle (the role pnpoly plays in the ear test)
bool inside_tri(const float x[3], const float y[3], float tx, float ty)
{
    bool in = false;
    for (int i = 0, j = 2; i < 3; j = i++)
        if (((y[i] > ty) != (y[j] > ty)) && (tx < (x[j] - x[i]) * (ty - y[i]) / (y[j] - y[i]) + x[i])) in = !in;
    return in;
}

struct Loader {
    ObjMesh &m;
    std::vector<float> &v;
    std::map<std::string, int> mat_by_name;
    std::string dir;
    int cur_mat = -1;

    explicit Loader(ObjMesh &mesh) : m(mesh), v(mesh.positions) {}

    void emit(const Corner &a, const Corner &b, const Corner &c)
    {
        const Corner t[3] = {a, b, c};
        for (const Corner &k : t) {
            m.position_triplets.push_back(k.v);
            m.normal_triplets.push_back(k.vn);
            m.texcoord_triplets.push_back(k.vt);
        }
        m.material_id.push_back(cur_mat + 1);   // src/main.cpp:185
    }

    bool coord(int vi, int axis, float *out) const
    {
        const size_t k = (size_t)vi * 3 + (size_t)axis;
        if (vi < 0 || k >= v.size()) return false;
        *out = v[k];
        return true;
    }

    // one `f` record -> triangles (tiny_obj_loader.h:1364-1560)
    void polygon(const std::vector<Corner> &f)
    {
        size_t n = f.size();
        if (n < 3) return;
        if (n == 3) { emit(f[0], f[1], f[2]); return; }
        // projection plane: drop the dominant axis of the first non-degenerate corner
        int ax0 = 1, ax1 = 2;
        for (size_t k = 0; k < n; k++) {
            const int i0 = f[k % n].v, i1 = f[(k + 1) % n].v, i2 = f[(k + 2) % n].v;
            float p0[3], p1[3], p2[3];
            bool ok = true;
            for (int a = 0; a < 3; a++) ok = ok && coord(i0, a, &p0[a]) && coord(i1, a, &p1[a]) && coord(i2, a, &p2[a]);
            if (!ok) continue;
            const float e0x = p1[0] - p0[0], e0y = p1[1] - p0[1], e0z = p1[2] - p0[2];
            const float e1x = p2[0] - p1[0], e1y = p2[1] - p1[1], e1z = p2[2] - p1[2];
            const float cx = std::fabs(e0y * e1z - e0z * e1y), cy = std::fabs(e0z * e1x - e0x * e1z),
                        cz = std::fabs(e0x * e1y - e0y * e1x);
            const float eps = std::numeric_limits<float>::epsilon();
            if (cx > eps || cy > eps || cz > eps) {
                if (!(cx > cy && cx > cz)) {
                    ax0 = 0;
                    if (cz > cx && cz > cy) ax1 = 1;
                }
                break;
            }
        }
        float area = 0;
        for (size_t k = 0; k < n; k++) {
            float x0, y0, x1, y1;
            if (!coord(f[k].v, ax0, &x0) || !coord(f[k].v, ax1, &y0) || !coord(f[(k + 1) % n].v, ax0, &x1) ||
                !coord(f[(k + 1) % n].v, ax1, &y1))
                continue;
            area += (x0 * y1 - y0 * x1) * 0.5f;
        }
        std::vector<Corner> rest = f;
        size_t guess = 0, budget = f.size(), prev = rest.size();
        while (rest.size() > 3 && budget > 0) {
            n = rest.size();
            if (guess >= n) guess -= n;
            if (prev != n) { prev = n; budget = n; } else { budget--; }
            Corner c[3];
            float x[3], y[3];
            for (int k = 0; k < 3; k++) {
                c[k] = rest[(guess + k) % n];
                if (!coord(c[k].v, ax0, &x[k]) || !coord(c[k].v, ax1, &y[k])) { x[k] = 0; y[k] = 0; }
            }
            const float cr = (x[1] - x[0]) * (y[2] - y[1]) - (y[1] - y[0]) * (x[2] - x[1]);
            if (cr * area < 0.0f) { guess++; continue; }   // reflex corner
            bool blocked = false;
            for (size_t o = 3; o < n && !blocked; o++) {
                float tx, ty;
                const size_t idx = (guess + o) % n;
                if (!coord(rest[idx].v, ax0, &tx) || !coord(rest[idx].v, ax1, &ty)) continue;
                blocked = inside_tri(x, y, tx, ty);
            }
            if (blocked) { guess++; continue; }
            emit(c[0], c[1], c[2]);
            rest.erase(rest.begin() + (long)((guess + 1) % n));
        }
        if (rest.size() == 3) emit(rest[0], rest[1], rest[2]);
    }

    // MTL: only what the path uses — material names and Kd (src/main.cpp:239-241)
    bool load_mtl(const std::string &name)
    {
        const std::string path = dir.empty() ? name : dir + "/" + name;
        FILE *fp = std::fopen(path.c_str(), "rb");
        if (!fp) return false;
        std::string text;
        char buf[65536];
        size_t got;
        while ((got = std::fread(buf, 1, sizeof buf, fp)) > 0) text.append(buf, got);
        std::fclose(fp);
        bool open = false;
        float kd[3] = {0, 0, 0};   // tinyobj InitMaterial: diffuse = 0
        std::string cur;
        auto flush = [&]() {
            if (!open) return;
            mat_by_name.insert(std::make_pair(cur, (int)m.material_names.size()));   // first definition wins
            m.material_names.push_back(cur);
            m.kd.insert(m.kd.end(), kd, kd + 3);
        };
        size_t pos = 0;
        while (pos < text.size()) {
            size_t eol = text.find('\n', pos);
            if (eol == std::string::npos) eol = text.size();
            std::string line = text.substr(pos, eol - pos);
            pos = eol + 1;
            while (!line.empty() && (line.back() == '\r' || line.back() == '\n')) line.pop_back();
            const char *p = line.c_str();
            p += std::strspn(p, " \t");
            if (*p == '\0' || *p == '#') continue;
            if (std::strncmp(p, "newmtl", 6) == 0 && is_space(p[6])) {
                flush();
                p += 7;
                // the name is the rest of the line, trailing blanks dropped
                p += std::strspn(p, " \t");
                std::string nm(p);
                while (!nm.empty() && is_space(nm.back())) nm.pop_back();
                cur = nm;
                kd[0] = kd[1] = kd[2] = 0;
                open = true;
                continue;
            }
            if (p[0] == 'K' && p[1] == 'd' && is_space(p[2])) {
                p += 2;
                kd[0] = parse_real(p); kd[1] = parse_real(p); kd[2] = parse_real(p);
                continue;
            }
        }
        flush();
        return true;
    }
};

}  // namespace

bool load_obj(const char *path, ObjMesh &mesh, std::string &err)
{
    FILE *fp = std::fopen(path, "rb");
    if (!fp) { err = std::string("cannot open ") + path; return false; }
    std::string text;
    {
        char buf[1 << 16];
        size_t got;
        while ((got = std::fread(buf, 1, sizeof buf, fp)) > 0) text.append(buf, got);
    }
    std::fclose(fp);
    mesh = ObjMesh();
    Loader L(mesh);
    {
        const std::string sp(path);
        const size_t slash = sp.find_last_of("/\\");
        L.dir = slash == std::string::npos ? std::string() : sp.substr(0, slash);
    }
    std::vector<Corner> face;
    size_t pos = 0, line_no = 0;
    while (pos < text.size()) {
        size_t eol = text.find('\n', pos);
        if (eol == std::string::npos) eol = text.size();
        std::string line = text.substr(pos, eol - pos);
        pos = eol + 1;
        line_no++;
        while (!line.empty() && (line.back() == '\n' || line.back() == '\r')) line.pop_back();
        const char *p = line.c_str();
        p += std::strspn(p, " \t");
        if (*p == '\0' || *p == '#') continue;
        if (p[0] == 'v' && is_space(p[1])) {
            p += 2;
            for (int k = 0; k < 3; k++) mesh.positions.push_back(parse_real(p));
            continue;
        }
        if (p[0] == 'v' && p[1] == 'n' && is_space(p[2])) {
            p += 3;
            for (int k = 0; k < 3; k++) mesh.normals.push_back(parse_real(p));
            continue;
        }
        if (p[0] == 'v' && p[1] == 't' && is_space(p[2])) {
            p += 3;
            for (int k = 0; k < 2; k++) mesh.texcoords.push_back(parse_real(p));
            continue;
        }
        if (p[0] == 'f' && is_space(p[1])) {
            p += 2;
            p += std::strspn(p, " \t");
            face.clear();
            while (!is_eol(*p)) {
                Corner c;
                if (!parse_corner(p, (int)(mesh.positions.size() / 3), (int)(mesh.normals.size() / 3),
                                  (int)(mesh.texcoords.size() / 2), &c)) {
                    err = "bad face index at line " + std::to_string(line_no);
                    return false;
                }
                face.push_back(c);
                p += std::strspn(p, " \t\r");
            }
            L.polygon(face);
            continue;
        }
        if (std::strncmp(p, "usemtl", 6) == 0) {
            p += 6;
            const std::string nm = parse_word(p);
            const auto it = L.mat_by_name.find(nm);
            L.cur_mat = it == L.mat_by_name.end() ? -1 : it->second;   // unknown name -> default material
            continue;
        }
        if (std::strncmp(p, "mtllib", 6) == 0 && is_space(p[6])) {
            p += 7;
            // several candidate files separated by blanks: the first one that opens is used
            std::string rest(p);
            size_t a = 0;
            while (a <= rest.size()) {
                size_t b = rest.find(' ', a);
                if (b == std::string::npos) b = rest.size();
                std::string nm = rest.substr(a, b - a);
                while (!nm.empty() && (nm.back() == '\r' || nm.back() == '\t')) nm.pop_back();
                if (!nm.empty() && L.load_mtl(nm)) break;
                a = b + 1;
            }
            continue;
        }
        // g / o / s / l / p / t and unknown records do not change the flattened output
    }
    // the reference would gather out of bounds on a bad index; reject positions,
    // keep -1 normals/texcoords (faces without them; wpt_scene_create falls back to geometric normals)
    const int32_t nv = (int32_t)(mesh.positions.size() / 3), nn = (int32_t)(mesh.normals.size() / 3);
    for (int32_t i : mesh.position_triplets)
        if (i < 0 || i >= nv) { err = "face references a vertex that does not exist"; return false; }
    for (int32_t &i : mesh.normal_triplets)
        if (i >= nn || i < -1) i = -1;
    return true;
}

}  // namespace wpt
