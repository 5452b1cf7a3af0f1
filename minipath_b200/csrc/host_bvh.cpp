// Host-side scene loading: OBJ → indexed triangles → compressed 8-wide BVH in the reference node format,
// flattened for the device.  This is the host half that "stays in Rust" in the reference
// (src/scene/triangle_bvh/building.rs); no Rust toolchain exists in the build image, so the host side is C++.
// It is not a GPU kernel and not part of the timed path (the reference bench excludes scene build,
// benches/render_teapot.rs:25-27).
//
// Compile with -ffp-contract=off: fused operations appear only where written as std::fmaf.
#include <algorithm>
#include <array>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>
#include <thread>
#include <tuple>
#include <vector>

#include "host_common.h"

namespace mp {

std::string& last_error_ref() {
    static thread_local std::string s;
    return s;
}
int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    last_error_ref() = buf;
    return code;
}

struct VertexData {  // building.rs:220-224
    V3 pos, tex, normal;
};
struct Tri {
    uint32_t v[3];
    uint32_t material;  // index into the material table (0 when there is none)
    V3 c;           // centroid and bounds, filled once by build_flat: the builders sort triangles many times and would
    V3 bmin, bmax;  // otherwise gather the three vertices (a cache miss each) at every level
};

// ---------------------------------------------------------------------------------------------------
// OBJ loading (building.rs:36-81 on top of the `obj` crate, which keeps polygons of any arity)
// ---------------------------------------------------------------------------------------------------
static bool parse_index_tuple(const std::string& tok, long nv, long nvt, long nvn, long out[3]) {
    // v, v/vt, v//vn, v/vt/vn ; 1-based, negative = relative to the end
    out[0] = out[1] = out[2] = -1;
    size_t pos = 0;
    for (int k = 0; k < 3 && pos <= tok.size(); ++k) {
        size_t slash = tok.find('/', pos);
        std::string part = tok.substr(pos, slash == std::string::npos ? std::string::npos : slash - pos);
        if (!part.empty()) {
            char* end = nullptr;
            long idx = std::strtol(part.c_str(), &end, 10);
            if (end == part.c_str()) return false;
            long count = k == 0 ? nv : (k == 1 ? nvt : nvn);
            idx = idx < 0 ? count + idx : idx - 1;
            if (idx < 0 || idx >= count) return false;
            out[k] = idx;
        }
        if (slash == std::string::npos) break;
        pos = slash + 1;
    }
    return out[0] >= 0;
}

// `newmtl` / `Kd` / `map_Kd` of a .mtl library, appended to `materials` in file order.  A material with a diffuse map gets the
// checker flag: the map images are not part of the data set, the interpolated texture coordinate drives a procedural pattern.
static int load_mtl(const std::string& path, std::vector<MpMaterial>& materials, std::map<std::string, uint32_t>& id_of) {
    std::ifstream f(path);
    if (!f) return fail(MP_EIO, "Failed to read file: %s", path.c_str());
    std::string line;
    MpMaterial* cur = nullptr;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string tag;
        if (!(ss >> tag)) continue;
        if (tag == "newmtl") {
            std::string name;
            ss >> name;
            MpMaterial m{};
            m.albedo[0] = m.albedo[1] = m.albedo[2] = MP_DEFAULT_ALBEDO;
            m.checker_scale = MP_DEFAULT_CHECKER_SCALE;
            id_of[name] = (uint32_t)materials.size();
            materials.push_back(m);
            cur = &materials.back();
        } else if (tag == "Kd" && cur) {
            double r = 0, g = 0, b = 0;
            ss >> r >> g >> b;
            cur->albedo[0] = (float)r, cur->albedo[1] = (float)g, cur->albedo[2] = (float)b;
        } else if (tag == "map_Kd" && cur) {
            cur->flags |= MP_MAT_CHECKER;
        }
    }
    return MP_OK;
}

static int load_obj(const char* path, int quad_mode, std::vector<Tri>& triangles, std::vector<VertexData>& vertices, std::vector<MpMaterial>& materials) {
    std::ifstream f(path);
    if (!f) return fail(MP_EIO, "Failed to read file: %s", path);
    std::map<std::string, uint32_t> material_id;
    uint32_t current_material = 0;
    std::vector<std::array<double, 3>> P, N;
    std::vector<std::array<double, 2>> T;
    std::map<std::tuple<long, long, long>, uint32_t> index_of;  // IndexMap: id = insertion index (building.rs:49-50)
    size_t skipped = 0;
    std::string line;
    while (std::getline(f, line)) {
        std::istringstream ss(line);
        std::string tag;
        if (!(ss >> tag)) continue;
        if (tag == "v") {
            std::array<double, 3> p{0, 0, 0};
            ss >> p[0] >> p[1] >> p[2];
            P.push_back(p);
        } else if (tag == "vn") {
            std::array<double, 3> p{0, 0, 0};
            ss >> p[0] >> p[1] >> p[2];
            N.push_back(p);
        } else if (tag == "vt") {
            std::array<double, 2> p{0, 0};
            ss >> p[0] >> p[1];
            T.push_back(p);
        } else if (tag == "mtllib") {
            if (materials.empty()) {  // entry 0: faces that precede every `usemtl`
                MpMaterial m{};
                m.albedo[0] = m.albedo[1] = m.albedo[2] = MP_DEFAULT_ALBEDO;
                m.checker_scale = MP_DEFAULT_CHECKER_SCALE;
                materials.push_back(m);
            }
            std::string name, dir(path);
            const size_t slash = dir.find_last_of('/');
            dir = slash == std::string::npos ? std::string() : dir.substr(0, slash + 1);
            while (ss >> name)
                if (int rc = load_mtl(name[0] == '/' ? name : dir + name, materials, material_id)) return rc;
        } else if (tag == "usemtl") {
            std::string name;
            ss >> name;
            auto it = material_id.find(name);
            if (it == material_id.end()) return fail(MP_EIO, "Failed to parse file: %s: unknown material '%s'", path, name.c_str());
            current_material = it->second;
        } else if (tag == "f") {
            std::vector<uint32_t> poly;
            std::string tok;
            bool ok = true;
            std::vector<std::tuple<long, long, long>> tuples;
            while (ss >> tok) {
                long idx[3];
                if (!parse_index_tuple(tok, (long)P.size(), (long)T.size(), (long)N.size(), idx)) {
                    ok = false;
                    break;
                }
                tuples.emplace_back(idx[0], idx[1], idx[2]);
            }
            if (!ok) return fail(MP_EIO, "Failed to parse file: %s: bad face '%s'", path, line.c_str());
            if (tuples.size() != 3 && !(quad_mode == 1 && tuples.size() > 3)) {
                ++skipped;  // "non-triangle primitive!" (building.rs:43-46)
                continue;
            }
            for (auto& t : tuples) {
                auto it = index_of.find(t);
                uint32_t id;
                if (it == index_of.end()) {
                    id = (uint32_t)vertices.size();
                    index_of.emplace(t, id);
                    VertexData vd;
                    auto& p = P[std::get<0>(t)];
                    vd.pos = v3((float)p[0], (float)p[1], (float)p[2]);  // f64 → f32 (building.rs:56)
                    vd.tex = v3(0, 0, 0);
                    if (std::get<1>(t) >= 0) vd.tex = v3((float)T[std::get<1>(t)][0], (float)T[std::get<1>(t)][1], 0.0f);
                    vd.normal = v3(0, 0, 0);
                    if (std::get<2>(t) >= 0) {
                        auto& n = N[std::get<2>(t)];
                        V3 nn = v3((float)n[0], (float)n[1], (float)n[2]);
                        // reference normalises unconditionally (building.rs:60-63); a present but zero-length
                        // normal would become NaN there — keep it zero so the triangle is flagged flat instead
                        if (dot(nn, nn) != 0.0f) nn = normalize(nn);
                        vd.normal = nn;
                    }
                    vertices.push_back(vd);
                } else {
                    id = it->second;
                }
                poly.push_back(id);
            }
            for (size_t k = 2; k < poly.size(); ++k) triangles.push_back(Tri{{poly[0], poly[k - 1], poly[k]}, current_material, V3{}, V3{}, V3{}});
        }
    }
    if (skipped) fprintf(stderr, "minipath_b200: %zu non-triangle primitive(s) skipped in %s\n", skipped, path);
    return MP_OK;
}

}  // namespace mp

#include "host_bvh_build.inc"
