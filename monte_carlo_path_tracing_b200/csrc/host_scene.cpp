// csrc/host_scene.cpp -- see host_scene.h.  Own OBJ/MTL/XML parsing, reference-shaped interface.
#include "host_scene.h"
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>

namespace mcpt {
namespace {

std::string dir_of(const std::string& path)
{
    size_t p = path.find_last_of("/\\");
    return p == std::string::npos ? std::string() : path.substr(0, p + 1);
}

std::string trim(const std::string& s)
{
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return std::string();
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}

// OBJ index: 1-based, negative = relative to the current count (tinyobj fixIndex)
int fix_index(int idx, int n)
{
    if (idx > 0) return idx - 1;
    if (idx == 0) return -1;
    return n + idx;
}

struct FaceVert { int v, vt, vn; };

// "v", "v/vt", "v//vn", "v/vt/vn"
FaceVert parse_face_vert(const char* tok, int nv, int nvt, int nvn)
{
    FaceVert r = {-1, -1, -1};
    char* end;
    long a = std::strtol(tok, &end, 10);
    r.v = fix_index((int)a, nv);
    if (*end != '/') return r;
    end++;
    if (*end != '/') {
        long b = std::strtol(end, &end, 10);
        r.vt = fix_index((int)b, nvt);
    }
    if (*end == '/') {
        end++;
        long c = std::strtol(end, &end, 10);
        r.vn = fix_index((int)c, nvn);
    }
    return r;
}

void load_mtl(const std::string& path, std::vector<material_t>& mats, std::map<std::string, int>& by_name)
{
    std::ifstream in(path);
    if (!in) throw Error(MCPT_ERR_IO, "cannot open material file " + path);
    std::string line;
    material_t cur;
    bool have = false;
    auto flush = [&]() {
        if (have) {
            by_name[cur.name] = (int)mats.size();  // later duplicates win in the map, like tinyobj
            mats.push_back(cur);
        }
    };
    while (std::getline(in, line)) {
        line = trim(line);
        if (line.empty() || line[0] == '#') continue;
        std::istringstream ss(line);
        std::string key;
        ss >> key;
        if (key == "newmtl") {
            flush();
            cur = material_t();
            have = true;
            std::string rest;
            std::getline(ss, rest);
            cur.name = trim(rest);
        } else if (key == "Kd") {
            double a = 0, b = 0, c = 0; ss >> a >> b >> c;
            cur.diffuse[0] = (float)a; cur.diffuse[1] = (float)b; cur.diffuse[2] = (float)c;
        } else if (key == "Ks") {
            double a = 0, b = 0, c = 0; ss >> a >> b >> c;
            cur.specular[0] = (float)a; cur.specular[1] = (float)b; cur.specular[2] = (float)c;
        } else if (key == "Ns") {
            double a = 1; ss >> a;
            cur.shininess = (float)a;
        }
    }
    flush();
}

double dot(const dvec& a, const dvec& b)
{
    double ans = 0;
    for (int i = 0; i < 3; i++) ans += a[i] * b[i];
    return ans;
}
dvec sub(const dvec& a, const dvec& b) { return {a[0] - b[0], a[1] - b[1], a[2] - b[2]}; }
dvec cross(const dvec& a, const dvec& b)
{
    return {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
}
double norm2(const dvec& a) { return std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]); }
dvec normalized(const dvec& a)
{
    double l = norm2(a);
    return {a[0] / l, a[1] / l, a[2] / l};
}

}  // namespace

// ---- Myobj ---------------------------------------------------------------------------------------
void Myobj::read()
{
    std::ifstream in(inputfile);
    if (!in) throw Error(MCPT_ERR_IO, "cannot open obj file " + inputfile);
    vertices.clear(); normals.clear(); shapes.clear(); materials.clear(); warning.clear();
    std::map<std::string, int> material_map;
    std::string base = dir_of(inputfile);
    shape_t shape;
    int material = -1;
    int ntex = 0;
    std::string line;
    std::vector<FaceVert> fv;
    auto push_tri = [&](const FaceVert& a, const FaceVert& b, const FaceVert& c) {
        const FaceVert* t[3] = {&a, &b, &c};
        for (int k = 0; k < 3; k++) {
            if (t[k]->v < 0 || (size_t)t[k]->v * 3 + 2 >= vertices.size())
                throw Error(MCPT_ERR_IO, inputfile + ": face with invalid vertex index");
            shape.vertex_index.push_back(t[k]->v);
            shape.normal_index.push_back(t[k]->vn);
        }
        shape.material_ids.push_back(material);
    };
    auto flush_shape = [&]() {
        if (shape.num_faces() > 0) shapes.push_back(shape);
        std::string keep;
        shape = shape_t();
    };
    size_t line_no = 0;
    while (std::getline(in, line)) {
        line_no++;
        const char* p = line.c_str();
        p += std::strspn(p, " \t");
        if (*p == 0 || *p == '#' || *p == '\r') continue;
        if (p[0] == 'v' && (p[1] == ' ' || p[1] == '\t')) {
            char* e;
            for (int k = 0; k < 3; k++) { vertices.push_back((float)std::strtod(k ? e : p + 2, &e)); }
        } else if (p[0] == 'v' && p[1] == 'n' && (p[2] == ' ' || p[2] == '\t')) {
            char* e;
            for (int k = 0; k < 3; k++) { normals.push_back((float)std::strtod(k ? e : p + 3, &e)); }
        } else if (p[0] == 'v' && p[1] == 't' && (p[2] == ' ' || p[2] == '\t')) {
            ntex++;
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            fv.clear();
            std::istringstream ss(p + 2);
            std::string tok;
            while (ss >> tok)
                fv.push_back(parse_face_vert(tok.c_str(), (int)(vertices.size() / 3), ntex, (int)(normals.size() / 3)));
            if (fv.size() < 3) continue;
            if (fv.size() == 3) {
                push_tri(fv[0], fv[1], fv[2]);
            } else if (fv.size() == 4) {
                // split along the shorter diagonal, computed in float like tinyobj does
                for (int k = 0; k < 4; k++)
                    if (fv[k].v < 0 || (size_t)fv[k].v * 3 + 2 >= vertices.size())
                        throw Error(MCPT_ERR_IO, inputfile + ": face with invalid vertex index");
                const float* v0 = &vertices[3 * (size_t)fv[0].v]; const float* v1 = &vertices[3 * (size_t)fv[1].v];
                const float* v2 = &vertices[3 * (size_t)fv[2].v]; const float* v3 = &vertices[3 * (size_t)fv[3].v];
                float e02x = v2[0] - v0[0], e02y = v2[1] - v0[1], e02z = v2[2] - v0[2];
                float e13x = v3[0] - v1[0], e13y = v3[1] - v1[1], e13z = v3[2] - v1[2];
                float sqr02 = e02x * e02x + e02y * e02y + e02z * e02z;
                float sqr13 = e13x * e13x + e13y * e13y + e13z * e13z;
                if (sqr02 < sqr13) { push_tri(fv[0], fv[1], fv[2]); push_tri(fv[0], fv[2], fv[3]); }
                else { push_tri(fv[0], fv[1], fv[3]); push_tri(fv[1], fv[2], fv[3]); }
            } else {
                // n-gons: plain fan (tinyobj ear-clips; identical for convex polygons up to the choice of fan apex)
                for (size_t k = 1; k + 1 < fv.size(); k++) push_tri(fv[0], fv[k], fv[k + 1]);
                warning += "polygon with >4 vertices fan-triangulated\n";
            }
        } else if (std::strncmp(p, "usemtl", 6) == 0) {
            std::string name = trim(p + 6);
            auto it = material_map.find(name);
            if (it != material_map.end()) material = it->second;
            else { material = -1; warning += "material [ '" + name + "' ] not found in .mtl\n"; }
        } else if (std::strncmp(p, "mtllib", 6) == 0 && (p[6] == ' ' || p[6] == '\t')) {
            std::istringstream ss(p + 7);
            std::string fn;
            bool ok = false;
            std::string last;
            while (ss >> fn) {
                try { load_mtl(base + fn, materials, material_map); ok = true; break; }
                catch (const Error& e) { last = e.what(); }
            }
            if (!ok) warning += "failed to load material file(s): " + last + "\n";
        } else if ((p[0] == 'g' || p[0] == 'o') && (p[1] == ' ' || p[1] == '\t' || p[1] == 0 || p[1] == '\r')) {
            flush_shape();
            shape.name = trim(p + 1);
        }
    }
    flush_shape();
    if (shapes.empty()) throw Error(MCPT_ERR_IO, inputfile + ": no faces");
}

size_t Myobj::num_prims() const
{
    size_t n = 0;
    for (auto& s : shapes) n += s.num_faces();
    return n;
}

std::array<dvec, 3> Myobj::get_vertexes_of_facet(size_t s, size_t f) const
{
    std::array<dvec, 3> pts;
    for (int i = 0; i < 3; i++) {
        size_t vi = (size_t)shapes[s].vertex_index[3 * f + i];
        pts[i] = {vertices[3 * vi + 0], vertices[3 * vi + 1], vertices[3 * vi + 2]};
    }
    return pts;
}

std::array<dvec, 3> Myobj::get_normals_of_facet(size_t s, size_t f) const
{
    std::array<dvec, 3> ns;
    for (int i = 0; i < 3; i++) {
        int ni = shapes[s].normal_index[3 * f + i];
        // the reference reads attrib.normals[3*(-1)+k] here (Myobj.cpp:671-673): undefined behaviour
        if (ni < 0 || (size_t)ni * 3 + 2 >= normals.size())
            throw Error(MCPT_ERR_IO, inputfile + ": face without a valid vn normal (the reference requires one per face vertex)");
        ns[i] = {normals[3 * (size_t)ni + 0], normals[3 * (size_t)ni + 1], normals[3 * (size_t)ni + 2]};
    }
    return ns;
}

dvec Myobj::get_unique_normal_of_facet(size_t s, size_t f) const
{
    auto v = get_vertexes_of_facet(s, f);
    auto n = get_normals_of_facet(s, f);
    float v9[9], n9[9];
    for (int i = 0; i < 3; i++)
        for (int c = 0; c < 3; c++) { v9[3 * i + c] = (float)v[i][c]; n9[3 * i + c] = (float)n[i][c]; }
    return unique_normal(v9, n9);
}

// Myobj.cpp:680-709: n = normalize((b-a)x(c-a)); keep n iff it agrees better with the sum of the
// normalised vertex normals than -n does; a tie (both sums equal) chooses -n (:702-708).
dvec unique_normal(const float* v9, const float* n9)
{
    dvec a = {v9[0], v9[1], v9[2]}, b = {v9[3], v9[4], v9[5]}, c = {v9[6], v9[7], v9[8]};
    dvec na = normalized({n9[0], n9[1], n9[2]}), nb = normalized({n9[3], n9[4], n9[5]}), nc = normalized({n9[6], n9[7], n9[8]});
    dvec n = normalized(cross(sub(b, a), sub(c, a)));
    dvec nr = {n[0] * -1, n[1] * -1, n[2] * -1};
    double w = dot(n, na) + dot(n, nb) + dot(n, nc);
    double wr = dot(nr, na) + dot(nr, nb) + dot(nr, nc);
    return (w > wr) ? n : nr;
}

// ---- Mylight -------------------------------------------------------------------------------------
namespace {
struct XmlTag {
    std::string name;
    std::map<std::string, std::string> attr;
    bool closing = false, self_closing = false;
};

// minimal tag scanner: elements, attributes in "..." or '...', comments, <? ?> and <! > skipped
bool next_tag(const std::string& s, size_t& pos, XmlTag& tag)
{
    for (;;) {
        size_t lt = s.find('<', pos);
        if (lt == std::string::npos) return false;
        if (s.compare(lt, 4, "<!--") == 0) {
            size_t e = s.find("-->", lt + 4);
            if (e == std::string::npos) return false;
            pos = e + 3;
            continue;
        }
        if (lt + 1 < s.size() && (s[lt + 1] == '?' || s[lt + 1] == '!')) {
            size_t e = s.find('>', lt);
            if (e == std::string::npos) return false;
            pos = e + 1;
            continue;
        }
        size_t i = lt + 1;
        tag = XmlTag();
        if (i < s.size() && s[i] == '/') { tag.closing = true; i++; }
        size_t n0 = i;
        while (i < s.size() && !std::isspace((unsigned char)s[i]) && s[i] != '>' && s[i] != '/') i++;
        tag.name = s.substr(n0, i - n0);
        while (i < s.size() && s[i] != '>') {
            while (i < s.size() && std::isspace((unsigned char)s[i])) i++;
            if (i >= s.size() || s[i] == '>') break;
            if (s[i] == '/') { tag.self_closing = true; i++; continue; }
            size_t k0 = i;
            while (i < s.size() && s[i] != '=' && !std::isspace((unsigned char)s[i]) && s[i] != '>' && s[i] != '/') i++;
            std::string key = s.substr(k0, i - k0);
            while (i < s.size() && std::isspace((unsigned char)s[i])) i++;
            std::string val;
            if (i < s.size() && s[i] == '=') {
                i++;
                while (i < s.size() && std::isspace((unsigned char)s[i])) i++;
                if (i < s.size() && (s[i] == '"' || s[i] == '\'')) {
                    char q = s[i++];
                    size_t v0 = i;
                    while (i < s.size() && s[i] != q) i++;
                    val = s.substr(v0, i - v0);
                    if (i < s.size()) i++;
                }
            }
            if (!key.empty()) tag.attr[key] = val;
        }
        pos = (i < s.size()) ? i + 1 : i;
        return true;
    }
}

// RadianceRGB(std::string) (RadianceRGB.cpp:17-27): three comma-separated stod
dvec parse_radiance(const std::string& str)
{
    std::stringstream ss(str);
    std::string item;
    dvec r{};
    for (int i = 0; i < 3; i++) {
        if (!std::getline(ss, item, ',')) throw Error(MCPT_ERR_IO, "bad radiance string '" + str + "'");
        try { r[i] = std::stod(item); }
        catch (...) { throw Error(MCPT_ERR_IO, "bad radiance string '" + str + "'"); }
    }
    return r;
}

double attr_d(const XmlTag& t, const char* k)
{
    auto it = t.attr.find(k);
    return it == t.attr.end() ? 0.0 : std::atof(it->second.c_str());
}
}  // namespace

void Mylight::read()
{
    std::ifstream in(inputfile);
    if (!in) throw Error(MCPT_ERR_IO, "cannot open light xml file " + inputfile);
    std::stringstream buf;
    buf << in.rdbuf();
    std::string s = buf.str();
    lightsRadiance.clear();
    camera = CameraDesc();
    size_t pos = 0;
    int depth = 0;
    bool in_camera = false;
    int camera_depth = 0;
    XmlTag t;
    while (next_tag(s, pos, t)) {
        if (t.closing) {
            depth--;
            if (in_camera && depth == camera_depth) in_camera = false;
            continue;
        }
        if (depth == 0 && t.name == "light") {  // doc.children("light"), Mylight.cpp:21
            std::string name = t.attr.count("mtlname") ? t.attr["mtlname"] : std::string();
            std::string rad = t.attr.count("radiance") ? t.attr["radiance"] : std::string();
            lightsRadiance[name] = parse_radiance(rad);
        } else if (depth == 0 && t.name == "camera") {
            camera.present = true;
            camera.width = (int)attr_d(t, "width");
            camera.height = (int)attr_d(t, "height");
            camera.fovy = attr_d(t, "fovy");
            if (!t.self_closing) { in_camera = true; camera_depth = depth; }
        } else if (in_camera && (t.name == "eye" || t.name == "lookat" || t.name == "up")) {
            dvec v = {attr_d(t, "x"), attr_d(t, "y"), attr_d(t, "z")};
            if (t.name == "eye") camera.eye = v;
            else if (t.name == "lookat") camera.lookat = v;
            else camera.up = v;
        }
        if (!t.self_closing) depth++;
    }
}

void Mylight::gather_light_triangles(const Myobj& obj)
{
    lightsTriangles.clear();
    islight.clear();
    for (size_t s = 0; s < obj.shapes.size(); s++)
        for (size_t f = 0; f < obj.shapes[s].num_faces(); f++) {
            int mid = obj.shapes[s].material_ids[f];
            if (mid < 0 || (size_t)mid >= obj.materials.size())  // materials.at() throws in the reference (Mylight.cpp:51)
                throw Error(MCPT_ERR_IO, obj.inputfile + ": face without a material (usemtl) -- the reference requires one");
            for (auto it = lightsRadiance.begin(); it != lightsRadiance.end(); ++it) {
                if (obj.materials[(size_t)mid].name == it->first) {
                    islight.insert_or_assign(triangle(s, f), it->second);
                    auto abc = obj.get_vertexes_of_facet(s, f);
                    dvec n = cross(sub(abc[1], abc[0]), sub(abc[2], abc[0]));  // Mylight.cpp:66-69
                    double inv = 1.0 / norm2(n);
                    n = {n[0] * inv, n[1] * inv, n[2] * inv};
                    double area = 0.5 * dot(cross(sub(abc[1], abc[0]), sub(abc[2], abc[0])), n);
                    lightsTriangles[it->first].push_back(lightTriangle{s, f, area});
                    break;
                }
            }
        }
}

// ---- flatten -------------------------------------------------------------------------------------
McptSceneDesc FlatScene::desc() const
{
    McptSceneDesc d;
    std::memset(&d, 0, sizeof d);
    d.num_prims = (uint32_t)material_id.size();
    d.num_shapes = (uint32_t)shape_offset.size() - 1;
    d.num_materials = (uint32_t)mat_ns.size();
    d.num_lights = (uint32_t)light_prim.size();
    d.num_classes = (uint32_t)(class_radiance.size() / 3);
    d.positions = positions.data(); d.normals = normals.data(); d.material_id = material_id.data();
    d.shape_offset = shape_offset.data(); d.mat_kd = mat_kd.data(); d.mat_ks = mat_ks.data(); d.mat_ns = mat_ns.data();
    d.light_prim = light_prim.data(); d.light_class = light_class.data(); d.class_radiance = class_radiance.data();
    return d;
}

FlatScene flatten(const Myobj& obj, const Mylight& lights)
{
    FlatScene fs;
    fs.shape_offset.push_back(0);
    for (size_t s = 0; s < obj.shapes.size(); s++) {
        for (size_t f = 0; f < obj.shapes[s].num_faces(); f++) {
            auto v = obj.get_vertexes_of_facet(s, f);
            auto n = obj.get_normals_of_facet(s, f);
            for (int k = 0; k < 3; k++)
                for (int c = 0; c < 3; c++) {
                    fs.positions.push_back((float)v[k][c]);
                    fs.normals.push_back((float)n[k][c]);
                }
            int mid = obj.shapes[s].material_ids[f];
            if (mid < 0 || (size_t)mid >= obj.materials.size())
                throw Error(MCPT_ERR_IO, obj.inputfile + ": face without a material (usemtl)");
            fs.material_id.push_back(mid);
        }
        fs.shape_offset.push_back((uint32_t)fs.material_id.size());
    }
    for (auto& m : obj.materials) {
        for (int c = 0; c < 3; c++) { fs.mat_kd.push_back(m.diffuse[c]); fs.mat_ks.push_back(m.specular[c]); }
        fs.mat_ns.push_back(m.shininess);
    }
    int ci = 0;
    for (auto& kv : lights.lightsRadiance) {  // std::map order == the reference's iteration order
        for (int c = 0; c < 3; c++) fs.class_radiance.push_back(kv.second[c]);
        auto it = lights.lightsTriangles.find(kv.first);
        if (it != lights.lightsTriangles.end())
            for (auto& lt : it->second) {
                fs.light_prim.push_back((int32_t)(fs.shape_offset[lt.s] + lt.f));
                fs.light_class.push_back(ci);
            }
        ci++;
    }
    return fs;
}

}  // namespace mcpt
