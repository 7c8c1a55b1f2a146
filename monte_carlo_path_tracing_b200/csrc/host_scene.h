// csrc/host_scene.h -- host-side mirror of the reference's scene interface.
//
// Same names, argument meaning and data the render loop reads from `Myobj veach` / `Mylight lights`
// (main.cpp:23-24): Myobj::{read, get_vertexes_of_facet, get_normals_of_facet,
// get_unique_normal_of_facet} (Myobj.h:92-117) and Mylight::{read, gather_light_triangles,
// lightsRadiance, lightsTriangles, islight} (Mylight.h:106-127).  The parsing is our own (the
// reference uses the vendored tiny_obj_loader / pugixml, which are not part of this tree): the OBJ/MTL
// subset tinyobj turns into attrib/shapes/materials for triangle meshes, and the XML subset
// Mylight::read iterates (document-level <light mtlname radiance/>) plus the <camera> node of
// README.md:339-343 that the reference hard-codes instead (main.cpp:507-517, F7).
// Unlike the reference (exit(1), Myobj.cpp:15-20) failures throw mcpt::Error; the C ABI maps it to a code.
#pragma once
#include <array>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/mcpt.h"

namespace mcpt {

struct Error : std::runtime_error {
    int code;
    Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

typedef std::array<double, 3> dvec;

// (s,f) key, Myobj.h:44-66
struct triangle {
    size_t s, f;
    triangle(size_t s_, size_t f_) : s(s_), f(f_) {}
    bool operator<(const triangle& b) const { return s != b.s ? s < b.s : f < b.f; }
    bool operator==(const triangle& b) const { return s == b.s && f == b.f; }
};

struct material_t {  // the fields of tinyobj::material_t the hot path reads (main.cpp:292,313)
    std::string name;
    float diffuse[3] = {0, 0, 0};
    float specular[3] = {0, 0, 0};
    float shininess = 1.0f;  // tinyobj InitMaterial default
};

struct shape_t {
    std::string name;
    std::vector<int> vertex_index;  // 3 per face
    std::vector<int> normal_index;  // 3 per face, -1 if the face had no vn
    std::vector<int> material_ids;  // per face
    size_t num_faces() const { return material_ids.size(); }
};

class Myobj {
public:
    std::string inputfile;
    std::vector<float> vertices;  // attrib.vertices (real_t = float, F2)
    std::vector<float> normals;   // attrib.normals
    std::vector<shape_t> shapes;
    std::vector<material_t> materials;
    std::string warning;

    explicit Myobj(std::string inputfile_) : inputfile(std::move(inputfile_)) {}
    void read();  // Myobj.cpp:10-28
    std::array<dvec, 3> get_vertexes_of_facet(size_t s, size_t f) const;  // Myobj.cpp:626-650
    std::array<dvec, 3> get_normals_of_facet(size_t s, size_t f) const;   // Myobj.cpp:653-678
    dvec get_unique_normal_of_facet(size_t s, size_t f) const;            // Myobj.cpp:680-709
    size_t num_prims() const;
};

struct lightTriangle {  // Mylight.h:10-24
    size_t s, f;
    double area;
};

struct CameraDesc {
    bool present = false;
    int width = 0, height = 0;
    double fovy = 0;
    dvec eye{}, lookat{}, up{};
};

class Mylight {
public:
    std::string inputfile;
    std::map<std::string, dvec> lightsRadiance;                        // Mylight.h:106
    std::map<std::string, std::vector<lightTriangle>> lightsTriangles;  // Mylight.h:108
    std::map<triangle, dvec> islight;                                   // Mylight.h:111
    CameraDesc camera;

    explicit Mylight(std::string inputfile_) : inputfile(std::move(inputfile_)) {}
    void read();                                     // Mylight.cpp:11-30 (+ <camera>)
    void gather_light_triangles(const Myobj& obj);  // Mylight.cpp:32-100
};

// Flat arrays in the McptSceneDesc layout; owns the storage desc() points into.
struct FlatScene {
    std::vector<float> positions, normals, mat_kd, mat_ks, mat_ns;
    std::vector<int32_t> material_id, light_prim, light_class;
    std::vector<uint32_t> shape_offset;
    std::vector<double> class_radiance;
    McptSceneDesc desc() const;
};
FlatScene flatten(const Myobj& obj, const Mylight& lights);

// geometric normal with the reference's orientation rule on raw arrays (used by scene_create)
dvec unique_normal(const float* v9, const float* n9);

}  // namespace mcpt
