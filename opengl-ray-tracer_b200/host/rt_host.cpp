// rt_host.cpp — headless CPU side: scenes, reference BVH builder, flattening.
// See rt_host.h for what it restates.  Plain fp32, built with -ffp-contract=off
// so the produced bytes are reproducible.
#include "rt_host.h"
#include "rt_anim_math.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <random>
#include <string>
#include <vector>

namespace {

struct V3 { float x, y, z; float operator[](int i) const { return i == 0 ? x : i == 1 ? y : z; } };
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline V3 normalize(V3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }
inline V3 vmin(V3 a, V3 b) { return V3{b.x < a.x ? b.x : a.x, b.y < a.y ? b.y : a.y, b.z < a.z ? b.z : a.z}; }
inline V3 vmax(V3 a, V3 b) { return V3{a.x < b.x ? b.x : a.x, a.y < b.y ? b.y : a.y, a.z < b.z ? b.z : a.z}; }
inline void put(float* d, V3 v) { d[0] = v.x; d[1] = v.y; d[2] = v.z; }

// material.hpp:23 defaults
struct Mat { V3 color{1, 1, 1}; float fresnel = 1.f, ambient = .4f, diffuse = 1.f, specular = .5f; int shininess = 32; };

// One record for all four shape classes (shapes/*.hpp).
struct Shape {
    int type = 0;
    Mat mat;
    V3 origin{0, 0, 0};
    bool animated = false;
    V3 center{0, 0, 0}; float radius = 0;          // Sphere   (sphere.hpp:27-33)
    V3 normal{0, 0, 0}; float d = 0;               // Plane    (plane.hpp:29-34)
    V3 start{0, 0, 0};  float width = 0, height = 0;   // Wall (wall.hpp:39-42)
    V3 a{0, 0, 0}, b{0, 0, 0}, c{0, 0, 0};         // Triangle (triangle.hpp:47-71)
};

Shape make_sphere(V3 c, float r) { Shape s; s.type = RT_SHAPE_SPHERE; s.center = c; s.radius = r; s.origin = c; return s; }
void set_plane(Shape& s, V3 n, V3 p) { s.normal = normalize(n); s.d = -dot(s.normal, p); s.origin = p; }
Shape make_plane(V3 n, V3 p) { Shape s; s.type = RT_SHAPE_PLANE; set_plane(s, n, p); return s; }
Shape make_wall(V3 st, float w, float h, V3 n) {
    Shape s; s.type = RT_SHAPE_WALL; set_plane(s, n, st); s.start = st; s.width = w; s.height = h; return s;
}
Shape make_triangle(V3 p1, V3 p2, V3 p3) {          // triangle.hpp:47-48, 83-92
    Shape s; s.type = RT_SHAPE_TRIANGLE; s.a = p1; s.b = p2; s.c = p3;
    set_plane(s, cross(p2 - p1, p3 - p1), p1);
    return s;
}
void invert_normal(Shape& s) { s.normal = -s.normal; s.d = -dot(s.normal, s.a); }   // triangle.hpp:94-97

V3 wall_end(const Shape& s) {                        // wall.hpp:16-31
    V3 t1, n = s.normal;
    if (std::fabs(n.x) > std::fabs(n.y)) t1 = normalize(v3(-n.z, 0, n.x));
    else t1 = normalize(v3(0, -n.z, n.y));
    V3 t2 = normalize(cross(n, t1));
    return s.start + (s.width * t1) + (s.height * t2);
}

struct BBox {                                        // BoundingBox.hpp:34-95
    V3 Min{INFINITY, INFINITY, INFINITY}, Max{-INFINITY, -INFINITY, -INFINITY};
    void grow(V3 p) { Min = vmin(Min, p); Max = vmax(Max, p); }
    void grow(const Shape& s) {
        if (s.type == RT_SHAPE_SPHERE) {
            grow(v3(s.center.x + s.radius, s.center.y + s.radius, s.center.z + s.radius));
            grow(v3(s.center.x - s.radius, s.center.y - s.radius, s.center.z - s.radius));
        } else if (s.type == RT_SHAPE_WALL) {
            grow(s.start); grow(wall_end(s));
        } else if (s.type == RT_SHAPE_TRIANGLE) {
            if (std::isfinite(s.a.x) && std::isfinite(s.b.x) && std::isfinite(s.c.x)) { grow(s.a); grow(s.b); grow(s.c); }
        }   // infinite planes are ignored (BoundingBox.hpp:87-95)
    }
    V3 center() const { return (Min + Max) * 0.5f; }
};

struct Node { BBox box; int left = -1, right = -1; std::vector<int> idx; };   // main.cpp:74-87

struct Wheel { std::vector<int> idx; V3 center{0, 0, 0}; };

}  // namespace

struct rh_scene {
    std::string name;
    rt_flat_light light{};
    rt_flat_camera cam{};
    std::vector<Shape> shapes;
    std::vector<std::unique_ptr<Node>> nodes;
    std::vector<rt_flat_shape> flatShapes;
    std::vector<rt_flat_node> flatNodes;
    std::vector<int32_t> flatIndices;
    std::vector<int32_t> animated;
    Wheel wheels[4];
    int bvhDepth = 15;
    int W = 800, H = 600;
};

namespace {

// main.cpp:1111-1173
void split(rh_scene& sc, std::unique_ptr<Node>& parent, int depth) {
    if (depth <= 0) { parent->left = parent->right = -1; return; }
    V3 size = parent->box.Max - parent->box.Min;
    int axis = size.x > (size.y < size.z ? size.z : size.y) ? 0 : size.y > size.z ? 1 : 2;
    float pos = parent->box.center()[axis];
    auto L = std::make_unique<Node>();
    auto R = std::make_unique<Node>();
    for (int i : parent->idx) {
        const Shape& s = sc.shapes[(size_t)i];
        V3 c = v3(0, 0, 0);                      // upstream leaves a plane's centre uninitialised; 0 here
        if (s.type == RT_SHAPE_SPHERE) c = s.center;
        else if (s.type == RT_SHAPE_WALL) c = (s.start + wall_end(s)) * 0.5f;
        else if (s.type == RT_SHAPE_TRIANGLE) c = (s.a + s.b + s.c) / 3.0f;
        Node& side = c[axis] < pos ? *L : *R;
        side.box.grow(s);
        side.idx.push_back(i);
    }
    if (L->idx.empty() || R->idx.empty()) { parent->left = parent->right = -1; return; }
    split(sc, L, depth - 1);
    split(sc, R, depth - 1);
    sc.nodes.push_back(std::move(L)); parent->left = (int)sc.nodes.size() - 1;
    sc.nodes.push_back(std::move(R)); parent->right = (int)sc.nodes.size() - 1;
}

// main.cpp:1175-1193
void build_bvh(rh_scene& sc, int maxDepth) {
    sc.nodes.clear();
    auto root = std::make_unique<Node>();
    for (int i = 0; i < (int)sc.shapes.size(); ++i) { root->box.grow(sc.shapes[(size_t)i]); root->idx.push_back(i); }
    split(sc, root, maxDepth);
    sc.nodes.push_back(std::move(root));
}

// main.cpp:1068-1077
void update_bvh(rh_scene& sc) {
    for (auto& n : sc.nodes)
        for (int i : n->idx)
            if (sc.shapes[(size_t)i].animated) n->box.grow(sc.shapes[(size_t)i]);
}

// main.cpp:955-979
void serialize_bvh(rh_scene& sc) {
    sc.flatNodes.clear(); sc.flatIndices.clear();
    sc.flatNodes.reserve(sc.nodes.size());
    for (auto& n : sc.nodes) {
        rt_flat_node f; std::memset(&f, 0, sizeof f);
        put(f.boundsMin, n->box.Min); put(f.boundsMax, n->box.Max);
        f.leftChild = n->left; f.rightChild = n->right;
        f.startShapeIdx = (int32_t)sc.flatIndices.size();
        f.numShapes = (int32_t)n->idx.size();
        sc.flatNodes.push_back(f);
        if (n->left == -1) for (int i : n->idx) sc.flatIndices.push_back(i);
    }
}

// main.cpp:994-1066 (foreign-type fields are zero here; upstream leaves scalars indeterminate)
rt_flat_shape serialize_shape(const Shape& s) {
    rt_flat_shape f; std::memset(&f, 0, sizeof f);
    f.type = s.type;
    put(f.material.color, s.mat.color);
    f.material.fresnelStrength = s.mat.fresnel; f.material.ambientStrength = s.mat.ambient;
    f.material.diffuseStrength = s.mat.diffuse; f.material.specularStrength = s.mat.specular;
    f.material.shininess = s.mat.shininess;
    if (s.type == RT_SHAPE_SPHERE) { put(f.sphereCenter, s.center); f.sphereRadius = s.radius; }
    else {
        put(f.planeNormal, s.normal); f.planeD = s.d;
        if (s.type == RT_SHAPE_WALL) { put(f.wallStart, s.start); f.wallWidth = s.width; f.wallHeight = s.height; }
        if (s.type == RT_SHAPE_TRIANGLE) { put(f.triP1, s.a); put(f.triP2, s.b); put(f.triP3, s.c); }
    }
    return f;
}

void serialize_all(rh_scene& sc) {
    sc.flatShapes.resize(sc.shapes.size());
    for (size_t i = 0; i < sc.shapes.size(); ++i) sc.flatShapes[i] = serialize_shape(sc.shapes[i]);
    serialize_bvh(sc);
}

// camera.hpp:139-163 + serializeCamera main.cpp:806-816
void look_at(rh_scene& sc, V3 pos, V3 target, float fov, float aspect) {
    const float RAD = 0.01745329251994329576923690768489f, DEG = 57.295779513082320876798154814105f;
    V3 dir = normalize(target - pos);
    float pitch = std::asin(dir.y) * DEG;
    float yaw = std::atan2(dir.z, dir.x) * DEG;
    V3 front = normalize(v3(std::cos(yaw * RAD) * std::cos(pitch * RAD), std::sin(pitch * RAD),
                            std::sin(yaw * RAD) * std::cos(pitch * RAD)));
    V3 right = normalize(cross(front, v3(0, 1, 0)));
    V3 up = normalize(cross(right, front));
    std::memset(&sc.cam, 0, sizeof sc.cam);
    put(sc.cam.Position, pos); put(sc.cam.Front, front); put(sc.cam.Up, up); put(sc.cam.Right, right);
    sc.cam.aspectRatio = aspect; sc.cam.fov = fov;
}

void set_light(rh_scene& sc, V3 pos, V3 base, float intensity) {      // light.hpp:22-35
    std::memset(&sc.light, 0, sizeof sc.light);
    put(sc.light.position, pos); put(sc.light.color, intensity * base);
}

// Deterministic uniform floats (upstream seeds from std::random_device, main.cpp:932-953).
struct Rng {
    std::mt19937 gen;
    explicit Rng(uint32_t seed) : gen(seed) {}
    float u01() { return (float)(gen() >> 8) * (1.0f / 16777216.0f); }
    float uni(float lo, float hi) { return lo + (hi - lo) * u01(); }
};

// Lat-long ellipsoid patch without the pole rows: nLon x nLat quads -> 2 triangles
// each, counter-clockwise seen from outside (the usual OBJ export convention).
void add_ellipsoid(rh_scene& sc, V3 c, V3 r, int nLon, int nLat, const Mat& m) {
    const float PI = 3.14159265358979323846f;
    auto P = [&](int i, int j) {
        float th = PI * (float)(j + 1) / (float)(nLat + 2);
        float ph = 2.0f * PI * (float)(i % nLon) / (float)nLon;
        return v3(c.x + r.x * std::sin(th) * std::cos(ph), c.y + r.y * std::cos(th), c.z + r.z * std::sin(th) * std::sin(ph));
    };
    for (int j = 0; j < nLat; ++j)
        for (int i = 0; i < nLon; ++i) {
            V3 p00 = P(i, j), p10 = P(i + 1, j), p01 = P(i, j + 1), p11 = P(i + 1, j + 1);
            Shape t1 = make_triangle(p00, p10, p11); t1.mat = m; sc.shapes.push_back(t1);
            Shape t2 = make_triangle(p00, p11, p01); t2.mat = m; sc.shapes.push_back(t2);
        }
}

// A wheel: axis z, nSeg segments x nRing bands on the tread + two cap fans.
void add_wheel(rh_scene& sc, int w, V3 c, float radius, float halfWidth, int nSeg, int nRing, const Mat& m) {
    const float PI = 3.14159265358979323846f;
    auto rim = [&](int i, float z) {
        float a = 2.0f * PI * (float)(i % nSeg) / (float)nSeg;
        return v3(c.x + radius * std::cos(a), c.y + radius * std::sin(a), z);
    };
    V3 sum = v3(0, 0, 0);
    size_t first = sc.shapes.size();
    auto push = [&](V3 a, V3 b, V3 cc) {
        Shape t = make_triangle(a, b, cc); t.mat = m; t.animated = true;
        sc.wheels[w].idx.push_back((int)sc.shapes.size());
        sc.animated.push_back((int32_t)sc.shapes.size());
        sc.shapes.push_back(t);
        sum = sum + a; sum = sum + b; sum = sum + cc;             // main.cpp:771-773
    };
    for (int k = 0; k < nRing; ++k) {
        float z0 = c.z - halfWidth + 2.0f * halfWidth * (float)k / (float)nRing;
        float z1 = c.z - halfWidth + 2.0f * halfWidth * (float)(k + 1) / (float)nRing;
        for (int i = 0; i < nSeg; ++i) {
            push(rim(i, z0), rim(i + 1, z0), rim(i + 1, z1));
            push(rim(i, z0), rim(i + 1, z1), rim(i, z1));
        }
    }
    for (int i = 0; i < nSeg; ++i) {
        push(v3(c.x, c.y, c.z - halfWidth), rim(i + 1, c.z - halfWidth), rim(i, c.z - halfWidth));
        push(v3(c.x, c.y, c.z + halfWidth), rim(i, c.z + halfWidth), rim(i + 1, c.z + halfWidth));
    }
    sc.wheels[w].center = sum / (float)((sc.shapes.size() - first) * 3);   // main.cpp:780-781
}

}  // namespace

namespace {

// generateScene1, main.cpp:583-716 (meshes: synthetic stand-ins, 968 + 240 triangles)
void gen_monkey(rh_scene& sc) {
    set_light(sc, v3(0, -14, 0), v3(1, 1, 1), 50);                                   // :591
    auto sphere = [&](V3 c, float r, V3 col, float fr, float am, float di, float sp, bool anim) {
        Shape s = make_sphere(c, r); s.mat.color = col; s.mat.fresnel = fr; s.mat.ambient = am;
        s.mat.diffuse = di; s.mat.specular = sp; s.animated = anim;
        if (anim) sc.animated.push_back((int32_t)sc.shapes.size());
        sc.shapes.push_back(s);
    };
    sphere(v3(0, 10, -8), 5.f, v3(0, 0.37f, 0), 0, 0.2f, 1, 0.1f, true);               // :594-600
    sphere(v3(12, 10, -8), 4.f, v3(0.58f, 0.18f, 0.48f), 0, 0, 0.5f, 0, true);         // :602-608
    sphere(v3(20, 7.5f, -8), 2.5f, v3(0.8f, 0.2f, 0.8f), 1, 0.06f, 0.06f, 0.5f, true); // :610-616
    sphere(v3(0, 23, -8), 1.5f, v3(0, 0.37f, 0), 0, 0, 0.5f, 0, false);                // :618-623
    {   // mirror wall :626-630
        Shape w = make_wall(v3(-15, 23, 10), 30, 20, v3(-1, 0.2f, 0));
        w.mat.fresnel = 1; w.mat.ambient = 0.1f; w.mat.diffuse = 0; w.mat.specular = 1;
        sc.shapes.push_back(w);
    }
    {   // inverted triangle :633-643
        Shape t = make_triangle(v3(-15, 20, 25), v3(-12, 20, 10), v3(-15, 0, 20));
        invert_normal(t);
        t.mat.color = v3(0.19f, 0.66f, 0.32f); t.mat.fresnel = 1; t.mat.ambient = 0.06f; t.mat.diffuse = 0.06f; t.mat.specular = 0.5f;
        sc.shapes.push_back(t);
    }
    Mat m1; m1.color = v3(179.f / 255, 165.f / 255, 61.f / 255); m1.fresnel = 1; m1.ambient = 0.2f; m1.diffuse = 0.8f; m1.specular = 0.1f;   // :655-660
    add_ellipsoid(sc, v3(0, 0, -30), v3(6, 6, 5), 22, 22, m1);                       // "monkey.obj" stand-in, 968
    Mat m2; m2.color = v3(0, 1, .9f); m2.fresnel = 1; m2.ambient = 0.2f; m2.diffuse = 0.8f; m2.specular = 0;                                // :672-677
    add_ellipsoid(sc, v3(50, 0, -30), v3(5, 5, 4), 12, 10, m2);                      // "lowpolymonkey.obj" stand-in, 240
    Rng rng(1);
    for (int i = 0; i < 25; ++i) {                                                   // :684-695
        float px = rng.uni(-40, 40), pz = rng.uni(-40, 40);
        Shape s = make_sphere(v3(px, 23, pz), 1.5f);
        s.mat.color = v3(rng.u01(), rng.u01(), rng.u01());
        sc.shapes.push_back(s);
    }
    {   // floor :699-701 (normal (0,1,0): NaN tangent basis in the shader)
        Shape w = make_wall(v3(-100, 25, -100), 210, 210, v3(0, 1, 0));
        w.mat.color = v3(0.65f, 0.17f, 0.35f); w.mat.specular = 0;
        sc.shapes.push_back(w);
    }
    look_at(sc, v3(30, -5, 40), sc.shapes[0].origin, 60, (float)sc.W / (float)sc.H);   // :587-588,704
    sc.bvhDepth = 15;                                                                // :712
}

// generateScene2, main.cpp:718-804 (car.obj stand-in: 3008 body + 960 wheel + 54 road triangles)
void gen_car(rh_scene& sc) {
    set_light(sc, v3(14.8f, -17, 17), v3(1, 1, 1), 26);                              // :725
    Mat body; body.color = v3(19.f / 255, 7.f / 255, 92.f / 255); body.specular = 0;     // :739-740
    add_ellipsoid(sc, v3(0, -3, 0), v3(10, 3, 4), 47, 32, body);
    Mat wheel; wheel.color = v3(.2f, .2f, .2f); wheel.specular = 0;                  // :746-747
    const V3 wc[4] = {v3(-6, 0, -3.5f), v3(6, 0, -3.5f), v3(-6, 0, 3.5f), v3(6, 0, 3.5f)};
    for (int w = 0; w < 4; ++w) add_wheel(sc, w, wc[w], 1.5f, 0.5f, 24, 4, wheel);
    Mat road; road.color = v3(0, 0, 0); road.specular = .25f;                        // :750-751
    for (int j = 0; j < 3; ++j)
        for (int i = 0; i < 9; ++i) {
            float x0 = -45.f + 10.f * (float)i, x1 = x0 + 10.f, z0 = -15.f + 10.f * (float)j, z1 = z0 + 10.f, y = 1.5f;
            // wound so that the plane normal is +y: hit from above in this y-down world
            Shape t1 = make_triangle(v3(x0, y, z0), v3(x1, y, z1), v3(x1, y, z0)); t1.mat = road; sc.shapes.push_back(t1);
            Shape t2 = make_triangle(v3(x0, y, z0), v3(x0, y, z1), v3(x1, y, z1)); t2.mat = road; sc.shapes.push_back(t2);
        }
    Rng rng(1234);
    for (int i = 0; i < 100; ++i) {                                                  // :788-795
        float px = rng.uni(-30, 30), py = rng.uni(-15, 0);
        Shape s = make_sphere(v3(px, py, -10), 1.5f);
        s.mat.color = v3(rng.u01(), rng.u01(), rng.u01());
        sc.shapes.push_back(s);
    }
    look_at(sc, v3(0, -10, 40), v3(0, 0, 0), 60, (float)sc.W / (float)sc.H);         // :721-722,797
    sc.bvhDepth = 25;                                                                // :800
}

// BASELINE.json configs[3]: triangle soup + spheres, deep BVH
void gen_soup(rh_scene& sc, int nTri) {
    set_light(sc, v3(14.8f, -17, 17), v3(1, 1, 1), 26);
    Rng rng(42);
    sc.shapes.reserve((size_t)nTri + (size_t)nTri / 100);
    for (int i = 0; i < nTri; ++i) {
        V3 a = v3(rng.uni(-30, 30), rng.uni(-15, 5), rng.uni(-20, 10));
        V3 b = a + v3(rng.uni(-.15f, .15f), rng.uni(-.15f, .15f), rng.uni(-.15f, .15f));
        V3 c = a + v3(rng.uni(-.15f, .15f), rng.uni(-.15f, .15f), rng.uni(-.15f, .15f));
        Shape t = make_triangle(a, b, c);
        t.mat.color = v3(rng.u01(), rng.u01(), rng.u01()); t.mat.specular = (i % 4 == 0) ? 0.3f : 0.f;
        sc.shapes.push_back(t);
    }
    for (int i = 0; i < nTri / 100; ++i) {
        Shape s = make_sphere(v3(rng.uni(-30, 30), rng.uni(-15, 0), -10), 0.15f);
        s.mat.color = v3(rng.u01(), rng.u01(), rng.u01());
        sc.shapes.push_back(s);
    }
    look_at(sc, v3(0, -10, 40), v3(0, 0, 0), 60, (float)sc.W / (float)sc.H);
    sc.bvhDepth = 40;
}

// generateScene3, main.cpp:1196-1229 (upstream never builds a BVH for it; one is built here)
void gen_triangle(rh_scene& sc) {
    set_light(sc, v3(14.8f, -17, 17), v3(1, 1, 1), 26);
    sc.shapes.push_back(make_triangle(v3(0, 0, 0), v3(5, 0, 0), v3(2.5f, -5, 0)));
    look_at(sc, v3(0, -10, 40), v3(0, 0, 0), 60, (float)sc.W / (float)sc.H);
    sc.bvhDepth = 15;
}

// Small scene touching every shape type and the documented edge cases.
void gen_toy(rh_scene& sc) {
    set_light(sc, v3(3, -12, 14), v3(1, 1, 1), 30);
    Shape s0 = make_sphere(v3(0, 0, 0), 3.f); s0.mat.color = v3(0.9f, 0.3f, 0.2f); s0.animated = true;
    sc.animated.push_back(0); sc.shapes.push_back(s0);
    Shape s1 = make_sphere(v3(6, 1, -2), 2.f); s1.mat.color = v3(0.2f, 0.4f, 0.9f); s1.mat.specular = 0; sc.shapes.push_back(s1);
    Shape s2 = make_sphere(v3(-5, -2, 3), 1.5f); s2.mat.fresnel = 0.7f; sc.shapes.push_back(s2);
    Shape fl = make_wall(v3(-40, 4, -40), 80, 80, v3(0, 1, 0)); fl.mat.color = v3(0.5f, 0.5f, 0.3f); fl.mat.specular = 0.2f;
    sc.shapes.push_back(fl);                                        // NaN-basis floor
    Shape mw = make_wall(v3(-12, 4, -6), 10, 9, v3(-1, 0.1f, -0.3f)); mw.mat.specular = 1; mw.mat.diffuse = 0.1f; sc.shapes.push_back(mw);
    Shape t0 = make_triangle(v3(-3, 3, 6), v3(4, 3, 6), v3(0, -6, 5)); t0.mat.color = v3(0.1f, 0.8f, 0.3f); sc.shapes.push_back(t0);
    Shape t1 = make_triangle(v3(8, 3, 4), v3(3, 3, 8), v3(7, -4, 6)); invert_normal(t1); t1.mat.specular = 0; sc.shapes.push_back(t1);
    Shape dg = make_triangle(v3(1, 1, 1), v3(2, 2, 2), v3(3, 3, 3)); sc.shapes.push_back(dg);   // degenerate: NaN normal
    look_at(sc, v3(2, -6, 22), v3(0, 0, 0), 60, (float)sc.W / (float)sc.H);
    sc.bvhDepth = 8;
}

void finish(rh_scene& sc, int depth) {
    if (depth > 0) sc.bvhDepth = depth;
    build_bvh(sc, sc.bvhDepth);
    serialize_all(sc);
}

// updateWheelAnimations main.cpp:1084-1109: transform = translate(center) * rotate(rotationSpeed * deltaTime, (0,0,1)) *
// translate(-center) as a GLM mat4, every vertex through mat4 * vec4 (rt_anim_math.h restates GLM's operation order and is
// pinned to the vendored GLM bit for bit); m_normal / d are NOT refreshed upstream (stale planes), reproduced.
void rotate_wheels(rh_scene& sc, float dt) {
    const float ang = 1.0f * dt, cs = std::cos(ang), sn = std::sin(ang);
    const float axis[3] = {0.f, 0.f, 1.f};                        // main.cpp:779
    for (Wheel& w : sc.wheels) {
        const float center[3] = {w.center.x, w.center.y, w.center.z};
        float m[16];
        rta_wheel_transform(center, cs, sn, axis, m);
        for (int i : w.idx) {
            Shape& t = sc.shapes[(size_t)i];
            for (V3* p : {&t.a, &t.b, &t.c}) {
                const float in[3] = {p->x, p->y, p->z};
                float out[3];
                rta_apply(m, in, out);
                p->x = out[0]; p->y = out[1]; p->z = out[2];
            }
        }
    }
}

}  // namespace

extern "C" {

rh_scene* rh_scene_create(const char* name, int width, int height, int n_override, int bvh_depth) {
    if (!name || width <= 0 || height <= 0) return nullptr;
    auto sc = std::make_unique<rh_scene>();
    sc->name = name; sc->W = width; sc->H = height;
    if (sc->name == "monkey") gen_monkey(*sc);
    else if (sc->name == "car") gen_car(*sc);
    else if (sc->name == "soup") gen_soup(*sc, n_override > 0 ? n_override : 1000000);
    else if (sc->name == "triangle") gen_triangle(*sc);
    else if (sc->name == "toy") gen_toy(*sc);
    else return nullptr;
    finish(*sc, bvh_depth);
    return sc.release();
}

rh_scene* rh_scene_from_shapes(const rt_flat_shape* shapes, int nShapes, const rt_flat_light* light,
                               const rt_flat_camera* camera, int bvh_depth) {
    if ((nShapes > 0 && !shapes) || !light || !camera) return nullptr;
    auto sc = std::make_unique<rh_scene>();
    sc->name = "custom"; sc->light = *light; sc->cam = *camera;
    for (int i = 0; i < nShapes; ++i) {
        const rt_flat_shape& f = shapes[i];
        Shape s; s.type = f.type;
        s.mat.color = v3(f.material.color); s.mat.fresnel = f.material.fresnelStrength; s.mat.ambient = f.material.ambientStrength;
        s.mat.diffuse = f.material.diffuseStrength; s.mat.specular = f.material.specularStrength; s.mat.shininess = f.material.shininess;
        s.center = v3(f.sphereCenter); s.radius = f.sphereRadius; s.normal = v3(f.planeNormal); s.d = f.planeD;
        s.start = v3(f.wallStart); s.width = f.wallWidth; s.height = f.wallHeight;
        s.a = v3(f.triP1); s.b = v3(f.triP2); s.c = v3(f.triP3);
        sc->shapes.push_back(s);
    }
    build_bvh(*sc, bvh_depth > 0 ? bvh_depth : 15);
    sc->flatShapes.assign(shapes, shapes + nShapes);      // caller's bytes pass through untouched
    serialize_bvh(*sc);
    return sc.release();
}

// ---- Wavefront OBJ -> flat triangles: the reference's model path (model.hpp:49-168 + mesh.hpp:163-189) without assimp ----
// The reference reads its meshes through assimp with aiProcess_Triangulate and WITHOUT aiProcess_JoinIdenticalVertices:
// the OBJ importer then hands out one vertex per face corner and one mesh per object / group / material run that has
// faces; main.cpp offsets every vertex by mesh.origin (mesh.hpp:171-173) and builds `Triangle(a, b, c)` from the corners
// (main.cpp:652-654, 669-671, 760-762) — the inverted normal of mesh2triangles is dropped on that way, so the plane of a
// scene triangle is always the one its corners span (triangle.hpp:47-48, 83-92).  apply_rule != 0 keeps mesh2triangles'
// own rule instead (a triangle whose normal has a positive dot product with Mesh::center(), mesh.hpp:51-61 =
// (origin + sum over the face corners of (origin + p)) / corners, is inverted: mesh.hpp:178-184).
// Polygons become fans (v0, vi, vi+1), as assimp triangulates a convex polygon; `v` and `f` records with v, v/vt, v//vn,
// v/vt/vn and negative (relative) indices are understood, everything else is skipped.
int rh_obj_triangles(const char* path, const float origin[3], const rt_flat_shape* material_from, int apply_rule,
                     rt_flat_shape* out, int cap, int32_t* mesh_first, int mesh_cap, int* n_meshes) {
    if (!path || !origin) return -1;
    FILE* fp = std::fopen(path, "rb");
    if (!fp) return -2;
    Mat mat;
    if (material_from) {
        const rt_flat_material& m = material_from->material;
        mat.color = v3(m.color); mat.fresnel = m.fresnelStrength; mat.ambient = m.ambientStrength; mat.diffuse = m.diffuseStrength;
        mat.specular = m.specularStrength; mat.shininess = m.shininess;
    }
    const V3 org = v3(origin[0], origin[1], origin[2]);
    std::vector<V3> verts;
    std::vector<Shape> tris;
    std::vector<V3> raw;                      // the face corners as read (3 per triangle): assimp's vertex array
    std::vector<int32_t> first;               // first triangle of every mesh
    bool meshOpen = false, bad = false;
    char line[4096];
    while (std::fgets(line, sizeof line, fp)) {
        const char* q = line;
        while (*q == ' ' || *q == '\t') ++q;
        if (q[0] == 'v' && (q[1] == ' ' || q[1] == '\t')) {
            float x = 0, y = 0, z = 0;
            if (std::sscanf(q + 1, "%f %f %f", &x, &y, &z) != 3) { bad = true; break; }
            verts.push_back(v3(x, y, z));
        } else if (q[0] == 'f' && (q[1] == ' ' || q[1] == '\t')) {
            std::vector<int> idx;
            const char* r = q + 1;
            for (;;) {
                while (*r == ' ' || *r == '\t') ++r;
                if (*r == 0 || *r == '\n' || *r == '\r' || *r == '#') break;
                char* end = nullptr;
                const long v = std::strtol(r, &end, 10);
                if (end == r) { bad = true; break; }
                const long at = v > 0 ? v - 1 : (long)verts.size() + v;     // negative: relative to the vertices read so far
                if (v == 0 || at < 0 || at >= (long)verts.size()) { bad = true; break; }
                idx.push_back((int)at);
                r = end;
                while (*r && *r != ' ' && *r != '\t' && *r != '\n' && *r != '\r') ++r;   // skip /vt/vn
            }
            if (bad) break;
            if (idx.size() < 3) continue;
            if (!meshOpen) { first.push_back((int32_t)tris.size()); meshOpen = true; }
            for (size_t i = 1; i + 1 < idx.size(); ++i) {
                Shape t = make_triangle(verts[(size_t)idx[0]] + org, verts[(size_t)idx[i]] + org, verts[(size_t)idx[i + 1]] + org);
                t.mat = mat;
                tris.push_back(t);
                raw.push_back(verts[(size_t)idx[0]]); raw.push_back(verts[(size_t)idx[i]]); raw.push_back(verts[(size_t)idx[i + 1]]);
            }
        } else if ((q[0] == 'o' || q[0] == 'g') && (q[1] == ' ' || q[1] == '\t' || q[1] == '\n' || q[1] == '\r')) {
            meshOpen = false;                 // the next face opens a new mesh
        } else if (std::strncmp(q, "usemtl", 6) == 0) {
            meshOpen = false;
        }
    }
    std::fclose(fp);
    if (bad) return -3;
    first.push_back((int32_t)tris.size());
    if (apply_rule) {
        for (size_t m = 0; m + 1 < first.size(); ++m) {
            const size_t t0 = (size_t)first[m], t1 = (size_t)first[m + 1];
            V3 c = org;                        // Mesh::center(), vertex by vertex in assimp's order (the face corners)
            for (size_t k = 3 * t0; k < 3 * t1; ++k) c = c + (org + raw[k]);
            c = c / (float)((t1 - t0) * 3);
            for (size_t t = t0; t < t1; ++t) if (dot(tris[t].normal, c) > 0.0f) invert_normal(tris[t]);
        }
    }
    const int n = (int)tris.size();
    for (int i = 0; i < n && i < cap && out; ++i) out[i] = serialize_shape(tris[(size_t)i]);
    const int nm = (int)first.size() - 1;
    if (n_meshes) *n_meshes = nm;
    for (int m = 0; m <= nm && m < mesh_cap && mesh_first; ++m) mesh_first[m] = first[(size_t)m];
    return n;
}

void rh_scene_destroy(rh_scene* s) { delete s; }
const rt_flat_light* rh_light(const rh_scene* s) { return &s->light; }
const rt_flat_camera* rh_camera(const rh_scene* s) { return &s->cam; }
const rt_flat_shape* rh_shapes(const rh_scene* s, int* n) { if (n) *n = (int)s->flatShapes.size(); return s->flatShapes.data(); }
const rt_flat_node* rh_nodes(const rh_scene* s, int* n) { if (n) *n = (int)s->flatNodes.size(); return s->flatNodes.data(); }
const int32_t* rh_indices(const rh_scene* s, int* n) { if (n) *n = (int)s->flatIndices.size(); return s->flatIndices.data(); }
const int32_t* rh_animated(const rh_scene* s, int* n) { if (n) *n = (int)s->animated.size(); return s->animated.data(); }

void rh_bvh_stats(const rh_scene* s, int out[4]) {
    int leaves = 0, maxLeaf = 0;
    for (auto& n : s->nodes) if (n->left == -1) { ++leaves; if ((int)n->idx.size() > maxLeaf) maxLeaf = (int)n->idx.size(); }
    std::vector<int> depth(s->nodes.size(), 0);
    int maxDepth = 0;
    for (int i = (int)s->nodes.size() - 1; i >= 0; --i) {       // parents have larger indices than children
        const Node& n = *s->nodes[(size_t)i];
        if (depth[(size_t)i] > maxDepth) maxDepth = depth[(size_t)i];
        if (n.left != -1) { depth[(size_t)n.left] = depth[(size_t)i] + 1; depth[(size_t)n.right] = depth[(size_t)i] + 1; }
    }
    out[0] = (int)s->nodes.size(); out[1] = leaves; out[2] = maxLeaf; out[3] = maxDepth;
}

int rh_scene_animate(rh_scene* s, float t, float dt, int mode) {
    if (s->name == "monkey") {                                   // main.cpp:440-446, 1079-1082
        const float amp[3] = {10, 7, 15}, freq[3] = {1, 0.8f, 1.5f};
        for (int k = 0; k < 3; ++k) s->shapes[(size_t)k].center.y = s->shapes[(size_t)k].origin.y + amp[k] * std::sin(freq[k] * t);
    } else if (s->name == "car") {
        rotate_wheels(*s, dt);
    } else if (s->name == "toy") {
        s->shapes[0].center.y = s->shapes[0].origin.y + 2.f * std::sin(t);
    }
    for (int32_t i : s->animated) s->flatShapes[(size_t)i] = serialize_shape(s->shapes[(size_t)i]);   // updateScene :981-992
    if (mode == 1) build_bvh(*s, s->bvhDepth); else update_bvh(*s);
    serialize_bvh(*s);                                           // :342
    return (int)s->animated.size();
}

int rh_anim_count(const rh_scene* s) {
    if (s->name == "monkey") return 3;
    if (s->name == "car") return 4;
    if (s->name == "toy") return 1;
    return 0;
}
const int32_t* rh_anim_get(const rh_scene* s, int k, int* kind, int* count, float params[4]) {
    static thread_local std::vector<int32_t> list;
    list.clear();
    if (k < 0 || k >= rh_anim_count(s)) return nullptr;
    if (s->name == "car") {
        const Wheel& w = s->wheels[k];
        for (int i : w.idx) list.push_back(i);
        *kind = 1; *count = (int)list.size();
        params[0] = w.center.x; params[1] = w.center.y; params[2] = w.center.z; params[3] = 1.0f;
        return list.data();
    }
    const float amp[3] = {10, 7, 15}, freq[3] = {1, 0.8f, 1.5f};
    const bool toy = s->name == "toy";
    list.push_back(k);
    *kind = 0; *count = 1;
    params[0] = s->shapes[(size_t)k].origin.y; params[1] = toy ? 2.f : amp[k]; params[2] = toy ? 1.f : freq[k]; params[3] = 0.f;
    return list.data();
}

void rh_set_camera(rh_scene* s, const float pos[3], const float target[3], float fov, float aspect) {
    look_at(*s, v3(pos), v3(target), fov, aspect);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// The interactive camera: the reference's Camera class (camera.hpp:30-165) as its render loop drives it every frame
// (processInput main.cpp:509-533, mouse_callback :536-551) and serializeCamera (:806-816), whose 80 bytes are the camera
// sub-upload (:328-335) = rt_update_camera's argument.  Pinned bit for bit to the reference class compiled in place with
// g++ (oracle/_ref/libref_camera.so, tests/test_camera.py).  One fact of that build decides the last bit: camera.hpp calls
// cos / sin / asin / atan2 UNQUALIFIED on floats, which g++ resolves to the C library's double functions (only std:: has the
// float overloads), so the products and glm::degrees of LookAt are evaluated in double and rounded once on assignment.
// (look_at above, which places the synthetic scenes' cameras, keeps its float evaluation: the golden frames depend on it.)
// ---------------------------------------------------------------------------
struct rh_cam {
    V3 Position, Front, Up, Right, WorldUp;
    float Yaw, Pitch, MovementSpeed, MouseSensitivity, Zoom, fov, aspectRatio;
};

namespace {

const float CAM_YAW = -90.0f, CAM_PITCH = 0.0f, CAM_SPEED = 15.f, CAM_SPEEDAMPLIFIER = 2.f, CAM_SENSITIVITY = 0.1f, CAM_ZOOM = 45.0f;   // camera.hpp:20-26
inline float cam_radians(float deg) { return deg * 0.01745329251994329576923690768489f; }        // glm::radians<float>

void update_camera_vectors(rh_cam& c) {                                                        // camera.hpp:152-163
    V3 front;
    front.x = (float)(std::cos((double)cam_radians(c.Yaw)) * std::cos((double)cam_radians(c.Pitch)));
    front.y = (float)std::sin((double)cam_radians(c.Pitch));
    front.z = (float)(std::sin((double)cam_radians(c.Yaw)) * std::cos((double)cam_radians(c.Pitch)));
    c.Front = normalize(front);
    c.Right = normalize(cross(c.Front, c.WorldUp));
    c.Up = normalize(cross(c.Right, c.Front));
}

}  // namespace

extern "C" {

rh_cam* rh_cam_create(const float pos[3], float aspect) {
    rh_cam* c = new rh_cam();
    // Camera() (camera.hpp:49-56): position 0, up +y, yaw -90, pitch 0, fov 60, aspect 1; then Position / aspectRatio as
    // generateScene1/2/3 set them (main.cpp:587-588, 721-722, 1198-1200)
    c->Front = v3(0.0f, 0.0f, -1.0f); c->MovementSpeed = CAM_SPEED; c->MouseSensitivity = CAM_SENSITIVITY; c->Zoom = CAM_ZOOM;
    c->fov = 60.0f; c->aspectRatio = 1.0f;
    c->Position = v3(0.0f, 0.0f, 0.0f); c->WorldUp = v3(0.0f, 1.0f, 0.0f); c->Yaw = CAM_YAW; c->Pitch = CAM_PITCH;
    update_camera_vectors(*c);
    if (pos) c->Position = v3(pos);
    c->aspectRatio = aspect;
    return c;
}
void rh_cam_destroy(rh_cam* c) { delete c; }

void rh_cam_look_at(rh_cam* c, const float target[3]) {                                     // camera.hpp:139-150
    const V3 direction = normalize(v3(target) - c->Position);
    c->Pitch = (float)(std::asin((double)direction.y) * 57.295779513082320876798154814105);       // glm::degrees<double>
    c->Yaw = (float)(std::atan2((double)direction.z, (double)direction.x) * 57.295779513082320876798154814105);
    update_camera_vectors(*c);
}

void rh_cam_keyboard(rh_cam* c, int direction, float deltaTime) {                           // camera.hpp:75-90
    const float velocity = c->MovementSpeed * deltaTime;
    if (direction == 0) c->Position = c->Position + c->Front * velocity;        // FORWARD
    if (direction == 1) c->Position = c->Position - c->Front * velocity;        // BACKWARD
    if (direction == 2) c->Position = c->Position - c->Right * velocity;        // LEFT
    if (direction == 3) c->Position = c->Position + c->Right * velocity;        // RIGHT
    if (direction == 4) c->Position = c->Position - c->Up * velocity;           // UP   (the reference's sign: its Up axis points down the screen)
    if (direction == 5) c->Position = c->Position + c->Up * velocity;           // DOWN
}

void rh_cam_mouse(rh_cam* c, float xoffset, float yoffset, int constrainPitch) {            // camera.hpp:93-112
    xoffset *= c->MouseSensitivity;
    yoffset *= c->MouseSensitivity;
    c->Yaw += xoffset;
    c->Pitch += yoffset;
    if (constrainPitch) {
        if (c->Pitch > 89.0f) c->Pitch = 89.0f;
        if (c->Pitch < -89.0f) c->Pitch = -89.0f;
    }
    update_camera_vectors(*c);
}

void rh_cam_scroll(rh_cam* c, float yoffset) {                                              // camera.hpp:115-122
    c->Zoom -= yoffset;
    if (c->Zoom < 1.0f) c->Zoom = 1.0f;
    if (c->Zoom > 45.0f) c->Zoom = 45.0f;
}

void rh_cam_shift(rh_cam* c, int pressed) { c->MovementSpeed = pressed ? CAM_SPEED * CAM_SPEEDAMPLIFIER : CAM_SPEED; }   // main.cpp:528

void rh_cam_state(const rh_cam* c, float out[4]) { out[0] = c->Yaw; out[1] = c->Pitch; out[2] = c->Zoom; out[3] = c->MovementSpeed; }

void rh_cam_serialize(const rh_cam* c, rt_flat_camera* out) {                               // serializeCamera main.cpp:806-816
    std::memset(out, 0, sizeof *out);
    put(out->Position, c->Position); out->aspectRatio = c->aspectRatio;
    put(out->Front, c->Front); put(out->Up, c->Up); put(out->Right, c->Right);
    out->fov = c->fov;
}

}  // extern "C"
