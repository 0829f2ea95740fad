/*
 * oracle_rt.cpp — CPU ORACLE.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A scalar fp32 restatement, function by function, of the reference's compute
 * shader src/shaders/gpu_shader.comp (the sole specification of the hot path;
 * the reference has no tests and no golden vectors).
 *
 * Parity status: the reference cannot be run through a graphics API on these
 * machines ("unavailable on this box").  The restatement is pinned instead
 * against the reference shader FILE ITSELF compiled as C++ with the vendored
 * GLM (oracle/build_ref.sh -> oracle/_ref/libref_shader.so, built from
 * /root/reference in place, never copied) and against the reference's own C++
 * shape classes; see tests/test_oracle_vs_ref.py and tests/golden/.
 *
 * Rules:
 *   - fp32 everywhere; build with -O2 -ffp-contract=off, no fast-math;
 *   - GLSL built-ins by their spec definitions, evaluated left to right:
 *       dot(a,b)      = a.x*b.x + a.y*b.y + a.z*b.z
 *       length(v)     = sqrt(dot(v,v));  distance(a,b) = length(a-b)
 *       normalize(v)  = v * (1/sqrt(dot(v,v)))
 *       reflect(I,N)  = I - N*dot(N,I)*2      (x2 is exact, order-free)
 *       mix(x,y,a)    = x*(1-a) + y*a
 *       min(x,y)      = y<x ? y : x ;  max(x,y) = x<y ? y : x   (NaN: first arg)
 *       radians(d)    = d * 0.017453292519943295f
 *       pow(x,int n)  = powf(x,(float)n)   (implicit int->float conversion)
 *   - every shape field is read only for the shape's own type.
 *
 * Build: see oracle/Makefile.
 */
#include "oracle_rt.h"

#include <cmath>
#include <cstring>
#include <thread>
#include <utility>
#include <vector>

namespace {

struct V3 { float x, y, z; };

inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }
inline V3 operator+(V3 a, V3 b) { return V3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 operator-(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator*(V3 a, V3 b) { return V3{a.x * b.x, a.y * b.y, a.z * b.z}; }
inline V3 operator*(V3 a, float s) { return V3{a.x * s, a.y * s, a.z * s}; }
inline V3 operator*(float s, V3 a) { return V3{s * a.x, s * a.y, s * a.z}; }
inline V3 operator/(V3 a, float s) { return V3{a.x / s, a.y / s, a.z / s}; }
inline V3 operator-(V3 a) { return V3{-a.x, -a.y, -a.z}; }

inline float g_min(float x, float y) { return y < x ? y : x; }
inline float g_max(float x, float y) { return x < y ? y : x; }
inline float g_dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float g_length(V3 v) { return std::sqrt(g_dot(v, v)); }
inline float g_distance(V3 a, V3 b) { return g_length(a - b); }
inline V3 g_normalize(V3 v) { float inv = 1.0f / std::sqrt(g_dot(v, v)); return v * inv; }
inline V3 g_cross(V3 a, V3 b) {
    return V3{a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y};
}
inline V3 g_reflect(V3 I, V3 N) { float d = g_dot(N, I); return I - N * d * 2.0f; }
inline float g_mix(float x, float y, float a) { return x * (1.0f - a) + y * a; }
inline float g_clamp(float x, float lo, float hi) { return g_min(g_max(x, lo), hi); }

enum : unsigned { NONE = 0, INNER = 1, OUTER = 2 };   // gpu_shader.comp:95-97

struct Ray { V3 start, dir; };                         // gpu_shader.comp:89-92

struct Hit {                                           // gpu_shader.comp:98-103 (+ shape id AOV)
    unsigned type = NONE;
    V3 point{0, 0, 0};
    V3 normal{0, 0, 0};
    const rt_flat_material* mat = nullptr;
    int shape = -1;
    float dist = 1e20f;
};

struct Scene {
    const rt_flat_light* light;
    const rt_flat_camera* cam;
    const rt_flat_shape* shapes; int nShapes;
    const rt_flat_node* nodes;   int nNodes;
    const int32_t* indices;      int nIndices;
    bool useMT;
};

struct Counters {
    uint64_t rays = 0, primary = 0, shadow = 0, reflect = 0, node_visits = 0, index_reads = 0;
    uint64_t tests[4] = {0, 0, 0, 0};
    uint64_t tri_stage2 = 0, ray_hits = 0;
};

inline V3 point_on_ray(const Ray& r, float t) { return r.start + t * r.dir; }   // :134-136

// gpu_shader.comp:170-195
unsigned tri_moller_trumbore(const rt_flat_shape& s, const Ray& ray, V3& hp) {
    V3 p1 = v3(s.triP1);
    V3 edge1 = v3(s.triP2) - p1;
    V3 edge2 = v3(s.triP3) - p1;
    V3 h = g_cross(ray.dir, edge2);
    float a = g_dot(edge1, h);
    if (std::fabs(a) < 1e-5f) return NONE;
    float f = 1.0f / a;
    V3 sv = ray.start - p1;
    float u = f * g_dot(sv, h);
    if (u < 0.0f || u > 1.0f) return NONE;
    V3 q = g_cross(sv, edge1);
    float v = f * g_dot(ray.dir, q);
    if (v < 0.0f || u + v > 1.0f) return NONE;
    float t = f * g_dot(edge2, q);
    if (t > 0.0f) { hp = point_on_ray(ray, t); return INNER; }
    return NONE;
}

// gpu_shader.comp:196-240 (twin: src/shapes/triangle.hpp:100-130)
unsigned tri_barycentric(const rt_flat_shape& s, const Ray& ray, V3& hp, Counters* c) {
    V3 n = v3(s.planeNormal);
    float np = g_dot(n, ray.dir);
    if (np == 0.0f) return NONE;
    float t = -(s.planeD + g_dot(n, ray.start)) / np;
    if (!(t > 0.0f)) return NONE;
    hp = point_on_ray(ray, t);
    if (np <= 0.0f) return OUTER;
    if (c) c->tri_stage2++;
    V3 p1 = v3(s.triP1);
    V3 edge1 = v3(s.triP2) - p1;
    V3 edge2 = v3(s.triP3) - p1;
    V3 toPoint = hp - p1;
    float d00 = g_dot(edge1, edge1);
    float d01 = g_dot(edge1, edge2);
    float d11 = g_dot(edge2, edge2);
    float d20 = g_dot(toPoint, edge1);
    float d21 = g_dot(toPoint, edge2);
    float denom = d00 * d11 - d01 * d01;
    float v = (d11 * d20 - d01 * d21) / denom;
    float w = (d00 * d21 - d01 * d20) / denom;
    float u = 1.0f - v - w;
    if (u < 0.0f || v < 0.0f || w < 0.0f) return NONE;
    return INNER;
}

// gpu_shader.comp:242-328
unsigned get_intersection(const rt_flat_shape& s, const Ray& ray, bool useMT, V3& hp, Counters* c) {
    if (c && s.type >= 0 && s.type <= 3) c->tests[s.type]++;
    if (s.type == RT_SHAPE_SPHERE) {                       // :246-271 (twin sphere.hpp:41-60)
        V3 ctr = v3(s.sphereCenter);
        float aa = g_dot(ray.dir, ray.dir);
        float bb = 2.0f * g_dot(ray.dir, ray.start - ctr);
        float cc = g_dot(ray.start - ctr, ray.start - ctr) - s.sphereRadius * s.sphereRadius;
        float D = bb * bb - 4.0f * aa * cc;
        if (D > 0.0f) {
            float sD = std::sqrt(D);
            float t1 = (-bb - sD) / (2.0f * aa);
            if (t1 > 0.0f) { hp = point_on_ray(ray, t1); return INNER; }
            float t2 = (-bb + sD) / (2.0f * aa);
            if (t2 > 0.0f) return OUTER;
        }
        return NONE;
    }
    if (s.type == RT_SHAPE_PLANE) {                        // :272-286 (twin plane.hpp:44-57)
        V3 n = v3(s.planeNormal);
        float np = g_dot(n, ray.dir);
        if (np == 0.0f) return NONE;
        float t = -(s.planeD + g_dot(n, ray.start)) / np;
        if (!(t > 0.0f)) return NONE;
        if (np <= 0.0f) return OUTER;
        hp = point_on_ray(ray, t);
        return INNER;
    }
    if (s.type == RT_SHAPE_WALL) {                         // :287-316 (twin wall.hpp:46-68)
        V3 n = v3(s.planeNormal);
        float np = g_dot(n, ray.dir);
        if (np == 0.0f) return NONE;
        float t = -(s.planeD + g_dot(n, ray.start)) / np;
        if (!(t > 0.0f)) return NONE;
        if (np <= 0.0f) return OUTER;
        hp = point_on_ray(ray, t);
        V3 u = g_normalize(g_cross(n, v3(0, 1, 0)));
        if (g_length(u) < 1e-5f) u = g_normalize(g_cross(n, v3(1, 0, 0)));   // false for NaN
        V3 v = g_normalize(g_cross(n, u));
        V3 local = hp - v3(s.wallStart);
        float uProj = g_dot(local, u);
        float vProj = g_dot(local, v);
        if (uProj < 0.0f || uProj > s.wallWidth || vProj < 0.0f || vProj > s.wallHeight) return NONE;
        return INNER;
    }
    if (s.type == RT_SHAPE_TRIANGLE) {                     // :317-325
        return useMT ? tri_moller_trumbore(s, ray, hp) : tri_barycentric(s, ray, hp, c);
    }
    return NONE;
}

// gpu_shader.comp:64-71
inline V3 normal_from_shape(const rt_flat_shape& s, V3 p) {
    if (s.type == RT_SHAPE_SPHERE) return g_normalize(p - v3(s.sphereCenter));
    return v3(s.planeNormal);
}

// gpu_shader.comp:364-377
inline bool ray_intersects_aabb(const Ray& ray, V3 bmin, V3 bmax) {
    V3 inv = v3(1.0f / ray.dir.x, 1.0f / ray.dir.y, 1.0f / ray.dir.z);
    V3 t0 = (bmin - ray.start) * inv;
    V3 t1 = (bmax - ray.start) * inv;
    V3 mn = v3(g_min(t0.x, t1.x), g_min(t0.y, t1.y), g_min(t0.z, t1.z));
    V3 mx = v3(g_max(t0.x, t1.x), g_max(t0.y, t1.y), g_max(t0.z, t1.z));
    float tMin = g_max(g_max(mn.x, mn.y), mn.z);
    float tMax = g_min(g_min(mx.x, mx.y), mx.z);
    return tMax >= tMin && tMax > 0.0f;
}

// closest-hit update shared by the BVH and brute-force paths (:408-419, :541-553)
inline void consider(const Scene& sc, int shapeIdx, const Ray& ray, Hit& best, Counters* c) {
    const rt_flat_shape& s = sc.shapes[shapeIdx];
    V3 hp{0, 0, 0};
    if (get_intersection(s, ray, sc.useMT, hp, c) == INNER) {
        float dist = g_distance(ray.start, hp);
        if (dist < best.dist) {
            best.dist = dist;
            best.type = INNER;
            best.point = hp;
            best.normal = normal_from_shape(s, hp);
            best.mat = &s.material;
            best.shape = shapeIdx;
        }
    }
}

// gpu_shader.comp:380-430.  An empty node buffer (root index -1, undefined in
// the reference) is defined here as "no hit".
Hit intersect_scene2(const Scene& sc, const Ray& ray, Counters* c) {
    Hit best;
    if (c) c->rays++;
    if (sc.nNodes <= 0) return best;
    int stack[64];
    int sp = 0;
    stack[sp++] = sc.nNodes - 1;
    while (sp > 0) {
        const rt_flat_node& node = sc.nodes[stack[--sp]];
        if (c) c->node_visits++;
        if (!ray_intersects_aabb(ray, v3(node.boundsMin), v3(node.boundsMax))) continue;
        if (node.leftChild == -1) {
            for (int i = 0; i < node.numShapes; ++i) {
                if (c) c->index_reads++;
                consider(sc, sc.indices[node.startShapeIdx + i], ray, best, c);
            }
        } else {
            stack[sp++] = node.leftChild;
            stack[sp++] = node.rightChild;
        }
    }
    if (c && best.type == INNER) c->ray_hits++;
    return best;
}

// gpu_shader.comp:331-361
V3 phong(V3 point, V3 normal, V3 viewDir, const rt_flat_light& L, const rt_flat_material& m) {
    V3 lpos = v3(L.position);
    float distToLight = g_distance(lpos, point);
    V3 lightColor = v3(L.color) / distToLight;
    V3 ambient = m.ambientStrength * lightColor;
    V3 lightDir = g_normalize(lpos - point);
    float diff = g_max(g_dot(normal, lightDir), 0.0f);
    V3 diffuse = (m.diffuseStrength * diff) * lightColor;
    V3 specular = v3(0, 0, 0);
    if (diff > 0.0f) {
        V3 reflectDir = g_reflect(-lightDir, normal);
        float spec = std::pow(g_max(g_dot(viewDir, reflectDir), 0.0f), (float)m.shininess);
        specular = (m.specularStrength * spec) * lightColor;
    }
    return ((ambient + diffuse) + specular) * v3(m.color);
}

// gpu_shader.comp:155-168 (twin: src/camera.hpp:124-137)
Ray get_ray(const rt_flat_camera& cam, float ndcX, float ndcY) {
    float h = 2.0f * std::tan((cam.fov / 2.0f) * 0.017453292519943295f);
    float w = h * cam.aspectRatio;
    V3 P = v3(cam.Position);
    V3 ipp = ((P + v3(cam.Front)) + (ndcX * w / 2.0f) * v3(cam.Right)) + (ndcY * h / 2.0f) * v3(cam.Up);
    Ray r;
    r.start = P;
    r.dir = g_normalize(ipp - P);
    return r;
}

struct PixelOut { float rgba[4]; int id; float dist; };

// reflection / Fresnel step shared by both paths (:495-516, :595-616).
// Returns false when the bounce loop must stop.
inline bool bounce(Ray& ray, V3 hitPoint, V3 hitNormal, const rt_flat_material& m, V3 phongColor,
                   bool useFresnel, V3& attenuation, V3& acc) {
    if (!(m.specularStrength > 0.0f)) return false;
    V3 reflectDir = g_reflect(ray.dir, hitNormal);
    ray.start = hitPoint + hitNormal * 1e-3f;
    ray.dir = reflectDir;
    if (useFresnel) {
        float fresnel = std::pow(1.0f - g_max(g_dot(-ray.dir, hitNormal), 0.0f), 5.0f);
        fresnel = g_clamp(fresnel, 0.0f, 0.8f);
        float rw = m.fresnelStrength * fresnel;
        float mw = 1.0f - rw;
        V3 hc = v3(m.color);
        attenuation = attenuation * v3(g_mix(hc.x, 1.0f, rw), g_mix(hc.y, 1.0f, rw), g_mix(hc.z, 1.0f, rw));
        acc = acc + (mw * hc) * phongColor;
    } else {
        attenuation = attenuation * m.specularStrength;
    }
    return true;
}

// gpu_shader.comp:433-624
void shade_pixel(const Scene& sc, const rt_params& p, int x, int y, PixelOut& out, Counters* c) {
    const float W = p.screenRes[0], H = p.screenRes[1];
    float a = (float)y / H;
    V3 bg = v3(g_mix(0.05f, 0.5f, a), g_mix(0.07f, 0.7f, a), g_mix(0.1f, 1.0f, a));   // :436
    Ray ray = get_ray(*sc.cam, 2.0f * (float)x / W - 1.0f, 1.0f - 2.0f * (float)y / H);   // :440-443
    V3 acc = v3(0, 0, 0), att = v3(1, 1, 1);
    const V3 lpos = v3(sc.light->position);
    out.id = -1;
    out.dist = 1e20f;

    for (int depth = 0; depth < p.maxBounces; ++depth) {
        Hit hit;
        if (p.useBVH) {                                            // :450-517
            hit = intersect_scene2(sc, ray, c);
            if (c) { if (depth == 0) c->primary++; else c->reflect++; }
        } else {                                                   // :527-554
            if (c) { c->rays++; if (depth == 0) c->primary++; else c->reflect++; }
            for (int i = 0; i < sc.nShapes; ++i) consider(sc, i, ray, hit, c);
            if (c && hit.type == INNER) c->ray_hits++;
        }
        if (depth == 0) { out.id = hit.shape; out.dist = hit.dist; }
        if (hit.type != INNER) { acc = acc + att * bg; break; }    // :454-458, :556-560

        V3 hitPoint = hit.point, hitNormal = hit.normal;
        const rt_flat_material& m = *hit.mat;

        bool inShadow = false;
        Ray sh;
        sh.dir = g_normalize(lpos - hitPoint);
        if (p.useBVH) {                                            // :466-480
            sh.start = hitPoint + hitNormal * 1e-3f;
            Hit shHit = intersect_scene2(sc, sh, c);
            if (c) c->shadow++;
            if (shHit.type != NONE) {
                float lightDist = g_distance(lpos, hitPoint);
                float hitDist = g_distance(sh.start, shHit.point);
                if (hitDist < lightDist) inShadow = true;
            }
        } else {                                                   // :562-580
            sh.start = hitPoint + hitNormal * 1e-5f;
            if (c) { c->rays++; c->shadow++; }
            for (int i = 0; i < sc.nShapes; ++i) {
                V3 hp{0, 0, 0};
                if (get_intersection(sc.shapes[i], sh, sc.useMT, hp, c) == INNER) {
                    float lightDist = g_distance(lpos, hitPoint);
                    float hitDist = g_distance(sh.start, hp);
                    if (hitDist < lightDist) { inShadow = true; break; }
                }
            }
        }

        V3 phongColor = phong(hitPoint, hitNormal, ray.dir, *sc.light, m);   // :483-488
        if (inShadow) phongColor = phongColor * 0.3f;                        // :491
        acc = acc + att * phongColor;                                        // :493
        if (!bounce(ray, hitPoint, hitNormal, m, phongColor, p.useFresnel != 0, att, acc)) break;
    }
    out.rgba[0] = acc.x; out.rgba[1] = acc.y; out.rgba[2] = acc.z; out.rgba[3] = 1.0f;   // :437,:519,:623
}

}  // namespace

// ---------------------------------------------------------------------------
// C interface
// ---------------------------------------------------------------------------
extern "C" {

int oracle_render(const rt_flat_light* light, const rt_flat_camera* camera,
                  const rt_flat_shape* shapes, int nShapes,
                  const rt_flat_node* nodes, int nNodes,
                  const int32_t* indices, int nIndices,
                  const rt_params* params, int W, int H,
                  int row_step, int row_phase,
                  float* rgba, int32_t* ids, float* dists,
                  oracle_counters* counters, int threads) {
    if (!light || !camera || !params || !rgba || W <= 0 || H <= 0 || row_step <= 0) return -1;
    if ((nShapes > 0 && !shapes) || (nNodes > 0 && !nodes) || (nIndices > 0 && !indices)) return -1;
    Scene sc{light, camera, shapes, nShapes, nodes, nNodes, indices, nIndices, params->useMollerTrumbore != 0};
    std::vector<int> rows;
    for (int y = 0; y < H; ++y) if (y % row_step == row_phase) rows.push_back(y);
    if (threads < 1) threads = 1;
    std::vector<Counters> cs((size_t)threads);
    const bool count = counters != nullptr;
    auto work = [&](int tid) {
        Counters* c = count ? &cs[(size_t)tid] : nullptr;
        for (size_t r = (size_t)tid; r < rows.size(); r += (size_t)threads) {
            int y = rows[r];
            for (int x = 0; x < W; ++x) {
                PixelOut o;
                shade_pixel(sc, *params, x, y, o, c);
                size_t px = (size_t)y * (size_t)W + (size_t)x;
                std::memcpy(rgba + 4 * px, o.rgba, sizeof o.rgba);
                if (ids) ids[px] = o.id;
                if (dists) dists[px] = o.dist;
            }
        }
    };
    if (threads == 1) {
        work(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& t : pool) t.join();
    }
    if (counters) {
        std::memset(counters, 0, sizeof *counters);
        for (const Counters& c : cs) {
            counters->rays += c.rays; counters->primary_rays += c.primary; counters->shadow_rays += c.shadow;
            counters->reflect_rays += c.reflect; counters->node_visits += c.node_visits;
            counters->index_reads += c.index_reads;
            for (int k = 0; k < 4; ++k) counters->tests[k] += c.tests[k];
            counters->tri_stage2 += c.tri_stage2; counters->ray_hits += c.ray_hits;
        }
        // b(sphere)=20, b(plane)=20, b(wall)=40, b(tri,MT)=40,
        // b(tri,Bary)=20 rejected at the plane stage else 56; +48 B per hit ray.
        const uint64_t tri = counters->tests[3];
        const uint64_t tri_bytes = sc.useMT ? 40 * tri : 20 * tri + 36 * counters->tri_stage2;
        counters->alg_bytes = 48 * counters->node_visits + 4 * counters->index_reads +
                              20 * counters->tests[0] + 20 * counters->tests[1] + 40 * counters->tests[2] +
                              tri_bytes + 48 * counters->ray_hits;
    }
    return 0;
}

int oracle_intersect_shape(const rt_flat_shape* shape, const float start[3], const float dir[3],
                           int useMollerTrumbore, float hit_point[3]) {
    Ray r{v3(start), v3(dir)};
    V3 hp{0, 0, 0};
    unsigned t = get_intersection(*shape, r, useMollerTrumbore != 0, hp, nullptr);
    if (t == INNER && hit_point) { hit_point[0] = hp.x; hit_point[1] = hp.y; hit_point[2] = hp.z; }
    return (int)t;
}

int oracle_intersect_scene(const rt_flat_shape* shapes, int nShapes,
                           const rt_flat_node* nodes, int nNodes,
                           const int32_t* indices, int nIndices,
                           const float start[3], const float dir[3], int useMollerTrumbore,
                           float hit_point[3], float hit_normal[3], float* dist) {
    Scene sc{nullptr, nullptr, shapes, nShapes, nodes, nNodes, indices, nIndices, useMollerTrumbore != 0};
    Ray r{v3(start), v3(dir)};
    Hit h = intersect_scene2(sc, r, nullptr);
    if (h.type == INNER) {
        if (hit_point) { hit_point[0] = h.point.x; hit_point[1] = h.point.y; hit_point[2] = h.point.z; }
        if (hit_normal) { hit_normal[0] = h.normal.x; hit_normal[1] = h.normal.y; hit_normal[2] = h.normal.z; }
    }
    if (dist) *dist = h.dist;
    return h.shape;
}

void oracle_intersect_scene_batch(const rt_flat_shape* shapes, int nShapes, const rt_flat_node* nodes, int nNodes,
                                  const int32_t* indices, int nIndices, int n, const float* starts, const float* dirs,
                                  int useMollerTrumbore, int32_t* out_shape, float* out_dist, float* out_point) {
    Scene sc{nullptr, nullptr, shapes, nShapes, nodes, nNodes, indices, nIndices, useMollerTrumbore != 0};
    for (int i = 0; i < n; ++i) {
        Ray r{v3(starts + 3 * (size_t)i), v3(dirs + 3 * (size_t)i)};
        Hit h = intersect_scene2(sc, r, nullptr);
        out_shape[i] = h.shape;
        out_dist[i] = h.dist;
        if (out_point) {
            const bool in = h.type == INNER;
            out_point[3 * (size_t)i] = in ? h.point.x : 0.f; out_point[3 * (size_t)i + 1] = in ? h.point.y : 0.f; out_point[3 * (size_t)i + 2] = in ? h.point.z : 0.f;
        }
    }
}

// Tie pixels (north_star: "pixels whose t-values tie within 1 ulp ... are counted and reported"): the primary
// ray's INNER candidates in the reference's visit order (:390-427); a pixel is a tie pixel when a candidate of a
// DIFFERENT shape lies within 1 ulp of the winning closestDist (:410-411 decides those by visit order alone).
int64_t oracle_primary_ties(const rt_flat_camera* camera, const rt_flat_shape* shapes, int nShapes,
                            const rt_flat_node* nodes, int nNodes, const int32_t* indices, int nIndices,
                            const rt_params* params, int W, int H, int row_step, int row_phase,
                            uint8_t* ties, int threads) {
    if (!camera || !params || W <= 0 || H <= 0 || row_step <= 0) return -1;
    Scene sc{nullptr, camera, shapes, nShapes, nodes, nNodes, indices, nIndices, params->useMollerTrumbore != 0};
    std::vector<int> rows;
    for (int y = 0; y < H; ++y) if (y % row_step == row_phase) rows.push_back(y);
    if (threads < 1) threads = 1;
    std::vector<int64_t> found((size_t)threads, 0);
    auto ulps = [](float a, float b) {
        int32_t ia, ib; std::memcpy(&ia, &a, 4); std::memcpy(&ib, &b, 4);
        return ia > ib ? (int64_t)ia - ib : (int64_t)ib - ia;      // both distances are positive floats
    };
    auto work = [&](int tid) {
        std::vector<std::pair<float, int>> cand;
        for (size_t r = (size_t)tid; r < rows.size(); r += (size_t)threads) {
            const int y = rows[r];
            for (int x = 0; x < W; ++x) {
                Ray ray = get_ray(*sc.cam, 2.0f * (float)x / params->screenRes[0] - 1.0f, 1.0f - 2.0f * (float)y / params->screenRes[1]);
                cand.clear();
                auto test = [&](int shapeIdx) {
                    V3 hp{0, 0, 0};
                    if (get_intersection(sc.shapes[shapeIdx], ray, sc.useMT, hp, nullptr) == INNER)
                        cand.emplace_back(g_distance(ray.start, hp), shapeIdx);
                };
                if (params->useBVH) {
                    int stack[64]; int sp = 0;
                    if (sc.nNodes > 0) stack[sp++] = sc.nNodes - 1;
                    while (sp > 0) {
                        const rt_flat_node& node = sc.nodes[stack[--sp]];
                        if (!ray_intersects_aabb(ray, v3(node.boundsMin), v3(node.boundsMax))) continue;
                        if (node.leftChild == -1) { for (int i = 0; i < node.numShapes; ++i) test(sc.indices[node.startShapeIdx + i]); }
                        else { stack[sp++] = node.leftChild; stack[sp++] = node.rightChild; }
                    }
                } else {
                    for (int i = 0; i < sc.nShapes; ++i) test(i);
                }
                float best = 1e20f; int who = -1;
                for (auto& c : cand) if (c.first < best) { best = c.first; who = c.second; }
                bool tie = false;
                if (who >= 0) for (auto& c : cand) if (c.second != who && c.first < 1e20f && ulps(c.first, best) <= 1) tie = true;
                if (ties) ties[(size_t)y * (size_t)W + (size_t)x] = tie ? 1 : 0;
                found[(size_t)tid] += tie ? 1 : 0;
            }
        }
    };
    if (threads == 1) work(0);
    else {
        std::vector<std::thread> pool;
        for (int t = 0; t < threads; ++t) pool.emplace_back(work, t);
        for (auto& t : pool) t.join();
    }
    int64_t n = 0;
    for (int64_t f : found) n += f;
    return n;
}

void oracle_present(const float* rgba, uint8_t* rgba8, int64_t pixels) {
    for (int64_t i = 0; i < pixels; ++i) {
        for (int k = 0; k < 4; ++k) {
            float v = k == 3 ? 1.0f : rgba[4 * i + k];
            v = g_min(g_max(v, 0.0f), 1.0f);          // NaN -> NaN by GLSL select; map NaN to 0 below
            if (!(v == v)) v = 0.0f;
            rgba8[4 * i + k] = (uint8_t)(int)(v * 255.0f + 0.5f);
        }
    }
}

}  // extern "C"
