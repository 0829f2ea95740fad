// cull_probe.cpp — EXPERIMENT (host only, estimate): how many node visits would distance culling save,
// in list order and with an ideal near-child-first order?  Builds on the test emulator.
//   g++ -std=c++17 -O2 -ffp-contract=off -fPIC -shared -I/usr/local/cuda/include -o build_var/libcull_probe.so tools/cull_probe.cpp
#include "../tests/emu/walk_emu.cpp"

#include <algorithm>
#include <cmath>

namespace {
struct PR { int slot; float dist; V3 p; };

inline bool slab_t(const Ray& r, const float4& a, const float4& b, float& tmin) {
    float t0 = 0.f, t1 = 3e38f;
    const float o[3] = {r.o.x, r.o.y, r.o.z}, d[3] = {r.d.x, r.d.y, r.d.z};
    const float lo[3] = {a.x, a.y, a.z}, hi[3] = {b.x, b.y, b.z};
    for (int i = 0; i < 3; ++i) {
        float inv = 1.0f / d[i];
        float ta = (lo[i] - o[i]) * inv, tb = (hi[i] - o[i]) * inv;
        if (ta > tb) std::swap(ta, tb);
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
    }
    tmin = t0;
    return t0 <= t1;
}

struct Walker {
    const Emu& E; const std::vector<float4>& N; const std::vector<int>& CW; const std::vector<int>& LNK;
    Ray r; float best; int slot; V3 bp; bool any; uint64_t visits = 0, prims = 0; bool done = false; float len;
    int end_of(int k) const { return CW[(size_t)k] < 0 ? LNK[(size_t)k] : k + 1; }
    void leaf(int k) {
        const int cw = CW[(size_t)k], lnk = LNK[(size_t)k], c = cw & RT_LEAF_COUNT_MASK;
        prims += (uint64_t)c;
        for (int s = lnk; s < lnk + c; ++s) {
            V3 hp;
            if (!test_prim<false>(E.prims[(size_t)s], r, hp)) continue;
            const float dist = distance3(r.o, hp);
            if (dist < best) { best = dist; slot = s; bp = hp; if (any) { done = true; return; } }
        }
    }
    // mode 0: list order, no culling; 1: list order + culling; 2: near child first + culling
    void list(int mode) {
        const int n = (int)CW.size();
        int k = 0;
        while (k < n && !done) {
            float tm; visits++;
            bool hit = slab_t(r, N[2 * (size_t)k], N[2 * (size_t)k + 1], tm);
            if (mode == 1 && hit && tm * len > best * 1.0001f + 1e-3f) hit = false;
            if (CW[(size_t)k] < 0) { k = hit ? k + 1 : LNK[(size_t)k]; continue; }
            if (hit) leaf(k);
            k++;
        }
    }
    void ordered(int k) {      // node k's box already passed
        if (done) return;
        if (CW[(size_t)k] >= 0) { leaf(k); return; }
        int ch[64]; float tm[64]; int nc = 0;
        for (int c = k + 1; c < end_of(k) && nc < 64; c = end_of(c)) {
            visits++;
            float t;
            if (slab_t(r, N[2 * (size_t)c], N[2 * (size_t)c + 1], t)) { ch[nc] = c; tm[nc] = t; nc++; }
        }
        int idx[64]; for (int i = 0; i < nc; ++i) idx[i] = i;
        std::sort(idx, idx + nc, [&](int a, int b) { return tm[a] < tm[b]; });
        for (int i = 0; i < nc && !done; ++i) {
            if (tm[idx[i]] * len > best * 1.0001f + 1e-3f) continue;
            ordered(ch[idx[i]]);
        }
    }
};

PR run(const Emu& E, V3 o, V3 d, bool any, float bound, int mode, uint64_t* cnt) {
    Ray r = make_ray(o, d, E.boxesFinite, E.omax);
    const bool useB = r.sub && !E.nodesB.empty();
    Walker w{E, useB ? E.nodesB : E.nodes, useB ? E.L.cwB : E.L.cw, useB ? E.L.linkB : E.L.link, r, any ? bound : 1e20f, -1, mk(0.f, 0.f, 0.f), any};
    w.len = std::sqrt(d.x * d.x + d.y * d.y + d.z * d.z);
    if (mode < 2) w.list(mode);
    else {
        // the list is a forest at top level: roots in sequence
        const int n = (int)w.CW.size();
        for (int k = 0; k < n && !w.done; k = w.end_of(k)) {
            float t; w.visits++;
            if (slab_t(r, w.N[2 * (size_t)k], w.N[2 * (size_t)k + 1], t) && !(t * w.len > w.best * 1.0001f + 1e-3f)) w.ordered(k);
        }
    }
    cnt[0] += w.visits; cnt[1] += w.prims;
    return PR{w.slot, w.best, w.bp};
}
}  // namespace

extern "C" void probe_frame(void* e, const rt_flat_shape* shapes, int n, const float* origins, const float* dirs,
                            const float* light, int bounces, int mode, uint64_t out[8]) {
    const Emu& E = *static_cast<Emu*>(e);
    for (int i = 0; i < 8; ++i) out[i] = 0;
    const V3 lpos = mk(light[0], light[1], light[2]);
    for (int i = 0; i < n; ++i) {
        V3 o = mk(origins + 3 * i), d = mk(dirs + 3 * i);
        for (int depth = 0; depth < bounces; ++depth) {
            PR h = run(E, o, d, false, 0.f, mode, out + (depth == 0 ? 0 : 4));
            out[depth == 0 ? 2 : 6]++;
            if (h.slot < 0) break;
            const rt_flat_shape* sh = shapes + E.L.pshape[(size_t)h.slot];
            V3 nrm = sh->type == RT_SHAPE_SPHERE ? normalize3(h.p - mk(sh->sphereCenter)) : mk(sh->planeNormal);
            const float ld = distance3(lpos, h.p);
            run(E, h.p + nrm * 1e-3f, normalize3(lpos - h.p), true, ld, mode, out + 4);
            out[6]++;
            if (!(sh->material.specularStrength > 0.f)) break;
            d = reflect3(d, nrm); o = h.p + nrm * 1e-3f;
        }
    }
}
