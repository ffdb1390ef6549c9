// TEST INFRASTRUCTURE ONLY (see oracle/README.md): a brute-force stand-in for the handful of Embree 3 entry points the
// reference's instancer/src/instancer.cpp calls, so that the reference's own C++ (hit sorting, patch toggling, segment
// walk, sample placement, parameter textures, shadows, mesh shading, patch distribution) can be compiled and run here
// without Embree.  Everything in this file is OUR code, not the reference's and not Embree's: there is no BVH (every
// query visits every primitive, in geometry-id order), intersections are plain Moeller-Trumbore in double precision.
// What it keeps of Embree's documented behaviour: the ray / hit / context structs' field layout, closest-hit semantics of
// rtcIntersect1 with a context filter that may reject a hit (valid[0] = 0) and sees the candidate distance in
// ray->tfar, any-hit semantics of rtcOccluded1 (tfar = -inf), instance transforms (the filter sees the ray and the
// normal in the instance's local space), quads split into the triangles (v0, v1, v3) and (v2, v3, v1), the geometric
// normal Ng = (v1 - v0) x (v2 - v0), barycentrics P = (1 - u - v) v0 + u v1 + v v2, sequential geometry ids.
#pragma once

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#define RTC_INVALID_GEOMETRY_ID ((unsigned int)-1)
#define RTC_MAX_INSTANCE_LEVEL_COUNT 1

enum RTCError { RTC_ERROR_NONE = 0, RTC_ERROR_UNKNOWN = 1, RTC_ERROR_INVALID_ARGUMENT = 2, RTC_ERROR_INVALID_OPERATION = 3 };
enum RTCSceneFlags { RTC_SCENE_FLAG_NONE = 0, RTC_SCENE_FLAG_DYNAMIC = 1, RTC_SCENE_FLAG_COMPACT = 2, RTC_SCENE_FLAG_ROBUST = 4,
                     RTC_SCENE_FLAG_CONTEXT_FILTER_FUNCTION = 8 };
inline RTCSceneFlags operator|(RTCSceneFlags a, RTCSceneFlags b) { return (RTCSceneFlags)((int)a | (int)b); }
enum RTCGeometryType { RTC_GEOMETRY_TYPE_TRIANGLE = 0, RTC_GEOMETRY_TYPE_QUAD = 1, RTC_GEOMETRY_TYPE_INSTANCE = 121 };
enum RTCBufferType { RTC_BUFFER_TYPE_INDEX = 0, RTC_BUFFER_TYPE_VERTEX = 1 };
enum RTCFormat { RTC_FORMAT_UINT3 = 0x5003, RTC_FORMAT_UINT4 = 0x5004, RTC_FORMAT_FLOAT3 = 0x9003, RTC_FORMAT_FLOAT3X4_ROW_MAJOR = 0x9134 };

struct RTCRay { float org_x, org_y, org_z, tnear, dir_x, dir_y, dir_z, time, tfar; unsigned int mask, id, flags; };
struct RTCHit { float Ng_x, Ng_y, Ng_z, u, v; unsigned int primID, geomID, instID[RTC_MAX_INSTANCE_LEVEL_COUNT]; };
struct RTCRayHit { RTCRay ray; RTCHit hit; };
struct RTCRayN;
struct RTCHitN;
struct RTCIntersectContext;
struct RTCFilterFunctionNArguments {
    int* valid; void* geometryUserPtr; RTCIntersectContext* context; RTCRayN* ray; RTCHitN* hit; unsigned int N;
};
typedef void (*RTCFilterFunctionN)(const RTCFilterFunctionNArguments* args);
struct RTCIntersectContext { int flags; RTCFilterFunctionN filter; unsigned int instID[RTC_MAX_INSTANCE_LEVEL_COUNT]; };
inline void rtcInitIntersectContext(RTCIntersectContext* c) { c->flags = 0; c->filter = nullptr; c->instID[0] = RTC_INVALID_GEOMETRY_ID; }

struct RTCPointQuery { float x, y, z, time, radius; };
struct RTCPointQueryContext { unsigned int instStackSize; };
inline void rtcInitPointQueryContext(RTCPointQueryContext* c) { c->instStackSize = 0; }
struct RTCPointQueryFunctionArguments {
    RTCPointQuery* query; void* userPtr; unsigned int primID, geomID; RTCPointQueryContext* context; float similarityScale;
};
typedef bool (*RTCPointQueryFunction)(RTCPointQueryFunctionArguments* args);

struct ShimScene;
struct ShimGeometry {
    RTCGeometryType type;
    const char* vtx = nullptr; size_t vtx_stride = 0, n_vtx = 0;
    const char* idx = nullptr; size_t idx_stride = 0, n_prim = 0;
    ShimScene* instanced = nullptr;
    double fwd[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0}, inv[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};   // 3x4 row-major
};
struct ShimScene { std::vector<ShimGeometry*> geoms; };       // index = geometry id (nullptr: id not in use)
struct ShimDevice { int unused; };
typedef ShimDevice* RTCDevice;
typedef ShimScene* RTCScene;
typedef ShimGeometry* RTCGeometry;
typedef void (*RTCErrorFunction)(void* userPtr, enum RTCError code, const char* str);

inline RTCDevice rtcNewDevice(const char*) { return new ShimDevice(); }
inline RTCError rtcGetDeviceError(RTCDevice) { return RTC_ERROR_NONE; }
inline void rtcSetDeviceErrorFunction(RTCDevice, RTCErrorFunction, void*) {}
inline RTCScene rtcNewScene(RTCDevice) { return new ShimScene(); }
inline void rtcSetSceneFlags(RTCScene, RTCSceneFlags) {}
inline void rtcCommitScene(RTCScene) {}
inline void rtcReleaseScene(RTCScene) {}                      // (objects are never freed: a short-lived test tool)
inline RTCGeometry rtcNewGeometry(RTCDevice, RTCGeometryType type) { ShimGeometry* g = new ShimGeometry(); g->type = type; return g; }
inline void rtcCommitGeometry(RTCGeometry) {}
inline void rtcReleaseGeometry(RTCGeometry) {}
inline void* rtcSetNewGeometryBuffer(RTCGeometry g, RTCBufferType type, unsigned int, RTCFormat, size_t byteStride, size_t itemCount) {
    char* p = (char*)calloc(itemCount + 1, byteStride);
    if (type == RTC_BUFFER_TYPE_VERTEX) { g->vtx = p; g->vtx_stride = byteStride; g->n_vtx = itemCount; }
    else { g->idx = p; g->idx_stride = byteStride; g->n_prim = itemCount; }
    return p;
}
inline void rtcSetSharedGeometryBuffer(RTCGeometry g, RTCBufferType type, unsigned int, RTCFormat, const void* ptr, size_t byteOffset,
                                       size_t byteStride, size_t itemCount) {
    const char* p = (const char*)ptr + byteOffset;
    if (type == RTC_BUFFER_TYPE_VERTEX) { g->vtx = p; g->vtx_stride = byteStride; g->n_vtx = itemCount; }
    else { g->idx = p; g->idx_stride = byteStride; g->n_prim = itemCount; }
}
inline unsigned int rtcAttachGeometry(RTCScene s, RTCGeometry g) { s->geoms.push_back(g); return (unsigned int)s->geoms.size() - 1; }
inline void rtcAttachGeometryByID(RTCScene s, RTCGeometry g, unsigned int id) {
    if (s->geoms.size() <= id) s->geoms.resize((size_t)id + 1, nullptr);
    s->geoms[id] = g;
}
inline void rtcSetGeometryInstancedScene(RTCGeometry g, RTCScene s) { g->instanced = s; }
inline void rtcSetGeometryTransform(RTCGeometry g, unsigned int, RTCFormat, const void* xfm) {
    const float* m = (const float*)xfm;                        // 3x4 row-major, local -> world
    for (int i = 0; i < 12; ++i) g->fwd[i] = m[i];
    const double a = m[0], b = m[1], c = m[2], d = m[4], e = m[5], f = m[6], gg = m[8], h = m[9], k = m[10];
    const double det = a * (e * k - f * h) - b * (d * k - f * gg) + c * (d * h - e * gg);
    const double r[9] = {(e * k - f * h) / det, (c * h - b * k) / det, (b * f - c * e) / det,
                         (f * gg - d * k) / det, (a * k - c * gg) / det, (c * d - a * f) / det,
                         (d * h - e * gg) / det, (b * gg - a * h) / det, (a * e - b * d) / det};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) g->inv[4 * i + j] = r[3 * i + j];
        g->inv[4 * i + 3] = -(r[3 * i] * m[3] + r[3 * i + 1] * m[7] + r[3 * i + 2] * m[11]);
    }
}

namespace embree_shim {

struct Vec { double x, y, z; };
inline Vec sub(Vec a, Vec b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec cross(Vec a, Vec b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline double dot(Vec a, Vec b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec vertex(const ShimGeometry* g, unsigned int i) {
    const float* p = (const float*)(g->vtx + (size_t)i * g->vtx_stride);
    return {p[0], p[1], p[2]};
}

// ray / triangle (v0, v1, v2): t, barycentrics (u for v1, v for v2), Ng = (v1 - v0) x (v2 - v0)
inline bool tri(Vec o, Vec d, Vec v0, Vec v1, Vec v2, double& t, double& u, double& v, Vec& ng) {
    const Vec e1 = sub(v1, v0), e2 = sub(v2, v0);
    const Vec p = cross(d, e2);
    const double det = dot(e1, p);
    if (det == 0.0) return false;
    const Vec s = sub(o, v0);
    u = dot(s, p) / det;
    if (u < 0.0 || u > 1.0) return false;
    const Vec q = cross(s, e1);
    v = dot(d, q) / det;
    if (v < 0.0 || u + v > 1.0) return false;
    t = dot(e2, q) / det;
    ng = cross(e1, e2);
    return true;
}

struct Candidate { double t, u, v; Vec ng; unsigned int prim; };

// every hit of the (local) ray with the primitives of one triangle / quad geometry
inline void collect(const ShimGeometry* g, Vec o, Vec d, std::vector<Candidate>& out) {
    for (unsigned int p = 0; p < g->n_prim; ++p) {
        const unsigned int* ix = (const unsigned int*)(g->idx + (size_t)p * g->idx_stride);
        Candidate c;
        c.prim = p;
        if (g->type == RTC_GEOMETRY_TYPE_TRIANGLE) {
            if (tri(o, d, vertex(g, ix[0]), vertex(g, ix[1]), vertex(g, ix[2]), c.t, c.u, c.v, c.ng)) out.push_back(c);
        } else {                                               // quad = triangles (v0, v1, v3) and (v2, v3, v1)
            const Vec v0 = vertex(g, ix[0]), v1 = vertex(g, ix[1]), v2 = vertex(g, ix[2]), v3 = vertex(g, ix[3]);
            if (tri(o, d, v0, v1, v3, c.t, c.u, c.v, c.ng)) out.push_back(c);
            else if (tri(o, d, v2, v3, v1, c.t, c.u, c.v, c.ng)) { c.u = 1.0 - c.u; c.v = 1.0 - c.v; out.push_back(c); }
        }
    }
}

// returns true when traversal must stop (occlusion query satisfied)
inline bool visit(const ShimGeometry* g, unsigned int geom_id, unsigned int inst_id, const RTCRay& world, Vec o, Vec d,
                  RTCIntersectContext* ctx, RTCRayHit* rh, RTCRay* occ) {
    std::vector<Candidate> cands;
    collect(g, o, d, cands);
    for (const Candidate& c : cands) {
        RTCRay& live = rh ? rh->ray : *occ;
        const float t = (float)c.t;
        if (!(t >= live.tnear && t <= live.tfar)) continue;
        RTCRay r = world;                                      // what the filter sees: the ray in the geometry's space
        r.org_x = (float)o.x; r.org_y = (float)o.y; r.org_z = (float)o.z;
        r.dir_x = (float)d.x; r.dir_y = (float)d.y; r.dir_z = (float)d.z;
        r.tnear = live.tnear; r.tfar = t;
        RTCHit h;
        h.Ng_x = (float)c.ng.x; h.Ng_y = (float)c.ng.y; h.Ng_z = (float)c.ng.z;
        h.u = (float)c.u; h.v = (float)c.v;
        h.primID = c.prim; h.geomID = geom_id; h.instID[0] = inst_id;
        int valid = -1;
        if (ctx && ctx->filter) {
            RTCFilterFunctionNArguments a = {&valid, nullptr, ctx, (RTCRayN*)&r, (RTCHitN*)&h, 1};
            ctx->filter(&a);
        }
        if (valid == 0) continue;                              // rejected: traversal goes on
        if (occ) { occ->tfar = -std::numeric_limits<float>::infinity(); return true; }
        rh->ray.tfar = t;
        rh->hit = h;
    }
    return false;
}

inline void traverse(ShimScene* s, RTCIntersectContext* ctx, RTCRayHit* rh, RTCRay* occ) {
    const RTCRay world = rh ? rh->ray : *occ;
    const Vec o = {world.org_x, world.org_y, world.org_z}, d = {world.dir_x, world.dir_y, world.dir_z};
    for (unsigned int id = 0; id < s->geoms.size(); ++id) {
        const ShimGeometry* g = s->geoms[id];
        if (!g) continue;
        if (g->type == RTC_GEOMETRY_TYPE_INSTANCE) {
            const double* m = g->inv;
            const Vec ol = {m[0] * o.x + m[1] * o.y + m[2] * o.z + m[3], m[4] * o.x + m[5] * o.y + m[6] * o.z + m[7],
                            m[8] * o.x + m[9] * o.y + m[10] * o.z + m[11]};
            const Vec dl = {m[0] * d.x + m[1] * d.y + m[2] * d.z, m[4] * d.x + m[5] * d.y + m[6] * d.z,
                            m[8] * d.x + m[9] * d.y + m[10] * d.z};
            for (unsigned int id2 = 0; id2 < g->instanced->geoms.size(); ++id2)
                if (g->instanced->geoms[id2] && visit(g->instanced->geoms[id2], id2, id, world, ol, dl, ctx, rh, occ)) return;
        } else if (visit(g, id, RTC_INVALID_GEOMETRY_ID, world, o, d, ctx, rh, occ)) {
            return;
        }
    }
}

}  // namespace embree_shim

inline void rtcIntersect1(RTCScene s, RTCIntersectContext* ctx, RTCRayHit* rh) { embree_shim::traverse(s, ctx, rh, nullptr); }
inline void rtcOccluded1(RTCScene s, RTCIntersectContext* ctx, RTCRay* ray) { embree_shim::traverse(s, ctx, nullptr, ray); }
inline bool rtcPointQuery(RTCScene s, RTCPointQuery* q, RTCPointQueryContext* c, RTCPointQueryFunction fn, void* userPtr) {
    bool changed = false;
    for (unsigned int id = 0; id < s->geoms.size(); ++id) {
        const ShimGeometry* g = s->geoms[id];
        if (!g || g->type == RTC_GEOMETRY_TYPE_INSTANCE) continue;
        for (unsigned int p = 0; p < g->n_prim; ++p) {
            RTCPointQueryFunctionArguments a = {q, userPtr, p, id, c, 1.0f};
            changed |= fn(&a);
        }
    }
    return changed;
}
