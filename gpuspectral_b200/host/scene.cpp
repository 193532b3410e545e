// scene.cpp — Scene methods, matrix helpers, built-in meshes and procedural scenes.
#include "scene.h"

#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>

namespace gsbhost {

// ---- matrices (column-major) -------------------------------------------------------------
void matIdentity(float* m) {
    for (int i = 0; i < 16; i++) m[i] = (i % 5 == 0) ? 1.0f : 0.0f;
}
void matMul(const float* a, const float* b, float* out) {
    float r[16];
    for (int c = 0; c < 4; c++)
        for (int row = 0; row < 4; row++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += (double)a[k * 4 + row] * (double)b[c * 4 + k];
            r[c * 4 + row] = (float)s;
        }
    std::memcpy(out, r, sizeof(r));
}
void matTranslate(float x, float y, float z, float* out) {
    matIdentity(out);
    out[12] = x; out[13] = y; out[14] = z;
}
void matScale(float x, float y, float z, float* out) {
    matIdentity(out);
    out[0] = x; out[5] = y; out[10] = z;
}
void matRotate(float ax, float ay, float az, float degrees, float* out) {
    double len = std::sqrt((double)ax * ax + (double)ay * ay + (double)az * az);
    matIdentity(out);
    if (len == 0.0) return;
    const double x = ax / len, y = ay / len, z = az / len;
    const double th = degrees * 3.14159265358979323846 / 180.0;
    const double c = std::cos(th), s = std::sin(th), t = 1.0 - c;
    // Rodrigues rotation matrix R(row, col), stored column-major
    const double R[3][3] = {{t * x * x + c, t * x * y - s * z, t * x * z + s * y},
                            {t * x * y + s * z, t * y * y + c, t * y * z - s * x},
                            {t * x * z - s * y, t * y * z + s * x, t * z * z + c}};
    for (int r = 0; r < 3; r++)
        for (int col = 0; col < 3; col++) out[col * 4 + r] = (float)R[r][col];
}
void matLookAt(const float* origin, const float* target, const float* up, float* out) {
    // Mitsuba's camera-to-world: +z looks at the target, x = left
    double d[3] = {(double)target[0] - origin[0], (double)target[1] - origin[1], (double)target[2] - origin[2]};
    double dl = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    for (double& v : d) v /= dl;
    double l[3] = {up[1] * d[2] - up[2] * d[1], up[2] * d[0] - up[0] * d[2], up[0] * d[1] - up[1] * d[0]};
    double ll = std::sqrt(l[0] * l[0] + l[1] * l[1] + l[2] * l[2]);
    for (double& v : l) v /= ll;
    double u[3] = {d[1] * l[2] - d[2] * l[1], d[2] * l[0] - d[0] * l[2], d[0] * l[1] - d[1] * l[0]};
    matIdentity(out);
    for (int r = 0; r < 3; r++) {
        out[0 * 4 + r] = (float)l[r];
        out[1 * 4 + r] = (float)u[r];
        out[2 * 4 + r] = (float)d[r];
        out[3 * 4 + r] = origin[r];
    }
}
void matReferenceCamera(const float* origin, const float* target, const float* up, float* out) {
    // The reference maps a camera-space direction d to flipY(toWorld * (-x, y, z)) in WORLD space
    // (raygen.rgen:20-35).  For a camera that is not upright this is not a rigid look-at, so the
    // matrix that makes the reference's formula look along (target - origin) is F * T * D with
    // T the true look-at frame, D = diag(1,-1,1) on camera axes and F = diag(1,-1,1) on world axes.
    matLookAt(origin, target, up, out);
    for (int r = 0; r < 3; r++) out[1 * 4 + r] = -out[1 * 4 + r];  // D: negate the up column
    for (int c = 0; c < 3; c++) out[c * 4 + 1] = -out[c * 4 + 1];  // F: negate world y of the rotation
}
void matFromRowMajor(const float* rm, float* out) {
    float r[16];
    for (int row = 0; row < 4; row++)
        for (int c = 0; c < 4; c++) r[c * 4 + row] = rm[row * 4 + c];
    std::memcpy(out, r, sizeof(r));
}

// ---- Scene ---------------------------------------------------------------------------------
void Scene::addRenderObject(const RenderObject& obj) {
    renderObjects.push_back(obj);
    const Material& mat = materials.at(obj.material);
    const bool emitting = mat.emission[0] != 0.0f || mat.emission[1] != 0.0f || mat.emission[2] != 0.0f;
    if (!emitting) return;
    const MeshData& m = meshes.at(obj.mesh);
    const float* T = obj.transform;
    for (uint32_t v = 0; v + 2 < m.vertexCount(); v += 3) {
        gsb_triangle_light light;
        std::memset(&light, 0, sizeof(light));
        for (int k = 0; k < 3; k++) {
            const float* p = &m.positions[3 * (size_t)(v + k)];
            // glm mat4 * vec4(p, 1): column sums in float, as the reference's host code does
            for (int r = 0; r < 4; r++)
                light.positions[k][r] = T[0 * 4 + r] * p[0] + T[1 * 4 + r] * p[1] + T[2 * 4 + r] * p[2] + T[3 * 4 + r];
        }
        light.radiance[0] = mat.emission[0];
        light.radiance[1] = mat.emission[1];
        light.radiance[2] = mat.emission[2];
        light.radiance[3] = 1.0f;
        triangleLights.push_back(light);
    }
}

std::vector<gsb_instance> Scene::instanceRecords() const {
    std::vector<gsb_instance> out(renderObjects.size());
    for (size_t i = 0; i < renderObjects.size(); i++) {
        const RenderObject& o = renderObjects[i];
        const Material& m = materials.at(o.material);
        gsb_instance& r = out[i];
        std::memcpy(r.transform, o.transform, sizeof(r.transform));
        r.mesh = o.mesh;
        r.emission[0] = m.emission[0]; r.emission[1] = m.emission[1]; r.emission[2] = m.emission[2];
        r.bsdf = m.bsdf;
        r.twofaced = m.twofaced ? 1u : 0u;
    }
    return out;
}

const void* Scene::bsdfData(uint32_t type, uint32_t* count, uint32_t* stride) const {
#define GSBH_CASE(T, vec) case T: *count = (uint32_t)vec.size(); *stride = (uint32_t)sizeof(vec[0]); return vec.empty() ? nullptr : (const void*)vec.data();
    switch (type) {
        GSBH_CASE(GSB_BSDF_DIFFUSE, diffuseBSDFs)
        GSBH_CASE(GSB_BSDF_SMOOTH_DIELECTRIC, smoothDielectricBSDFs)
        GSBH_CASE(GSB_BSDF_SMOOTH_CONDUCTOR, smoothConductorBSDFs)
        GSBH_CASE(GSB_BSDF_SMOOTH_PLASTIC, smoothPlasticBSDFs)
        GSBH_CASE(GSB_BSDF_ROUGH_CONDUCTOR, roughConductorBSDFs)
        GSBH_CASE(GSB_BSDF_SMOOTH_FLOOR, smoothFloorBSDFs)
        GSBH_CASE(GSB_BSDF_ROUGH_FLOOR, roughFloorBSDFs)
        GSBH_CASE(GSB_BSDF_ROUGH_PLASTIC, roughPlasticBSDFs)
        GSBH_CASE(GSB_BSDF_ROUGH_DIELECTRIC, roughDielectricBSDFs)
    }
#undef GSBH_CASE
    *count = 0;
    *stride = 0;
    return nullptr;
}

uint64_t Scene::instancedTriangles() const {
    uint64_t n = 0;
    for (auto& o : renderObjects) n += meshes.at(o.mesh).triangleCount();
    return n;
}

// ---- built-in meshes -----------------------------------------------------------------------
static void pushVertex(MeshData& m, const float* p, const float* n, float u = 0.0f, float v = 0.0f) {
    m.positions.insert(m.positions.end(), p, p + 3);
    m.normals.insert(m.normals.end(), n, n + 3);
    m.uvs.push_back(u);
    m.uvs.push_back(v);
}

MeshData makeRectangle() {
    // corners 1..4 = (-1,1) (1,1) (-1,-1) (1,-1) at z = 0; faces (1,3,2) (3,4,2); normal +z
    MeshData m;
    m.name = "builtin:rectangle";
    const float v[4][3] = {{-1, 1, 0}, {1, 1, 0}, {-1, -1, 0}, {1, -1, 0}};
    const float n[3] = {0, 0, 1};
    const int f[2][3] = {{0, 2, 1}, {2, 3, 1}};
    const float uv[4][2] = {{0, 1}, {1, 1}, {0, 0}, {1, 0}};   // the vt lines of assets/rect.obj
    for (auto& tri : f)
        for (int k : tri) pushVertex(m, v[k], n, uv[k][0], uv[k][1]);
    return m;
}

MeshData makeCube() {
    // six quads of [-1,1]^3 in the order +x, -x, +y, -y, +z, -z; each quad (a,b,c,d) gives
    // faces (a,c,b) (c,d,b), the winding of assets/box.obj
    MeshData m;
    m.name = "builtin:cube";
    const float q[6][4][3] = {
        {{1, 1, 1}, {1, 1, -1}, {1, -1, 1}, {1, -1, -1}},
        {{-1, 1, -1}, {-1, 1, 1}, {-1, -1, -1}, {-1, -1, 1}},
        {{-1, 1, -1}, {1, 1, -1}, {-1, 1, 1}, {1, 1, 1}},
        {{-1, -1, 1}, {1, -1, 1}, {-1, -1, -1}, {1, -1, -1}},
        {{-1, 1, 1}, {1, 1, 1}, {-1, -1, 1}, {1, -1, 1}},
        {{1, 1, -1}, {-1, 1, -1}, {1, -1, -1}, {-1, -1, -1}},
    };
    const float n[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    const int f[2][3] = {{0, 2, 1}, {2, 3, 1}};
    const float uv[4][2] = {{0, 1}, {1, 1}, {0, 0}, {1, 0}};   // every face carries the unit square
    for (int s = 0; s < 6; s++)
        for (auto& tri : f)
            for (int k : tri) pushVertex(m, q[s][k], n[s], uv[k][0], uv[k][1]);
    return m;
}

MeshData makeDisk(int segments) {
    // Mitsuba's `disk`: the unit disk in the xy plane, normal +z.  engine/Loader.cpp:278-279 maps the plugin to
    // "disk.obj", a file the reference never shipped; this fan of `segments` triangles (counter-clockwise seen from +z,
    // the winding of rect.obj) stands in for it.
    MeshData m;
    m.name = "builtin:disk";
    const float n[3] = {0, 0, 1}, c[3] = {0, 0, 0};
    for (int k = 0; k < segments; k++) {
        const double a0 = 2.0 * 3.14159265358979323846 * k / segments, a1 = 2.0 * 3.14159265358979323846 * (k + 1) / segments;
        const float p0[3] = {(float)std::cos(a0), (float)std::sin(a0), 0.0f};
        const float p1[3] = {k + 1 == segments ? 1.0f : (float)std::cos(a1), k + 1 == segments ? 0.0f : (float)std::sin(a1), 0.0f};
        pushVertex(m, c, n, 0.5f, 0.5f);
        pushVertex(m, p0, n, 0.5f + 0.5f * p0[0], 0.5f + 0.5f * p0[1]);
        pushVertex(m, p1, n, 0.5f + 0.5f * p1[0], 0.5f + 0.5f * p1[1]);
    }
    return m;
}

MeshData makeUvSphere(int stacks, int slices) {
    MeshData m;
    m.name = "builtin:uvsphere:" + std::to_string(stacks) + "x" + std::to_string(slices);
    auto point = [&](int i, int j, float* p) {
        const double th = 3.14159265358979323846 * i / stacks;
        const double ph = 2.0 * 3.14159265358979323846 * (j % slices) / slices;
        p[0] = (float)(std::sin(th) * std::cos(ph));
        p[1] = (float)std::cos(th);
        p[2] = (float)(std::sin(th) * std::sin(ph));
        if (i == 0) { p[0] = 0; p[1] = 1; p[2] = 0; }
        if (i == stacks) { p[0] = 0; p[1] = -1; p[2] = 0; }
    };
    for (int i = 0; i < stacks; i++)
        for (int j = 0; j < slices; j++) {
            float a[3], b[3], c[3], d[3];
            point(i, j, a); point(i + 1, j, b); point(i + 1, j + 1, c); point(i, j + 1, d);
            if (i != 0) { pushVertex(m, a, a); pushVertex(m, d, d); pushVertex(m, b, b); }
            if (i != stacks - 1) { pushVertex(m, d, d); pushVertex(m, c, c); pushVertex(m, b, b); }
        }
    return m;
}

MeshData makeDisplacedSphere(int subdiv, uint32_t seed, float amplitude) {
    // indexed icosphere -> radial displacement by a seeded sum of plane waves -> smooth normals -> soup
    std::vector<double> V;  // xyz per vertex
    std::vector<uint32_t> F;
    const double t = (1.0 + std::sqrt(5.0)) / 2.0;
    const double iv[12][3] = {{-1, t, 0}, {1, t, 0}, {-1, -t, 0}, {1, -t, 0}, {0, -1, t}, {0, 1, t},
                              {0, -1, -t}, {0, 1, -t}, {t, 0, -1}, {t, 0, 1}, {-t, 0, -1}, {-t, 0, 1}};
    for (auto& p : iv) {
        double l = std::sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
        V.push_back(p[0] / l); V.push_back(p[1] / l); V.push_back(p[2] / l);
    }
    const uint32_t itri[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11}, {1, 5, 9}, {5, 11, 4},
                                  {11, 10, 2}, {10, 7, 6}, {7, 1, 8}, {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8},
                                  {3, 8, 9}, {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
    for (auto& f : itri) { F.push_back(f[0]); F.push_back(f[1]); F.push_back(f[2]); }
    for (int s = 0; s < subdiv; s++) {
        std::map<uint64_t, uint32_t> mid;
        auto midpoint = [&](uint32_t a, uint32_t b) {
            const uint64_t key = a < b ? ((uint64_t)a << 32) | b : ((uint64_t)b << 32) | a;
            auto it = mid.find(key);
            if (it != mid.end()) return it->second;
            double p[3] = {V[3 * a] + V[3 * b], V[3 * a + 1] + V[3 * b + 1], V[3 * a + 2] + V[3 * b + 2]};
            double l = std::sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
            const uint32_t id = (uint32_t)(V.size() / 3);
            V.push_back(p[0] / l); V.push_back(p[1] / l); V.push_back(p[2] / l);
            mid.emplace(key, id);
            return id;
        };
        std::vector<uint32_t> NF;
        NF.reserve(F.size() * 4);
        for (size_t f = 0; f < F.size(); f += 3) {
            uint32_t a = F[f], b = F[f + 1], c = F[f + 2];
            uint32_t ab = midpoint(a, b), bc = midpoint(b, c), ca = midpoint(c, a);
            const uint32_t nf[12] = {a, ab, ca, b, bc, ab, c, ca, bc, ab, bc, ca};
            NF.insert(NF.end(), nf, nf + 12);
        }
        F.swap(NF);
    }
    // seeded plane waves
    uint32_t st = seed * 747796405u + 2891336453u;
    auto rnd = [&]() {
        st = st * 747796405u + 2891336453u;
        uint32_t w = ((st >> ((st >> 28u) + 4u)) ^ st) * 277803737u;
        return (double)((w >> 22u) ^ w) / 4294967296.0;
    };
    const int waves = 6;
    double wd[waves][3], wf[waves], wp[waves];
    for (int k = 0; k < waves; k++) {
        double z = 2.0 * rnd() - 1.0, ph = 6.283185307179586 * rnd(), r = std::sqrt(1.0 - z * z);
        wd[k][0] = r * std::cos(ph); wd[k][1] = r * std::sin(ph); wd[k][2] = z;
        wf[k] = 2.0 + 6.0 * rnd();
        wp[k] = 6.283185307179586 * rnd();
    }
    const size_t nv = V.size() / 3;
    for (size_t v = 0; v < nv; v++) {
        double h = 0.0;
        for (int k = 0; k < waves; k++)
            h += std::sin(wf[k] * (wd[k][0] * V[3 * v] + wd[k][1] * V[3 * v + 1] + wd[k][2] * V[3 * v + 2]) + wp[k]);
        const double rr = 1.0 + amplitude * h / waves;
        V[3 * v] *= rr; V[3 * v + 1] *= rr; V[3 * v + 2] *= rr;
    }
    std::vector<double> N(V.size(), 0.0);
    for (size_t f = 0; f < F.size(); f += 3) {
        const double* a = &V[3 * F[f]];
        const double* b = &V[3 * F[f + 1]];
        const double* c = &V[3 * F[f + 2]];
        double e1[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e2[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        double n[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
        for (int k = 0; k < 3; k++)
            for (int a3 = 0; a3 < 3; a3++) N[3 * F[f + k] + a3] += n[a3];
    }
    MeshData m;
    m.name = "builtin:displaced:" + std::to_string(subdiv) + ":" + std::to_string(seed);
    m.positions.reserve(F.size() * 3);
    m.normals.reserve(F.size() * 3);
    for (size_t f = 0; f < F.size(); f++) {
        const uint32_t v = F[f];
        double l = std::sqrt(N[3 * v] * N[3 * v] + N[3 * v + 1] * N[3 * v + 1] + N[3 * v + 2] * N[3 * v + 2]);
        if (l == 0.0) l = 1.0;
        const float p[3] = {(float)V[3 * v], (float)V[3 * v + 1], (float)V[3 * v + 2]};
        const float n[3] = {(float)(N[3 * v] / l), (float)(N[3 * v + 1] / l), (float)(N[3 * v + 2] / l)};
        pushVertex(m, p, n);
    }
    return m;
}

// ---- built-in scenes -----------------------------------------------------------------------
Scene loadMitsubaSceneFromString(const std::string& xml, const std::string& baseDir);

// The numbers of assets/scenes/cornell-box/scene.xml (Bitterli's CC0 Cornell box): 5 walls,
// 2 boxes, 1 area light; 8 instances of 2 unique meshes, 36 triangles, 2 triangle lights.
static const char* kCornellXml = R"XML(<?xml version="1.0" encoding="utf-8"?>
<scene version="0.5.0">
  <sensor type="perspective">
    <float name="fov" value="19.5"/>
    <transform name="toWorld"><matrix value="-1 0 0 0 0 1 0 1 0 0 -1 6.8 0 0 0 1"/></transform>
    <film type="ldrfilm"><integer name="width" value="1024"/><integer name="height" value="1024"/></film>
  </sensor>
  <bsdf type="twosided" id="LeftWall"><bsdf type="diffuse"><rgb name="reflectance" value="0.63, 0.065, 0.05"/></bsdf></bsdf>
  <bsdf type="twosided" id="RightWall"><bsdf type="diffuse"><rgb name="reflectance" value="0.14, 0.45, 0.091"/></bsdf></bsdf>
  <bsdf type="twosided" id="Floor"><bsdf type="diffuse"><rgb name="reflectance" value="0.725, 0.71, 0.68"/></bsdf></bsdf>
  <bsdf type="twosided" id="Ceiling"><bsdf type="diffuse"><rgb name="reflectance" value="0.725, 0.71, 0.68"/></bsdf></bsdf>
  <bsdf type="twosided" id="BackWall"><bsdf type="diffuse"><rgb name="reflectance" value="0.725, 0.71, 0.68"/></bsdf></bsdf>
  <bsdf type="twosided" id="ShortBox"><bsdf type="diffuse"><rgb name="reflectance" value="0.725, 0.71, 0.68"/></bsdf></bsdf>
  <bsdf type="twosided" id="TallBox"><bsdf type="diffuse"><rgb name="reflectance" value="0.725, 0.71, 0.68"/></bsdf></bsdf>
  <bsdf type="twosided" id="Light"><bsdf type="diffuse"><rgb name="reflectance" value="0, 0, 0"/></bsdf></bsdf>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="-4.37114e-008 1 4.37114e-008 0 0 -8.74228e-008 2 0 1 4.37114e-008 1.91069e-015 0 0 0 0 1"/></transform>
    <ref id="Floor"/>
  </shape>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="-1 7.64274e-015 -1.74846e-007 0 8.74228e-008 8.74228e-008 -2 2 0 -1 -4.37114e-008 0 0 0 0 1"/></transform>
    <ref id="Ceiling"/>
  </shape>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="1.91069e-015 1 1.31134e-007 0 1 3.82137e-015 -8.74228e-008 1 -4.37114e-008 1.31134e-007 -2 -1 0 0 0 1"/></transform>
    <ref id="BackWall"/>
  </shape>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="4.37114e-008 -1.74846e-007 2 1 1 3.82137e-015 -8.74228e-008 1 3.82137e-015 1 2.18557e-007 0 0 0 0 1"/></transform>
    <ref id="RightWall"/>
  </shape>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="-4.37114e-008 8.74228e-008 -2 -1 1 3.82137e-015 -8.74228e-008 1 0 -1 -4.37114e-008 0 0 0 0 1"/></transform>
    <ref id="LeftWall"/>
  </shape>
  <shape type="cube">
    <transform name="toWorld"><matrix value="0.0851643 0.289542 1.31134e-008 0.328631 3.72265e-009 1.26563e-008 -0.3 0.3 -0.284951 0.0865363 5.73206e-016 0.374592 0 0 0 1"/></transform>
    <ref id="ShortBox"/>
  </shape>
  <shape type="cube">
    <transform name="toWorld"><matrix value="0.286776 0.098229 -2.29282e-015 -0.335439 -4.36233e-009 1.23382e-008 -0.6 0.6 -0.0997984 0.282266 2.62268e-008 -0.291415 0 0 0 1"/></transform>
    <ref id="TallBox"/>
  </shape>
  <shape type="rectangle">
    <transform name="toWorld"><matrix value="0.235 -1.66103e-008 -7.80685e-009 -0.005 -2.05444e-008 3.90343e-009 -0.0893 1.98 2.05444e-008 0.19 8.30516e-009 -0.03 0 0 0 1"/></transform>
    <ref id="Light"/>
    <emitter type="area"><rgb name="radiance" value="17, 12, 4"/></emitter>
  </shape>
</scene>
)XML";

Scene makeCornellBox() { return loadMitsubaSceneFromString(kCornellXml, "."); }

// BSDF gallery (BASELINE.json config 3): the Cornell shell without its boxes + seven spheres, one
// per non-diffuse material the reference's loader can produce (Loader.cpp:162-227) and one frosted-glass
// sphere for the rough dielectric extension.  Material
// constants follow assets/scenes/test3/scene.xml:57-95 where it has them.
Scene makeBsdfGallery() {
    Scene base = makeCornellBox();
    Scene s;
    s.camera = base.camera;
    s.filmWidth = 1920;
    s.filmHeight = 1080;
    s.meshes = base.meshes;  // 0 = rectangle, 1 = cube
    const uint32_t sphere = s.addMesh(makeUvSphere(48, 96));
    s.diffuseBSDFs = base.diffuseBSDFs;
    // walls + light: render objects 0..4 and 7 of the Cornell box
    const int keep[6] = {0, 1, 2, 3, 4, 7};
    for (int k : keep) {
        RenderObject o = base.renderObjects[k];
        o.material = s.addMaterial(base.materials[base.renderObjects[k].material]);
        s.addRenderObject(o);
    }
    struct Ball { float x, y, z, r; uint32_t bsdf; };
    const float ior = 1.5f;
    float R0 = (ior - 1.0f) / (ior + 1.0f);
    R0 *= R0;
    const float s2 = (float)std::sqrt(2.0f);
    gsb_rough_conductor_bsdf gold = {{0.143119f, 0.374957f, 1.44248f}, {3.98316f, 2.38572f, 1.60322f}, {1, 1, 1}, s2 * 0.1f, 0};
    gsb_rough_conductor_bsdf alu = {{1.65746f, 0.880369f, 0.521229f}, {9.22387f, 6.26952f, 4.837f}, {1, 1, 1}, s2 * 0.3f, 0};
    gsb_smooth_conductor_bsdf mirror = {0.0f, 1.0f};
    gsb_smooth_dielectric_bsdf glass = {1.5f, 1.0f};
    gsb_smooth_plastic_bsdf plastic = {{1.0f, 0.378676f, 0.0134734f}, ior, 1.0f, R0};
    gsb_rough_plastic_bsdf rplastic1 = {{0.434734f, 0.578676f, 1.0f}, ior, 1.0f, R0, s2 * 0.05f, 0};
    gsb_rough_dielectric_bsdf frosted = {1.5f, 1.0f, 0.15f, 0};
    const Ball balls[7] = {
        {-0.62f, 0.27f, 0.35f, 0.27f, s.addRoughConductorBSDF(gold)},
        {0.0f, 0.27f, 0.45f, 0.27f, s.addSmoothDielectricBSDF(glass)},
        {0.62f, 0.27f, 0.35f, 0.27f, s.addSmoothPlasticBSDF(plastic)},
        {-0.62f, 0.95f, -0.45f, 0.27f, s.addSmoothConductorBSDF(mirror)},
        {0.0f, 1.05f, -0.5f, 0.27f, s.addRoughPlasticBSDF(rplastic1)},
        {0.62f, 0.95f, -0.45f, 0.27f, s.addRoughConductorBSDF(alu)},
        {0.0f, 0.3f, -0.45f, 0.3f, s.addRoughDielectricBSDF(frosted)},
    };
    for (const Ball& b : balls) {
        Material m;
        m.bsdf = b.bsdf;
        m.hasBsdf = true;
        m.twofaced = (b.bsdf >> 16) != GSB_BSDF_SMOOTH_DIELECTRIC && (b.bsdf >> 16) != GSB_BSDF_ROUGH_DIELECTRIC;
        RenderObject o;
        float T[16], S[16];
        matTranslate(b.x, b.y, b.z, T);
        matScale(b.r, b.r, b.r, S);
        matMul(T, S, o.transform);
        o.mesh = sphere;
        o.material = s.addMaterial(m);
        s.addRenderObject(o);
    }
    // pedestals under the raised back-row spheres
    const float ped[3][4] = {{-0.62f, 0.34f, -0.45f, 0.2f}, {0.62f, 0.34f, -0.45f, 0.2f}, {0.0f, 0.0f, 0.0f, 0.0f}};
    for (int k = 0; k < 2; k++) {
        Material m;
        m.bsdf = GSB_BSDF_HANDLE(GSB_BSDF_DIFFUSE, 2);
        m.hasBsdf = true;
        m.twofaced = true;
        RenderObject o;
        float T[16], S[16];
        matTranslate(ped[k][0], ped[k][1], ped[k][2], T);
        matScale(ped[k][3], ped[k][1], ped[k][3], S);
        matMul(T, S, o.transform);
        o.mesh = 1;
        o.material = s.addMaterial(m);
        s.addRenderObject(o);
    }
    return s;
}

// The Cornell box with textured reflectances (extension, see gsb_texture): the walls are instances of the rectangle mesh
// (texture coordinates = the vt lines of assets/rect.obj) and the boxes instances of the cube mesh.  Records 0 (floor),
// 2 (back wall) and 6 (tall box) of the diffuse table — the order cornell-box/scene.xml lists its shapes in — get a
// checkerboard, a procedurally painted bitmap and a finer checkerboard.
Scene makeTexturedCornellBox() {
    Scene s = makeCornellBox();
    Scene::TextureData floorTex;
    std::memset(&floorTex.desc, 0, sizeof(floorTex.desc));
    floorTex.desc.type = GSB_TEXTURE_CHECKERBOARD;
    const float a0[3] = {0.75f, 0.72f, 0.68f}, a1[3] = {0.15f, 0.2f, 0.35f};
    std::memcpy(floorTex.desc.color0, a0, sizeof(a0));
    std::memcpy(floorTex.desc.color1, a1, sizeof(a1));
    floorTex.desc.uscale = floorTex.desc.vscale = 4.0f;
    Scene::TextureData wall;
    std::memset(&wall.desc, 0, sizeof(wall.desc));
    wall.desc.type = GSB_TEXTURE_BITMAP;
    wall.desc.width = 96; wall.desc.height = 64;
    wall.desc.uscale = wall.desc.vscale = 1.0f;
    wall.texels.resize((size_t)96 * 64 * 3);
    for (int y = 0; y < 64; y++)
        for (int x = 0; x < 96; x++) {
            const float u = (x + 0.5f) / 96.0f, v = (y + 0.5f) / 64.0f;
            const float ring = 0.5f + 0.5f * std::sin(40.0f * std::sqrt((u - 0.5f) * (u - 0.5f) * 2.25f + (v - 0.5f) * (v - 0.5f)));
            float* p = &wall.texels[3 * ((size_t)y * 96 + x)];
            p[0] = 0.15f + 0.7f * u; p[1] = 0.15f + 0.6f * ring; p[2] = 0.8f - 0.6f * v;
        }
    Scene::TextureData boxTex = floorTex;
    const float b0[3] = {0.8f, 0.3f, 0.2f}, b1[3] = {0.9f, 0.85f, 0.6f};
    std::memcpy(boxTex.desc.color0, b0, sizeof(b0));
    std::memcpy(boxTex.desc.color1, b1, sizeof(b1));
    boxTex.desc.uscale = 3.0f; boxTex.desc.vscale = 6.0f; boxTex.desc.uoffset = 0.25f;
    const uint32_t t0 = s.addTexture(std::move(floorTex)), t1 = s.addTexture(std::move(wall)), t2 = s.addTexture(std::move(boxTex));
    if (s.diffuseBSDFs.size() >= 7) {
        s.diffuseBSDFs[0].hasTexture = (int32_t)t0 + 1;
        s.diffuseBSDFs[2].hasTexture = (int32_t)t1 + 1;
        s.diffuseBSDFs[6].hasTexture = (int32_t)t2 + 1;
    }
    return s;
}

// Instanced high-geometry field (BASELINE.json config 4): `uniqueMeshes` displaced icospheres of
// 20*4^subdiv triangles each, instanced gridN x gridN times with seeded jitter/rotation/scale over
// a floor, lit by one quad light.  Defaults 16 / 7 / 13 give 169 instances x 327 680 = 55.4 M
// instanced triangles from 5.2 M unique ones.
Scene makeInstancedField(uint32_t seed, int uniqueMeshes, int subdiv, int gridN) {
    Scene s;
    s.filmWidth = 1920;
    s.filmHeight = 1080;
    uint32_t st = seed * 747796405u + 2891336453u;
    auto rnd = [&]() {
        st = st * 747796405u + 2891336453u;
        uint32_t w = ((st >> ((st >> 28u) + 4u)) ^ st) * 277803737u;
        return (float)((double)((w >> 22u) ^ w) / 4294967296.0);
    };
    const uint32_t rect = s.addMesh(makeRectangle());
    std::vector<uint32_t> blobs;
    for (int k = 0; k < uniqueMeshes; k++) blobs.push_back(s.addMesh(makeDisplacedSphere(subdiv, seed + 17u * (uint32_t)k, 0.25f)));
    const float extent = (float)gridN * 1.25f;
    // floor
    {
        Material m;
        m.bsdf = s.addDiffuseBSDF({{0.6f, 0.6f, 0.6f}, 0});
        m.hasBsdf = true;
        m.twofaced = true;
        RenderObject o;
        float S[16], R[16];
        matScale(extent * 2.0f, extent * 2.0f, 1.0f, S);
        matRotate(1, 0, 0, -90.0f, R);
        matMul(R, S, o.transform);
        o.mesh = rect;
        o.material = s.addMaterial(m);
        s.addRenderObject(o);
    }
    // light: quad facing down
    {
        Material m;
        m.bsdf = s.addDiffuseBSDF({{0.0f, 0.0f, 0.0f}, 0});
        m.hasBsdf = true;
        m.twofaced = true;
        m.emission[0] = 18.0f; m.emission[1] = 17.0f; m.emission[2] = 15.0f;
        RenderObject o;
        float S[16], R[16], T[16], RS[16];
        matScale(extent * 0.5f, extent * 0.5f, 1.0f, S);
        matRotate(1, 0, 0, 90.0f, R);
        matTranslate(0.0f, extent * 0.9f, 0.0f, T);
        matMul(R, S, RS);
        matMul(T, RS, o.transform);
        o.mesh = rect;
        o.material = s.addMaterial(m);
        s.addRenderObject(o);
    }
    // a handful of materials shared by the instances
    std::vector<uint32_t> mats;
    for (int k = 0; k < 6; k++) mats.push_back(s.addDiffuseBSDF({{0.25f + 0.6f * rnd(), 0.25f + 0.6f * rnd(), 0.25f + 0.6f * rnd()}, 0}));
    mats.push_back(s.addRoughConductorBSDF({{1.65746f, 0.880369f, 0.521229f}, {9.22387f, 6.26952f, 4.837f}, {1, 1, 1}, 0.2f, 0}));
    mats.push_back(s.addRoughPlasticBSDF({{0.8f, 0.3f, 0.2f}, 1.5f, 1.0f, 0.04f, 0.15f, 0}));
    for (int gz = 0; gz < gridN; gz++)
        for (int gx = 0; gx < gridN; gx++) {
            Material m;
            m.bsdf = mats[(size_t)(rnd() * (float)mats.size()) % mats.size()];
            m.hasBsdf = true;
            m.twofaced = true;
            RenderObject o;
            const float sc = 0.7f + 0.4f * rnd();
            const float px = ((float)gx - 0.5f * (float)(gridN - 1)) * 2.5f + (rnd() - 0.5f) * 0.6f;
            const float pz = ((float)gz - 0.5f * (float)(gridN - 1)) * 2.5f + (rnd() - 0.5f) * 0.6f;
            float S[16], R[16], T[16], RS[16];
            matScale(sc, sc, sc, S);
            matRotate(rnd() - 0.5f, rnd() + 0.1f, rnd() - 0.5f, 360.0f * rnd(), R);
            matTranslate(px, sc * 1.3f, pz, T);
            matMul(R, S, RS);
            matMul(T, RS, o.transform);
            o.mesh = blobs[(size_t)(rnd() * (float)blobs.size()) % blobs.size()];
            o.material = s.addMaterial(m);
            s.addRenderObject(o);
        }
    const float origin[3] = {0.0f, extent * 0.55f, extent * 1.35f};
    const float target[3] = {0.0f, 0.5f, 0.0f};
    const float up[3] = {0.0f, 1.0f, 0.0f};
    matReferenceCamera(origin, target, up, s.camera.toWorld);
    s.camera.fov = 40.0f * 3.14159265358979323846f / 180.f;
    return s;
}

// The instanced field under a procedural sky (environment emitter extension): a 128 x 64 latitude-longitude image — blue
// gradient above the horizon, dim ground colour below, and a small very bright sun 35 degrees up — rotated a little about the
// vertical axis.  The quad light of the field stays (dimmed), so light sampling chooses between triangle lights and the sky.
Scene makeFieldUnderSky(uint32_t seed, int uniqueMeshes, int subdiv, int gridN) {
    Scene s = makeInstancedField(seed, uniqueMeshes, subdiv, gridN);
    const uint32_t W = 128, H = 64;
    s.environmentTexels.resize((size_t)W * H * 3);
    const float kPi = 3.14159265358979323846f;
    const float sunTheta = (90.0f - 35.0f) * kPi / 180.0f, sunPhi = 0.6f;
    for (uint32_t y = 0; y < H; y++)
        for (uint32_t x = 0; x < W; x++) {
            const float th = (y + 0.5f) / H * kPi, ph = ((x + 0.5f) / W - 0.5f) * 2.0f * kPi;
            float* p = &s.environmentTexels[3 * ((size_t)y * W + x)];
            if (th < 0.5f * kPi) {
                const float t = th / (0.5f * kPi);
                p[0] = 0.25f + 0.45f * t; p[1] = 0.45f + 0.35f * t; p[2] = 0.9f;
            } else {
                p[0] = 0.12f; p[1] = 0.1f; p[2] = 0.08f;
            }
            const float dth = th - sunTheta, dph = std::remainder(ph - sunPhi, 2.0f * kPi) * std::sin(sunTheta);
            if (dth * dth + dph * dph < 0.05f * 0.05f) { p[0] = 900.0f; p[1] = 800.0f; p[2] = 600.0f; }
        }
    std::memset(&s.environment, 0, sizeof(s.environment));
    s.environment.type = GSB_ENVIRONMENT_LATLONG;
    s.environment.width = W;
    s.environment.height = H;
    s.environment.scale[0] = s.environment.scale[1] = s.environment.scale[2] = 1.0f;
    const float a = 0.4f, ca = std::cos(a), sa = std::sin(a);
    const float Rm[9] = {ca, 0, sa, 0, 1, 0, -sa, 0, ca};
    std::memcpy(s.environment.to_world, Rm, sizeof(Rm));
    return s;
}

}  // namespace gsbhost
