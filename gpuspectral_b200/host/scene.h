// scene.h — host-side scene model of the B200 path-tracing core.
//
// Mirrors what the reference keeps in renderer/Scene.h:16-186 (Scene, RenderObject, Material,
// BSDFHandle, the eight BSDF vectors, TriangleLight) and renderer/Camera.h, minus every
// Vulkan type: a mesh is plain float arrays (the de-interleaved position / normal streams of
// renderer/Mesh.cpp:5-22) and a BSDF handle is the packed uint32 the shaders see.
#pragma once

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/gpuspectral_b200.h"

namespace gsbhost {

// Non-indexed triangle soup: vertex i of triangle p is element 3p+i (Loader.cpp:42-63 emits
// indices 0..n-1, VulkanRays.cpp:17-35 consumes float3 stride 12).
struct MeshData {
    std::string name;
    std::vector<float> positions;  // 3 floats per vertex
    std::vector<float> normals;    // 3 floats per vertex
    std::vector<float> uvs;        // 2 floats per vertex, or empty (Mesh::Vertex::uv, renderer/Mesh.h:17)
    uint32_t vertexCount() const { return (uint32_t)(positions.size() / 3); }
    uint32_t triangleCount() const { return vertexCount() / 3; }
};

struct Material {  // Scene.h:99-104
    float emission[3] = {0.0f, 0.0f, 0.0f};
    bool twofaced = false;
    bool facenormals = false;
    uint32_t bsdf = GSB_BSDF_HANDLE(GSB_BSDF_DIFFUSE, 0);
    bool hasBsdf = false;
};

struct RenderObject {  // Scene.h:23-27
    float transform[16];  // column-major, object -> world
    uint32_t mesh = 0;
    int material = 0;
};

struct Camera {  // Camera.h: toWorld + fov; eye = toWorld[3] (Camera.cpp:33-35)
    float toWorld[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float fov = 0.6f;  // radians
};

struct Scene {
    uint64_t revision = 0;   // bump (touch()) after editing transforms / materials / lights in place: PathTracer re-uploads
    void touch() { revision++; }
    Camera camera;
    uint32_t filmWidth = 0, filmHeight = 0;  // from the XML film; the reference ignores it (frame size = window size)
    std::vector<MeshData> meshes;
    std::vector<RenderObject> renderObjects;
    std::vector<Material> materials;
    std::vector<gsb_triangle_light> triangleLights;
    std::vector<gsb_diffuse_bsdf> diffuseBSDFs;
    std::vector<gsb_smooth_dielectric_bsdf> smoothDielectricBSDFs;
    std::vector<gsb_smooth_conductor_bsdf> smoothConductorBSDFs;
    std::vector<gsb_smooth_plastic_bsdf> smoothPlasticBSDFs;
    std::vector<gsb_rough_conductor_bsdf> roughConductorBSDFs;
    std::vector<gsb_smooth_floor_bsdf> smoothFloorBSDFs;
    std::vector<gsb_rough_floor_bsdf> roughFloorBSDFs;
    std::vector<gsb_rough_plastic_bsdf> roughPlasticBSDFs;
    std::vector<gsb_rough_dielectric_bsdf> roughDielectricBSDFs;   // extension (no reference counterpart)
    std::vector<std::string> warnings;
    // reflectance textures (extension): descriptors as the C ABI takes them, texels owned here
    struct TextureData { gsb_texture desc; std::vector<float> texels; };
    std::vector<TextureData> textures;
    // environment emitter (extension): descriptor as the C ABI takes it, texels owned here; type 0 = none
    gsb_environment environment = {};
    std::vector<float> environmentTexels;
    gsb_environment environmentRecord() const { gsb_environment e = environment; e.texels = environmentTexels.empty() ? nullptr : environmentTexels.data(); return e; }
    uint32_t addTexture(TextureData&& t) { textures.push_back(std::move(t)); return (uint32_t)textures.size() - 1; }
    std::vector<gsb_texture> textureRecords() const {
        std::vector<gsb_texture> out;
        for (const TextureData& t : textures) { gsb_texture d = t.desc; d.texels = t.texels.empty() ? nullptr : t.texels.data(); out.push_back(d); }
        return out;
    }

    int addMaterial(const Material& m) { materials.push_back(m); return (int)materials.size() - 1; }
    uint32_t addMesh(MeshData&& m) { meshes.push_back(std::move(m)); return (uint32_t)meshes.size() - 1; }

    uint32_t addDiffuseBSDF(const gsb_diffuse_bsdf& b) { diffuseBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_DIFFUSE, diffuseBSDFs.size() - 1); }
    uint32_t addSmoothDielectricBSDF(const gsb_smooth_dielectric_bsdf& b) { smoothDielectricBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_SMOOTH_DIELECTRIC, smoothDielectricBSDFs.size() - 1); }
    uint32_t addSmoothConductorBSDF(const gsb_smooth_conductor_bsdf& b) { smoothConductorBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_SMOOTH_CONDUCTOR, smoothConductorBSDFs.size() - 1); }
    uint32_t addSmoothPlasticBSDF(const gsb_smooth_plastic_bsdf& b) { smoothPlasticBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_SMOOTH_PLASTIC, smoothPlasticBSDFs.size() - 1); }
    uint32_t addRoughConductorBSDF(const gsb_rough_conductor_bsdf& b) { roughConductorBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_ROUGH_CONDUCTOR, roughConductorBSDFs.size() - 1); }
    uint32_t addSmoothFloorBSDF(const gsb_smooth_floor_bsdf& b) { smoothFloorBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_SMOOTH_FLOOR, smoothFloorBSDFs.size() - 1); }
    uint32_t addRoughFloorBSDF(const gsb_rough_floor_bsdf& b) { roughFloorBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_ROUGH_FLOOR, roughFloorBSDFs.size() - 1); }
    uint32_t addRoughPlasticBSDF(const gsb_rough_plastic_bsdf& b) { roughPlasticBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_ROUGH_PLASTIC, roughPlasticBSDFs.size() - 1); }

    uint32_t addRoughDielectricBSDF(const gsb_rough_dielectric_bsdf& b) { roughDielectricBSDFs.push_back(b); return GSB_BSDF_HANDLE(GSB_BSDF_ROUGH_DIELECTRIC, roughDielectricBSDFs.size() - 1); }

    // Adds the object and, when its material emits, one world-space TriangleLight per mesh
    // triangle (Loader.cpp:310-330).
    void addRenderObject(const RenderObject& obj);

    // The instance records the C ABI consumes (PathTracer.cpp:58-70 without the device addresses).
    std::vector<gsb_instance> instanceRecords() const;

    const void* bsdfData(uint32_t type, uint32_t* count, uint32_t* stride) const;
    uint64_t instancedTriangles() const;
};

// ---- loaders / generators ---------------------------------------------------------------
// Mitsuba 0.5 / 2 XML (+ OBJ shapes) following the conventions of engine/Loader.cpp:145-349.
// Throws std::runtime_error on unreadable input.
Scene loadMitsubaScene(const std::string& xmlPath);
MeshData loadObjMesh(const std::string& objPath);
// Image files for bitmap textures: PFM, binary PPM (P6) and PNG (8-bit, non-interlaced; own inflate).  Returns linear RGB
// (8-bit sources go through the sRGB decoding curve), rows top to bottom.  Throws std::runtime_error.
void loadImageLinearRGB(const std::string& path, std::vector<float>& rgb, uint32_t& width, uint32_t& height);
// glTF 2.0 (.gltf / .glb), see gltf.cpp.  Throws std::runtime_error on unreadable input.
Scene loadGltfScene(const std::string& path);

Scene makeFieldUnderSky(uint32_t seed, int uniqueMeshes, int subdiv, int gridN);   // the instanced field lit by a latitude-longitude sky with a sun (environment extension)
Scene makeTexturedCornellBox();                 // the Cornell box with a checkerboard floor, a bitmap back wall and a checkerboard tall box (texture extension)
MeshData makeRectangle();                       // assets/rect.obj equivalent (2 triangles, z = 0, normal +z)
MeshData makeCube();
MeshData makeDisk(int segments = 64);                            // assets/box.obj equivalent (12 triangles, [-1,1]^3)
MeshData makeUvSphere(int stacks, int slices);  // unit sphere, smooth normals
MeshData makeDisplacedSphere(int subdiv, uint32_t seed, float amplitude);  // icosphere, 20*4^subdiv triangles

// Built-in scenes (no files needed): the Cornell box of assets/scenes/cornell-box/scene.xml,
// a BSDF gallery (BASELINE.json config 3) and a seeded instanced high-geometry scene (config 4).
Scene makeCornellBox();
Scene makeBsdfGallery();
Scene makeInstancedField(uint32_t seed, int uniqueMeshes, int subdiv, int gridN);

// ---- small matrix helpers (column-major float[16]) ---------------------------------------
void matIdentity(float* m);
void matMul(const float* a, const float* b, float* out);  // out = a * b
void matTranslate(float x, float y, float z, float* out);
void matScale(float x, float y, float z, float* out);
void matRotate(float ax, float ay, float az, float degrees, float* out);
void matLookAt(const float* origin, const float* target, const float* up, float* out);
void matReferenceCamera(const float* origin, const float* target, const float* up, float* out);
void matFromRowMajor(const float* rowMajor16, float* out);

}  // namespace gsbhost
