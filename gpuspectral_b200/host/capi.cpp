// capi.cpp — extern "C" wrappers of the host layer (include/gpuspectral_b200_host.h).
#include <cstring>
#include <exception>
#include <vector>
#include <stdexcept>
#include <string>

#include "../../include/gpuspectral_b200_host.h"
#include "path_tracer.h"
#include "scene.h"

namespace gsbhost {
Scene loadMitsubaSceneFromString(const std::string& xml, const std::string& baseDir);
}

struct gsbh_scene { gsbhost::Scene scene; };
struct gsbh_tracer { gsbhost::PathTracer tracer; explicit gsbh_tracer(int dev) : tracer(dev) {} };

static thread_local std::string g_hostError;

template <class F>
static int guarded(F&& f) {
    try {
        f();
        return 0;
    } catch (const std::exception& e) {
        g_hostError = e.what();
        return 1;
    }
}

extern "C" {

const char* gsbh_last_error(void) { return g_hostError.c_str(); }

gsbh_scene* gsbh_scene_load_xml(const char* path) {
    gsbh_scene* s = nullptr;
    guarded([&] { s = new gsbh_scene{gsbhost::loadMitsubaScene(path)}; });
    return s;
}

gsbh_scene* gsbh_scene_load_gltf(const char* path) {
    gsbh_scene* s = nullptr;
    guarded([&] { s = new gsbh_scene{gsbhost::loadGltfScene(path)}; });
    return s;
}

gsbh_scene* gsbh_scene_load_xml_string(const char* xml, const char* base_dir) {
    gsbh_scene* s = nullptr;
    guarded([&] { s = new gsbh_scene{gsbhost::loadMitsubaSceneFromString(xml, base_dir ? base_dir : ".")}; });
    return s;
}

gsbh_scene* gsbh_scene_builtin(const char* name, const int* params, int nparams) {
    gsbh_scene* s = nullptr;
    guarded([&] {
        const std::string n = name ? name : "";
        auto par = [&](int i, int def) { return (params && i < nparams) ? params[i] : def; };
        if (n == "cornell") s = new gsbh_scene{gsbhost::makeCornellBox()};
        else if (n == "field_env") s = new gsbh_scene{gsbhost::makeFieldUnderSky((uint32_t)par(0, 1234), par(1, 3), par(2, 3), par(3, 3))};
        else if (n == "cornell_textured") s = new gsbh_scene{gsbhost::makeTexturedCornellBox()};
        else if (n == "gallery") s = new gsbh_scene{gsbhost::makeBsdfGallery()};
        else if (n == "field") s = new gsbh_scene{gsbhost::makeInstancedField((uint32_t)par(0, 1234), par(1, 16), par(2, 7), par(3, 13))};
        else throw std::runtime_error("unknown built-in scene '" + n + "'");
    });
    return s;
}

void gsbh_scene_free(gsbh_scene* s) { delete s; }

uint32_t gsbh_scene_mesh_count(const gsbh_scene* s) { return (uint32_t)s->scene.meshes.size(); }
uint32_t gsbh_scene_mesh_vertex_count(const gsbh_scene* s, uint32_t m) { return s->scene.meshes.at(m).vertexCount(); }
const float* gsbh_scene_mesh_positions(const gsbh_scene* s, uint32_t m) { return s->scene.meshes.at(m).positions.data(); }
const float* gsbh_scene_mesh_normals(const gsbh_scene* s, uint32_t m) { return s->scene.meshes.at(m).normals.data(); }
const float* gsbh_scene_mesh_uvs(const gsbh_scene* s, uint32_t m) {
    const auto& mesh = s->scene.meshes.at(m);
    return mesh.uvs.size() == (size_t)mesh.vertexCount() * 2 ? mesh.uvs.data() : nullptr;
}
void gsbh_scene_environment(const gsbh_scene* s, gsb_environment* out) { *out = s->scene.environmentRecord(); }
uint32_t gsbh_scene_texture_count(const gsbh_scene* s) { return (uint32_t)s->scene.textures.size(); }
void gsbh_scene_copy_textures(const gsbh_scene* s, gsb_texture* out) {
    const auto v = s->scene.textureRecords();   // texel pointers stay valid as long as the scene lives
    if (!v.empty()) std::memcpy(out, v.data(), v.size() * sizeof(gsb_texture));
}
const char* gsbh_scene_mesh_name(const gsbh_scene* s, uint32_t m) { return s->scene.meshes.at(m).name.c_str(); }
uint32_t gsbh_scene_instance_count(const gsbh_scene* s) { return (uint32_t)s->scene.renderObjects.size(); }
void gsbh_scene_copy_instances(const gsbh_scene* s, gsb_instance* out) {
    auto v = s->scene.instanceRecords();
    if (!v.empty()) std::memcpy(out, v.data(), v.size() * sizeof(gsb_instance));
}
uint32_t gsbh_scene_bsdf_count(const gsbh_scene* s, uint32_t type) {
    uint32_t c = 0, st = 0;
    s->scene.bsdfData(type, &c, &st);
    return c;
}
const void* gsbh_scene_bsdf_data(const gsbh_scene* s, uint32_t type) {
    uint32_t c = 0, st = 0;
    return s->scene.bsdfData(type, &c, &st);
}
uint32_t gsbh_scene_bsdf_stride(uint32_t type) {
    static const uint32_t strides[GSB_BSDF_TYPE_COUNT] = {
        sizeof(gsb_diffuse_bsdf), sizeof(gsb_smooth_dielectric_bsdf), sizeof(gsb_smooth_conductor_bsdf), sizeof(gsb_smooth_plastic_bsdf),
        sizeof(gsb_rough_conductor_bsdf), sizeof(gsb_smooth_floor_bsdf), sizeof(gsb_rough_floor_bsdf), sizeof(gsb_rough_plastic_bsdf),
        sizeof(gsb_rough_dielectric_bsdf)};
    return type < GSB_BSDF_TYPE_COUNT ? strides[type] : 0;
}
uint32_t gsbh_scene_light_count(const gsbh_scene* s) { return (uint32_t)s->scene.triangleLights.size(); }
const gsb_triangle_light* gsbh_scene_lights(const gsbh_scene* s) { return s->scene.triangleLights.data(); }
void gsbh_scene_camera(const gsbh_scene* s, float eye[3], float to_world[16], float* fov) {
    const float* T = s->scene.camera.toWorld;
    std::memcpy(to_world, T, sizeof(float) * 16);
    eye[0] = T[12]; eye[1] = T[13]; eye[2] = T[14];
    *fov = s->scene.camera.fov;
}
void gsbh_scene_film(const gsbh_scene* s, uint32_t* w, uint32_t* h) { *w = s->scene.filmWidth; *h = s->scene.filmHeight; }
uint64_t gsbh_scene_instanced_triangles(const gsbh_scene* s) { return s->scene.instancedTriangles(); }
uint32_t gsbh_scene_warning_count(const gsbh_scene* s) { return (uint32_t)s->scene.warnings.size(); }
const char* gsbh_scene_warning(const gsbh_scene* s, uint32_t i) { return s->scene.warnings.at(i).c_str(); }

gsbh_tracer* gsbh_tracer_create(int device_ordinal) {
    gsbh_tracer* t = nullptr;
    guarded([&] { t = new gsbh_tracer(device_ordinal); });
    return t;
}
void gsbh_tracer_destroy(gsbh_tracer* t) { delete t; }
gsb_context* gsbh_tracer_context(gsbh_tracer* t) { return t->tracer.context(); }
int gsbh_tracer_setup(gsbh_tracer* t, uint32_t w, uint32_t h) { return guarded([&] { t->tracer.setup(w, h); }); }
int gsbh_tracer_set_params(gsbh_tracer* t, const gsb_params* p) { return guarded([&] { t->tracer.setParams(*p); }); }
int gsbh_tracer_set_timestamp(gsbh_tracer* t, int ts) { t->tracer.timestamp = ts; return 0; }
int gsbh_tracer_upload(gsbh_tracer* t, const gsbh_scene* s) { return guarded([&] { t->tracer.uploadScene(s->scene); }); }
int gsbh_tracer_create_render_pass(gsbh_tracer* t, const gsbh_scene* s) { return guarded([&] { t->tracer.createRenderPass(s->scene); }); }
int gsbh_tracer_render(gsbh_tracer* t, const gsbh_scene* s, uint32_t spp, uint32_t stride) {
    return guarded([&] { t->tracer.render(s->scene, spp, stride); });
}
int gsbh_tracer_read_hdr(gsbh_tracer* t, float* out) {
    return guarded([&] {
        if (gsb_read_hdr(t->tracer.context(), out) != GSB_OK) throw std::runtime_error(gsb_last_error());
    });
}
int gsbh_tracer_read_ldr(gsbh_tracer* t, uint8_t* out) {
    return guarded([&] {
        if (gsb_read_ldr(t->tracer.context(), out) != GSB_OK) throw std::runtime_error(gsb_last_error());
    });
}
int gsbh_tracer_synchronize(gsbh_tracer* t) { return guarded([&] { t->tracer.synchronize(); }); }

int gsbh_write_pfm(const char* path, const float* rgba, uint32_t w, uint32_t h) { return guarded([&] { gsbhost::writePFM(path, rgba, w, h); }); }
int gsbh_read_pfm(const char* path, float* rgba, uint32_t capacity_pixels, uint32_t* w, uint32_t* h) {
    return guarded([&] {
        std::vector<float> img;
        uint32_t ww = 0, hh = 0;
        gsbhost::readPFM(path, img, ww, hh);
        if (w) *w = ww;
        if (h) *h = hh;
        if (rgba) {
            if ((uint64_t)ww * hh > capacity_pixels) throw std::runtime_error("gsbh_read_pfm: buffer too small");
            std::memcpy(rgba, img.data(), img.size() * sizeof(float));
        }
    });
}
int gsbh_image_diff(const float* rgba_a, const float* rgba_b, uint32_t w, uint32_t h, double out5[5]) {
    return guarded([&] {
        const gsbhost::ImageDiff d = gsbhost::diffImages(rgba_a, rgba_b, w, h);
        out5[0] = d.relMse; out5[1] = d.rmse; out5[2] = d.meanA; out5[3] = d.meanB; out5[4] = d.maxAbs;
    });
}
int gsbh_write_png(const char* path, const uint8_t* rgba8, uint32_t w, uint32_t h) { return guarded([&] { gsbhost::writePNG(path, rgba8, w, h); }); }

}  // extern "C"
