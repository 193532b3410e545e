#include "path_tracer.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <stdexcept>

namespace gsbhost {

static void check(gsb_status st, const char* what) {
    if (st != GSB_OK) throw std::runtime_error(std::string(what) + ": " + gsb_last_error());
}

PathTracer::PathTracer(int device_ordinal) { check(gsb_create(device_ordinal, &ctx), "gsb_create"); }

PathTracer::~PathTracer() {
    if (ctx) gsb_destroy(ctx);
}

void PathTracer::setup(uint32_t width, uint32_t height) {
    w = width;
    h = height;
    timestamp = 0;
    check(gsb_clear(ctx), "gsb_clear");
}

void PathTracer::setParams(const gsb_params& p) { check(gsb_set_params(ctx, &p), "gsb_set_params"); }

// prepareScene (PathTracer.cpp:58-93) + getOrCreateBLAS (Renderer.cpp:122-131) + createTLAS,
// done once per scene instead of once per frame.
void PathTracer::uploadScene(const Scene& scene) {
    for (uint32_t mh : meshHandles) gsb_mesh_destroy(ctx, mh);
    meshHandles.clear();
    for (const MeshData& m : scene.meshes) {
        uint32_t handle = 0;
        const float* uv = m.uvs.size() == (size_t)m.vertexCount() * 2 ? m.uvs.data() : nullptr;
        check(gsb_mesh_create_uv(ctx, m.positions.data(), m.normals.data(), uv, m.vertexCount(), &handle), "gsb_mesh_create_uv");
        meshHandles.push_back(handle);
    }
    for (uint32_t type = 0; type < GSB_BSDF_TYPE_COUNT; type++) {
        uint32_t count = 0, stride = 0;
        const void* data = scene.bsdfData(type, &count, &stride);
        check(gsb_scene_set_bsdfs(ctx, type, data, count), "gsb_scene_set_bsdfs");
    }
    {
        const std::vector<gsb_texture> tex = scene.textureRecords();
        check(gsb_scene_set_textures(ctx, tex.data(), (uint32_t)tex.size()), "gsb_scene_set_textures");
    }
    {
        const gsb_environment env = scene.environmentRecord();
        check(gsb_scene_set_environment(ctx, &env), "gsb_scene_set_environment");
    }
    check(gsb_scene_set_lights(ctx, scene.triangleLights.data(), (uint32_t)scene.triangleLights.size()), "gsb_scene_set_lights");
    std::vector<gsb_instance> inst = scene.instanceRecords();
    for (gsb_instance& i : inst) i.mesh = meshHandles.at(i.mesh);
    check(gsb_scene_set_instances(ctx, inst.data(), (uint32_t)inst.size()), "gsb_scene_set_instances");
    const float* T = scene.camera.toWorld;
    const float eye[3] = {T[12], T[13], T[14]};  // Camera::getPosition = toWorld[3]
    check(gsb_set_camera(ctx, eye, T, scene.camera.fov), "gsb_set_camera");
    uploaded = &scene;
    uploadedObjects = scene.renderObjects.size();
    uploadedRevision = scene.revision;
}

void PathTracer::render(const Scene& scene, uint32_t spp, uint32_t stride) {
    if (w == 0 || h == 0) throw std::runtime_error("PathTracer::setup was not called");
    // The reference re-reads everything every frame (PathTracer.cpp:58-93).  Here the scene arrays are uploaded when the
    // Scene object, its object count or its revision (Scene::touch(), to be called after editing transforms, materials or
    // lights in place) changes; the camera is cheap host state and is set on every call.
    if (uploaded != &scene || uploadedObjects != scene.renderObjects.size() || uploadedRevision != scene.revision) uploadScene(scene);
    const float* T = scene.camera.toWorld;
    const float eye[3] = {T[12], T[13], T[14]};  // Camera::getPosition = toWorld[3]
    check(gsb_set_camera(ctx, eye, T, scene.camera.fov), "gsb_set_camera");
    check(gsb_render(ctx, w, h, (uint32_t)timestamp, spp, stride), "gsb_render");
    timestamp += (int)(spp * stride);
}

void PathTracer::createRenderPass(const Scene& scene) { render(scene, 1, 1); }

void PathTracer::readHDR(std::vector<float>& rgba) {
    rgba.resize((size_t)w * h * 4);
    check(gsb_read_hdr(ctx, rgba.data()), "gsb_read_hdr");
}

void PathTracer::readLDR(std::vector<uint8_t>& rgba8) {
    rgba8.resize((size_t)w * h * 4);
    check(gsb_read_ldr(ctx, rgba8.data()), "gsb_read_ldr");
}

void PathTracer::synchronize() { check(gsb_synchronize(ctx), "gsb_synchronize"); }

// ---- image writers -----------------------------------------------------------------------------
void writePFM(const std::string& path, const float* rgba, uint32_t width, uint32_t height) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::fprintf(f, "PF\n%u %u\n-1.0\n", width, height);  // little endian, rows bottom-to-top
    std::vector<float> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; y++) {
        const float* src = rgba + (size_t)(height - 1 - y) * width * 4;
        for (uint32_t x = 0; x < width; x++) {
            row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2];
        }
        std::fwrite(row.data(), sizeof(float), row.size(), f);
    }
    std::fclose(f);
}

void readPFM(const std::string& path, std::vector<float>& rgba, uint32_t& width, uint32_t& height) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot read " + path);
    char magic[3] = {0, 0, 0};
    int w = 0, h = 0;
    float scale = 0.0f;
    if (std::fscanf(f, "%2s %d %d %f", magic, &w, &h, &scale) != 4 || (std::strcmp(magic, "PF") && std::strcmp(magic, "Pf")) || w <= 0 || h <= 0 || scale == 0.0f) {
        std::fclose(f);
        throw std::runtime_error(path + ": not a PFM file");
    }
    std::fgetc(f);   // the single whitespace byte after the scale
    const int ch = magic[1] == 'F' ? 3 : 1;
    std::vector<float> row((size_t)w * ch);
    rgba.assign((size_t)w * h * 4, 1.0f);
    const bool swap = scale > 0.0f;   // positive scale = big endian
    for (int y = 0; y < h; y++) {
        if (std::fread(row.data(), sizeof(float), row.size(), f) != row.size()) {
            std::fclose(f);
            throw std::runtime_error(path + ": truncated PFM file");
        }
        if (swap)
            for (float& v : row) {
                unsigned char* b = reinterpret_cast<unsigned char*>(&v);
                std::swap(b[0], b[3]);
                std::swap(b[1], b[2]);
            }
        float* dst = rgba.data() + (size_t)(h - 1 - y) * w * 4;   // PFM rows run bottom-up
        for (int x = 0; x < w; x++)
            for (int c = 0; c < 3; c++) dst[4 * x + c] = row[(size_t)x * ch + (ch == 3 ? c : 0)];
    }
    std::fclose(f);
    width = (uint32_t)w;
    height = (uint32_t)h;
}

ImageDiff diffImages(const float* a, const float* b, uint32_t width, uint32_t height) {
    ImageDiff d = {0.0, 0.0, 0.0, 0.0, 0.0};
    const size_t n = (size_t)width * height;
    for (size_t i = 0; i < n; i++)
        for (int c = 0; c < 3; c++) {
            const double x = a[4 * i + c], y = b[4 * i + c], e = x - y;
            d.relMse += e * e / (y * y + 1e-2);
            d.rmse += e * e;
            d.meanA += x;
            d.meanB += y;
            d.maxAbs = std::max(d.maxAbs, std::fabs(e));
        }
    const double m = (double)(n ? n * 3 : 1);
    d.relMse /= m;
    d.rmse = std::sqrt(d.rmse / m);
    d.meanA /= m;
    d.meanB /= m;
    return d;
}

void writePPM(const std::string& path, const uint8_t* rgba8, uint32_t width, uint32_t height) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::fprintf(f, "P6\n%u %u\n255\n", width, height);
    std::vector<uint8_t> row((size_t)width * 3);
    for (uint32_t y = 0; y < height; y++) {
        const uint8_t* src = rgba8 + (size_t)y * width * 4;
        for (uint32_t x = 0; x < width; x++) {
            row[3 * x] = src[4 * x]; row[3 * x + 1] = src[4 * x + 1]; row[3 * x + 2] = src[4 * x + 2];
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    std::fclose(f);
}

// PNG with stored (uncompressed) deflate blocks: no zlib needed.
namespace {
uint32_t crcTable[256];
bool crcReady = false;
uint32_t crc32(uint32_t crc, const uint8_t* p, size_t n) {
    if (!crcReady) {
        for (uint32_t i = 0; i < 256; i++) {
            uint32_t c = i;
            for (int k = 0; k < 8; k++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1;
            crcTable[i] = c;
        }
        crcReady = true;
    }
    crc = ~crc;
    for (size_t i = 0; i < n; i++) crc = crcTable[(crc ^ p[i]) & 0xff] ^ (crc >> 8);
    return ~crc;
}
void be32(std::vector<uint8_t>& v, uint32_t x) {
    v.push_back((uint8_t)(x >> 24)); v.push_back((uint8_t)(x >> 16)); v.push_back((uint8_t)(x >> 8)); v.push_back((uint8_t)x);
}
void chunk(FILE* f, const char* tag, const std::vector<uint8_t>& data) {
    std::vector<uint8_t> buf;
    be32(buf, (uint32_t)data.size());
    buf.insert(buf.end(), tag, tag + 4);
    buf.insert(buf.end(), data.begin(), data.end());
    uint32_t c = crc32(0, buf.data() + 4, buf.size() - 4);
    be32(buf, c);
    std::fwrite(buf.data(), 1, buf.size(), f);
}
}  // namespace

void writePNG(const std::string& path, const uint8_t* rgba8, uint32_t width, uint32_t height) {
    FILE* f = std::fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    const uint8_t sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    std::fwrite(sig, 1, 8, f);
    std::vector<uint8_t> ihdr;
    be32(ihdr, width);
    be32(ihdr, height);
    ihdr.push_back(8); ihdr.push_back(6); ihdr.push_back(0); ihdr.push_back(0); ihdr.push_back(0);  // 8-bit RGBA
    chunk(f, "IHDR", ihdr);
    std::vector<uint8_t> srgb(1, 0);  // rendering intent: the bytes are sRGB-encoded
    chunk(f, "sRGB", srgb);
    std::vector<uint8_t> raw;
    raw.reserve((size_t)height * ((size_t)width * 4 + 1));
    for (uint32_t y = 0; y < height; y++) {
        raw.push_back(0);
        raw.insert(raw.end(), rgba8 + (size_t)y * width * 4, rgba8 + (size_t)(y + 1) * width * 4);
    }
    std::vector<uint8_t> z;
    z.push_back(0x78); z.push_back(0x01);
    uint32_t a = 1, b = 0;
    for (size_t i = 0; i < raw.size(); i++) { a = (a + raw[i]) % 65521u; b = (b + a) % 65521u; }
    size_t pos = 0;
    while (pos < raw.size() || raw.empty()) {
        size_t n = std::min<size_t>(65535, raw.size() - pos);
        z.push_back(pos + n >= raw.size() ? 1 : 0);
        z.push_back((uint8_t)(n & 0xff)); z.push_back((uint8_t)(n >> 8));
        z.push_back((uint8_t)(~n & 0xff)); z.push_back((uint8_t)((~n >> 8) & 0xff));
        z.insert(z.end(), raw.begin() + pos, raw.begin() + pos + n);
        pos += n;
        if (raw.empty()) break;
    }
    be32(z, (b << 16) | a);
    chunk(f, "IDAT", z);
    chunk(f, "IEND", {});
    std::fclose(f);
}

}  // namespace gsbhost
