// gltf.cpp — glTF 2.0 scene input (.gltf with external / data-URI buffers, and .glb).
//
// The reference only carries a dead include of a glTF parser (renderer/Renderer.cpp:4, no call site), so there is
// no reference behaviour to mirror; north_star lists "glTF scenes in" beside Mitsuba XML and OBJ.  This reader maps
// a glTF scene onto the same host Scene the Mitsuba loader fills (scene.h), with the conventions of
// engine/Loader.cpp where they apply: meshes are un-indexed triangle soups, one RenderObject per mesh primitive and
// node, emissive materials become one TriangleLight per triangle (Scene::addRenderObject).
//
// Supported: node hierarchy (matrix or TRS), triangle primitives (mode 4) with POSITION, optional NORMAL, optional
// indices (u8 / u16 / u32), float accessors with byteStride, the first perspective camera, materials through
// pbrMetallicRoughness factors (no textures, like the reference: uv is never used, rayhit.rchit:716),
// emissiveFactor (+ KHR_materials_emissive_strength), KHR_materials_transmission and KHR_materials_ior.
// Material mapping (documented in INTEGRATION.md): transmission > 0.5 -> smooth / rough dielectric; metallic >= 0.5 ->
// rough conductor (aluminium eta / k, reflectance = base colour) or a perfect mirror when smooth and white;
// otherwise roughness >= 0.7 -> diffuse, else rough plastic.
#include <cmath>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>

#include "scene.h"

namespace gsbhost {

namespace {

// ---- minimal JSON -----------------------------------------------------------------------------------------------
struct Json {
    enum Kind { Null, Bool, Num, Str, Arr, Obj } kind = Null;
    bool b = false;
    double num = 0.0;
    std::string str;
    std::vector<Json> arr;
    std::vector<std::pair<std::string, Json>> obj;

    const Json* get(const char* key) const {
        if (kind != Obj) return nullptr;
        for (auto& kv : obj)
            if (kv.first == key) return &kv.second;
        return nullptr;
    }
    double number(const char* key, double def) const {
        const Json* j = get(key);
        return (j && j->kind == Num) ? j->num : def;
    }
    int index(const char* key) const {   // -1 when absent
        const Json* j = get(key);
        return (j && j->kind == Num) ? (int)j->num : -1;
    }
    std::string string(const char* key, const std::string& def = "") const {
        const Json* j = get(key);
        return (j && j->kind == Str) ? j->str : def;
    }
    size_t size() const { return kind == Arr ? arr.size() : 0; }
};

struct JsonParser {
    const std::string& s;
    size_t i = 0;
    explicit JsonParser(const std::string& text) : s(text) {}
    [[noreturn]] void fail(const char* what) const { throw std::runtime_error(std::string("glTF JSON: ") + what + " at byte " + std::to_string(i)); }
    void ws() {
        while (i < s.size() && (s[i] == ' ' || s[i] == '\n' || s[i] == '\r' || s[i] == '\t')) i++;
    }
    Json parse() {
        ws();
        if (i >= s.size()) fail("unexpected end");
        Json j;
        const char c = s[i];
        if (c == '{') {
            j.kind = Json::Obj;
            i++;
            ws();
            if (i < s.size() && s[i] == '}') { i++; return j; }
            for (;;) {
                ws();
                if (i >= s.size() || s[i] != '"') fail("expected a key");
                std::string key = parseString();
                ws();
                if (i >= s.size() || s[i] != ':') fail("expected ':'");
                i++;
                j.obj.emplace_back(std::move(key), parse());
                ws();
                if (i < s.size() && s[i] == ',') { i++; continue; }
                if (i < s.size() && s[i] == '}') { i++; break; }
                fail("expected ',' or '}'");
            }
        } else if (c == '[') {
            j.kind = Json::Arr;
            i++;
            ws();
            if (i < s.size() && s[i] == ']') { i++; return j; }
            for (;;) {
                j.arr.push_back(parse());
                ws();
                if (i < s.size() && s[i] == ',') { i++; continue; }
                if (i < s.size() && s[i] == ']') { i++; break; }
                fail("expected ',' or ']'");
            }
        } else if (c == '"') {
            j.kind = Json::Str;
            j.str = parseString();
        } else if (s.compare(i, 4, "true") == 0) {
            j.kind = Json::Bool; j.b = true; i += 4;
        } else if (s.compare(i, 5, "false") == 0) {
            j.kind = Json::Bool; j.b = false; i += 5;
        } else if (s.compare(i, 4, "null") == 0) {
            i += 4;
        } else {
            size_t start = i;
            while (i < s.size() && (std::isdigit((unsigned char)s[i]) || s[i] == '-' || s[i] == '+' || s[i] == '.' || s[i] == 'e' || s[i] == 'E')) i++;
            if (i == start) fail("unexpected character");
            j.kind = Json::Num;
            j.num = std::strtod(s.substr(start, i - start).c_str(), nullptr);
        }
        return j;
    }
    std::string parseString() {
        std::string out;
        i++;  // opening quote
        while (i < s.size() && s[i] != '"') {
            if (s[i] == '\\' && i + 1 < s.size()) {
                const char e = s[i + 1];
                i += 2;
                switch (e) {
                    case 'n': out += '\n'; break;
                    case 't': out += '\t'; break;
                    case 'r': out += '\r'; break;
                    case 'b': out += '\b'; break;
                    case 'f': out += '\f'; break;
                    case 'u': out += '?'; i += 4; break;   // names only; no need for full unicode here
                    default: out += e; break;
                }
            } else {
                out += s[i++];
            }
        }
        if (i >= s.size()) fail("unterminated string");
        i++;
        return out;
    }
};

std::string readBinary(const std::string& path) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open " + path);
    std::ostringstream ss;
    ss << f.rdbuf();
    return ss.str();
}

std::string decodeBase64(const std::string& in) {
    std::string out;
    uint32_t acc = 0;
    int bits = 0;
    for (char c : in) {
        int v;
        if (c >= 'A' && c <= 'Z') v = c - 'A';
        else if (c >= 'a' && c <= 'z') v = c - 'a' + 26;
        else if (c >= '0' && c <= '9') v = c - '0' + 52;
        else if (c == '+' || c == '-') v = 62;
        else if (c == '/' || c == '_') v = 63;
        else continue;   // '=', whitespace
        acc = (acc << 6) | (uint32_t)v;
        bits += 6;
        if (bits >= 8) {
            bits -= 8;
            out += (char)((acc >> bits) & 0xffu);
        }
    }
    return out;
}

std::string dirOfPath(const std::string& path) {
    const size_t k = path.find_last_of("/\\");
    return k == std::string::npos ? std::string(".") : path.substr(0, k);
}

// ---- document ---------------------------------------------------------------------------------------------------
struct Doc {
    Json root;
    std::vector<std::string> buffers;
    std::string baseDir;
    Scene* scene = nullptr;
    std::map<int, uint32_t> materialBsdf;                // glTF material -> host material index
    std::map<std::pair<int, int>, uint32_t> meshCache;   // (mesh, primitive) -> host mesh
    bool haveCamera = false;

    const Json& at(const char* array, int i) const {
        const Json* a = root.get(array);
        if (!a || i < 0 || (size_t)i >= a->size()) throw std::runtime_error(std::string("glTF: index out of range in '") + array + "'");
        return a->arr[(size_t)i];
    }
};

struct View {   // one accessor resolved to bytes
    const uint8_t* data = nullptr;
    size_t count = 0, stride = 0;
    int componentType = 0, components = 0;
};

View accessorView(const Doc& d, int accessor) {
    const Json& a = d.at("accessors", accessor);
    if (a.get("sparse")) throw std::runtime_error("glTF: sparse accessors are not supported");
    const Json& bv = d.at("bufferViews", a.index("bufferView"));
    const int buffer = bv.index("buffer");
    if (buffer < 0 || (size_t)buffer >= d.buffers.size()) throw std::runtime_error("glTF: bufferView without a buffer");
    View v;
    v.componentType = (int)a.number("componentType", 0);
    const std::string type = a.string("type");
    v.components = type == "SCALAR" ? 1 : type == "VEC2" ? 2 : type == "VEC3" ? 3 : type == "VEC4" ? 4 : type == "MAT4" ? 16 : 0;
    const size_t compSize = (v.componentType == 5120 || v.componentType == 5121) ? 1 : (v.componentType == 5122 || v.componentType == 5123) ? 2 : 4;
    if (!v.components) throw std::runtime_error("glTF: unsupported accessor type " + type);
    v.count = (size_t)a.number("count", 0);
    const size_t offset = (size_t)bv.number("byteOffset", 0) + (size_t)a.number("byteOffset", 0);
    v.stride = (size_t)bv.number("byteStride", 0);
    if (!v.stride) v.stride = compSize * (size_t)v.components;
    const std::string& buf = d.buffers[(size_t)buffer];
    if (v.count && offset + (v.count - 1) * v.stride + compSize * (size_t)v.components > buf.size())
        throw std::runtime_error("glTF: accessor reaches past the end of its buffer");
    v.data = reinterpret_cast<const uint8_t*>(buf.data()) + offset;
    return v;
}

void readVec3(const View& v, size_t i, float out[3]) {
    if (v.componentType != 5126 || v.components < 3) throw std::runtime_error("glTF: POSITION / NORMAL must be float VEC3");
    std::memcpy(out, v.data + i * v.stride, 12);
}

uint32_t readIndex(const View& v, size_t i) {
    const uint8_t* p = v.data + i * v.stride;
    switch (v.componentType) {
        case 5121: return *p;
        case 5123: { uint16_t x; std::memcpy(&x, p, 2); return x; }
        case 5125: { uint32_t x; std::memcpy(&x, p, 4); return x; }
    }
    throw std::runtime_error("glTF: indices must be unsigned byte / short / int");
}

// one triangle-list primitive -> un-indexed soup (Mesh.cpp:7-22 of the reference consumes exactly that)
uint32_t primitiveMesh(Doc& d, int meshIndex, int primIndex, const Json& prim, const std::string& name) {
    auto key = std::make_pair(meshIndex, primIndex);
    auto it = d.meshCache.find(key);
    if (it != d.meshCache.end()) return it->second;
    const Json* attrs = prim.get("attributes");
    const int posAcc = attrs ? attrs->index("POSITION") : -1;
    if (posAcc < 0) throw std::runtime_error("glTF: primitive without POSITION");
    const View pos = accessorView(d, posAcc);
    const int nrmAcc = attrs->index("NORMAL");
    View nrm;
    if (nrmAcc >= 0) nrm = accessorView(d, nrmAcc);
    const int idxAcc = prim.index("indices");
    View idx;
    if (idxAcc >= 0) idx = accessorView(d, idxAcc);
    const size_t corners = idxAcc >= 0 ? idx.count : pos.count;
    MeshData m;
    m.name = "gltf:" + name + ":" + std::to_string(primIndex);
    m.positions.reserve(corners * 3);
    m.normals.reserve(corners * 3);
    for (size_t t = 0; t + 2 < corners; t += 3) {
        float p[3][3], n[3][3];
        uint32_t vi[3];
        for (int k = 0; k < 3; k++) {
            vi[k] = idxAcc >= 0 ? readIndex(idx, t + k) : (uint32_t)(t + k);
            if (vi[k] >= pos.count) throw std::runtime_error("glTF: vertex index out of range");
            readVec3(pos, vi[k], p[k]);
        }
        float fn[3] = {(p[1][1] - p[0][1]) * (p[2][2] - p[0][2]) - (p[1][2] - p[0][2]) * (p[2][1] - p[0][1]),
                       (p[1][2] - p[0][2]) * (p[2][0] - p[0][0]) - (p[1][0] - p[0][0]) * (p[2][2] - p[0][2]),
                       (p[1][0] - p[0][0]) * (p[2][1] - p[0][1]) - (p[1][1] - p[0][1]) * (p[2][0] - p[0][0])};
        const float fl = std::sqrt(fn[0] * fn[0] + fn[1] * fn[1] + fn[2] * fn[2]);
        if (fl > 0.0f) { fn[0] /= fl; fn[1] /= fl; fn[2] /= fl; }
        for (int k = 0; k < 3; k++) {
            bool have = false;
            if (nrmAcc >= 0 && vi[k] < nrm.count) {
                readVec3(nrm, vi[k], n[k]);
                have = n[k][0] * n[k][0] + n[k][1] * n[k][1] + n[k][2] * n[k][2] > 1e-12f;
            }
            if (!have) { n[k][0] = fn[0]; n[k][1] = fn[1]; n[k][2] = fn[2]; }
            m.positions.insert(m.positions.end(), p[k], p[k] + 3);
            m.normals.insert(m.normals.end(), n[k], n[k] + 3);
        }
    }
    const uint32_t id = d.scene->addMesh(std::move(m));
    d.meshCache[key] = id;
    return id;
}

// pbrMetallicRoughness factors -> one of the BSDFs of the shader library (see the header comment)
uint32_t materialFor(Doc& d, int material) {
    auto it = d.materialBsdf.find(material);
    if (it != d.materialBsdf.end()) return it->second;
    Scene& s = *d.scene;
    Material m;
    m.twofaced = true;
    m.hasBsdf = true;
    float base[3] = {0.8f, 0.8f, 0.8f};
    float metallic = 0.0f, roughness = 1.0f, transmission = 0.0f, ior = 1.5f, strength = 1.0f;
    if (material >= 0) {
        const Json& jm = d.at("materials", material);
        metallic = 1.0f;   // glTF defaults when pbrMetallicRoughness is present but a factor is not
        if (const Json* pbr = jm.get("pbrMetallicRoughness")) {
            if (const Json* c = pbr->get("baseColorFactor"))
                for (size_t k = 0; k < 3 && k < c->size(); k++) base[k] = (float)c->arr[k].num;
            else base[0] = base[1] = base[2] = 1.0f;
            metallic = (float)pbr->number("metallicFactor", 1.0);
            roughness = (float)pbr->number("roughnessFactor", 1.0);
            if (pbr->get("baseColorTexture") || pbr->get("metallicRoughnessTexture")) s.warnings.push_back("glTF: textures are ignored (factors only)");
        }
        if (const Json* ext = jm.get("extensions")) {
            if (const Json* t = ext->get("KHR_materials_transmission")) transmission = (float)t->number("transmissionFactor", 0.0);
            if (const Json* t = ext->get("KHR_materials_ior")) ior = (float)t->number("ior", 1.5);
            if (const Json* t = ext->get("KHR_materials_emissive_strength")) strength = (float)t->number("emissiveStrength", 1.0);
        }
        if (const Json* e = jm.get("emissiveFactor"))
            for (size_t k = 0; k < 3 && k < e->size(); k++) m.emission[k] = (float)e->arr[k].num * strength;
        if (jm.get("doubleSided") && jm.get("doubleSided")->kind == Json::Bool) m.twofaced = jm.get("doubleSided")->b;
    }
    const float alpha = std::fmax(roughness * roughness, 0.01f);
    float R0 = (ior - 1.0f) / (ior + 1.0f);
    R0 *= R0;
    if (transmission > 0.5f) {
        m.twofaced = false;   // the sign of the normal tells inside from outside
        if (roughness < 0.05f) m.bsdf = s.addSmoothDielectricBSDF({ior, 1.0f});
        else m.bsdf = s.addRoughDielectricBSDF({ior, 1.0f, alpha, 0});
    } else if (metallic >= 0.5f) {
        const bool white = base[0] > 0.98f && base[1] > 0.98f && base[2] > 0.98f;
        if (roughness < 0.05f && white) m.bsdf = s.addSmoothConductorBSDF({0.0f, 1.0f});   // iorIn = 0: Fr = 1 (Loader.cpp:193-195)
        else m.bsdf = s.addRoughConductorBSDF({{1.65746f, 0.880369f, 0.521229f}, {9.22387f, 6.26952f, 4.837f}, {base[0], base[1], base[2]}, alpha, 0});
    } else if (roughness >= 0.7f) {
        m.bsdf = s.addDiffuseBSDF({{base[0], base[1], base[2]}, 0});
    } else {
        m.bsdf = s.addRoughPlasticBSDF({{base[0], base[1], base[2]}, ior, 1.0f, R0, alpha, 0});
    }
    if (m.emission[0] > 0.0f || m.emission[1] > 0.0f || m.emission[2] > 0.0f) m.twofaced = true;
    const uint32_t id = (uint32_t)s.addMaterial(m);
    d.materialBsdf[material] = id;
    return id;
}

void nodeLocal(const Json& node, float* out) {
    matIdentity(out);
    if (const Json* m = node.get("matrix")) {
        if (m->size() == 16)
            for (int k = 0; k < 16; k++) out[k] = (float)m->arr[(size_t)k].num;   // glTF matrices are column-major
        return;
    }
    float T[16], R[16], S[16], TR[16];
    matIdentity(T); matIdentity(R); matIdentity(S);
    if (const Json* t = node.get("translation"))
        if (t->size() == 3) matTranslate((float)t->arr[0].num, (float)t->arr[1].num, (float)t->arr[2].num, T);
    if (const Json* s = node.get("scale"))
        if (s->size() == 3) matScale((float)s->arr[0].num, (float)s->arr[1].num, (float)s->arr[2].num, S);
    if (const Json* q = node.get("rotation"))
        if (q->size() == 4) {
            const double x = q->arr[0].num, y = q->arr[1].num, z = q->arr[2].num, w = q->arr[3].num;
            const double r[9] = {1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
                                 2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
                                 2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)};
            for (int row = 0; row < 3; row++)
                for (int col = 0; col < 3; col++) R[col * 4 + row] = (float)r[row * 3 + col];
        }
    matMul(T, R, TR);
    matMul(TR, S, out);
}

void visitNode(Doc& d, int nodeIndex, const float* parent, int depth) {
    if (depth > 64) throw std::runtime_error("glTF: node hierarchy deeper than 64 (cycle?)");
    const Json& node = d.at("nodes", nodeIndex);
    float local[16], world[16];
    nodeLocal(node, local);
    matMul(parent, local, world);
    Scene& s = *d.scene;
    const int mesh = node.index("mesh");
    if (mesh >= 0) {
        const Json& jm = d.at("meshes", mesh);
        const Json* prims = jm.get("primitives");
        for (size_t p = 0; prims && p < prims->size(); p++) {
            const Json& prim = prims->arr[p];
            const int mode = (int)prim.number("mode", 4);
            if (mode != 4) { s.warnings.push_back("glTF: primitive mode " + std::to_string(mode) + " skipped (triangles only)"); continue; }
            RenderObject o;
            std::memcpy(o.transform, world, sizeof(world));
            o.mesh = primitiveMesh(d, mesh, (int)p, prim, jm.string("name", std::to_string(mesh)));
            o.material = (int)materialFor(d, prim.index("material"));
            s.addRenderObject(o);
        }
    }
    const int cam = node.index("camera");
    if (cam >= 0 && !d.haveCamera) {
        const Json& jc = d.at("cameras", cam);
        if (const Json* persp = jc.get("perspective")) {
            // glTF cameras look down -Z with +Y up in their node's frame
            const float origin[3] = {world[12], world[13], world[14]};
            const float target[3] = {origin[0] - world[8], origin[1] - world[9], origin[2] - world[10]};
            const float up[3] = {world[4], world[5], world[6]};
            matReferenceCamera(origin, target, up, s.camera.toWorld);
            const double yfov = persp->number("yfov", 0.7);
            const double aspect = persp->number("aspectRatio", (double)s.filmWidth / (double)s.filmHeight);
            // the reference's fov spans the larger image axis (raygen.rgen:20-25)
            s.camera.fov = (float)(aspect > 1.0 ? 2.0 * std::atan(std::tan(yfov / 2.0) * aspect) : yfov);
            s.filmHeight = 1080;
            s.filmWidth = (uint32_t)std::lround(1080.0 * aspect);
            d.haveCamera = true;
        }
    }
    if (const Json* kids = node.get("children"))
        for (const Json& k : kids->arr) visitNode(d, (int)k.num, world, depth + 1);
}

// no camera in the file: look at the scene bounds from +Z
void frameScene(Scene& s) {
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    for (const RenderObject& o : s.renderObjects) {
        const MeshData& m = s.meshes[o.mesh];
        for (size_t v = 0; v + 2 < m.positions.size(); v += 3)
            for (int r = 0; r < 3; r++) {
                const float w = o.transform[0 * 4 + r] * m.positions[v] + o.transform[1 * 4 + r] * m.positions[v + 1] +
                                o.transform[2 * 4 + r] * m.positions[v + 2] + o.transform[3 * 4 + r];
                lo[r] = std::fmin(lo[r], w);
                hi[r] = std::fmax(hi[r], w);
            }
    }
    if (lo[0] > hi[0]) return;
    const float c[3] = {0.5f * (lo[0] + hi[0]), 0.5f * (lo[1] + hi[1]), 0.5f * (lo[2] + hi[2])};
    const float radius = 0.5f * std::sqrt((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1]) + (hi[2] - lo[2]) * (hi[2] - lo[2]));
    s.camera.fov = 0.8f;
    const float origin[3] = {c[0], c[1], c[2] + radius / std::tan(0.5f * 0.8f * 0.5625f) * 1.05f + 1e-3f};
    const float up[3] = {0.0f, 1.0f, 0.0f};
    matReferenceCamera(origin, c, up, s.camera.toWorld);
    s.warnings.push_back("glTF: no perspective camera, framing the scene bounds from +Z");
}

}  // namespace

Scene loadGltfScene(const std::string& path) {
    Doc d;
    d.baseDir = dirOfPath(path);
    std::string file = readBinary(path);
    std::string json, glbBin;
    if (file.size() >= 12 && std::memcmp(file.data(), "glTF", 4) == 0) {
        // GLB container: 12-byte header, then chunks {length, type, data}; JSON first, BIN optional
        size_t off = 12;
        while (off + 8 <= file.size()) {
            uint32_t len, type;
            std::memcpy(&len, file.data() + off, 4);
            std::memcpy(&type, file.data() + off + 4, 4);
            if (off + 8 + len > file.size()) throw std::runtime_error("glTF: truncated GLB chunk");
            if (type == 0x4E4F534Au) json.assign(file.data() + off + 8, len);
            else if (type == 0x004E4942u) glbBin.assign(file.data() + off + 8, len);
            off += 8 + ((size_t)len + 3) / 4 * 4;
        }
        if (json.empty()) throw std::runtime_error("glTF: GLB without a JSON chunk");
    } else {
        json.swap(file);
    }
    JsonParser parser(json);
    d.root = parser.parse();
    if (const Json* asset = d.root.get("asset"))
        if (asset->string("version", "2.0").substr(0, 1) != "2") throw std::runtime_error("glTF: only version 2.x is supported");
    if (const Json* bufs = d.root.get("buffers"))
        for (const Json& b : bufs->arr) {
            const std::string uri = b.string("uri");
            if (uri.empty()) d.buffers.push_back(glbBin);
            else if (uri.compare(0, 5, "data:") == 0) {
                const size_t comma = uri.find(',');
                d.buffers.push_back(comma == std::string::npos ? std::string() : decodeBase64(uri.substr(comma + 1)));
            } else d.buffers.push_back(readBinary(d.baseDir + "/" + uri));
        }
    Scene s;
    s.filmWidth = 1920;
    s.filmHeight = 1080;
    d.scene = &s;
    float identity[16];
    matIdentity(identity);
    const Json* scenes = d.root.get("scenes");
    int sceneIndex = d.root.index("scene");
    if (sceneIndex < 0) sceneIndex = 0;
    if (scenes && (size_t)sceneIndex < scenes->size()) {
        if (const Json* roots = scenes->arr[(size_t)sceneIndex].get("nodes"))
            for (const Json& n : roots->arr) visitNode(d, (int)n.num, identity, 0);
    } else if (const Json* nodes = d.root.get("nodes")) {
        // no scene object: every node that is nobody's child is a root
        std::vector<bool> child(nodes->size(), false);
        for (const Json& n : nodes->arr)
            if (const Json* kids = n.get("children"))
                for (const Json& k : kids->arr)
                    if ((size_t)k.num < child.size()) child[(size_t)k.num] = true;
        for (size_t n = 0; n < nodes->size(); n++)
            if (!child[n]) visitNode(d, (int)n, identity, 0);
    }
    if (!d.haveCamera) frameScene(s);
    if (s.triangleLights.empty()) s.warnings.push_back("glTF: no emissive material, the image will be black (punctual lights are not supported)");
    return s;
}

}  // namespace gsbhost
