// tpm_glue.cpp — the REFERENCE'S OWN scene parser (external/TinyParser-Mitsuba, the library engine/Loader.cpp:253-349 reads
// Mitsuba XML with) behind one C function that dumps the parsed object tree as JSON.  TEST INFRASTRUCTURE ONLY
// (oracle/_ref/libtpm.so, built by oracle/Makefile from the sources where they lie under /root/reference; nothing of the
// reference is copied into the repository).  tests/test_loader_vs_reference_cpu.py cross-checks the product's own XML
// reader + loader (gpuspectral_b200/host/xml.cpp, loader.cpp) against it on the reference's scene files: instance
// matrices, BSDF parameters, emitters, camera.
#include <tinyparser-mitsuba.h>

#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>

namespace tpm = tinyparser_mitsuba;

namespace {

void num(std::ostringstream& o, double v) {
    char buf[40];
    std::snprintf(buf, sizeof(buf), "%.9g", v);
    o << buf;
}

void str(std::ostringstream& o, const std::string& s) {
    o << '"';
    for (char c : s) {
        if (c == '"' || c == '\\') o << '\\' << c;
        else if ((unsigned char)c < 0x20) o << ' ';
        else o << c;
    }
    o << '"';
}

void property(std::ostringstream& o, const tpm::Property& p) {
    switch (p.type()) {
    case tpm::PT_BOOL: o << "{\"t\":\"bool\",\"v\":" << (p.getBool() ? "true" : "false") << "}"; break;
    case tpm::PT_INTEGER: o << "{\"t\":\"int\",\"v\":" << (long long)p.getInteger() << "}"; break;
    case tpm::PT_NUMBER: o << "{\"t\":\"num\",\"v\":"; num(o, p.getNumber()); o << "}"; break;
    case tpm::PT_STRING: o << "{\"t\":\"str\",\"v\":"; str(o, p.getString()); o << "}"; break;
    case tpm::PT_COLOR: {
        const tpm::Color& c = p.getColor();
        o << "{\"t\":\"rgb\",\"v\":["; num(o, c.r); o << ","; num(o, c.g); o << ","; num(o, c.b); o << "]}";
        break;
    }
    case tpm::PT_VECTOR: {
        const tpm::Vector& v = p.getVector();
        o << "{\"t\":\"vec\",\"v\":["; num(o, v.x); o << ","; num(o, v.y); o << ","; num(o, v.z); o << "]}";
        break;
    }
    case tpm::PT_TRANSFORM: {
        const tpm::Transform& t = p.getTransform();
        o << "{\"t\":\"transform\",\"v\":[";          // row major, as the library stores it
        for (int i = 0; i < 16; i++) { if (i) o << ","; num(o, t.matrix[i]); }
        o << "]}";
        break;
    }
    default: o << "{\"t\":\"other\"}"; break;
    }
}

void object(std::ostringstream& o, const tpm::Object& obj) {
    o << "{\"type\":" << (int)obj.type() << ",\"plugin\":";
    str(o, obj.pluginType());
    o << ",\"id\":";
    str(o, obj.id());
    o << ",\"props\":{";
    bool first = true;
    for (const auto& kv : obj.properties()) {
        if (!first) o << ",";
        first = false;
        str(o, kv.first);
        o << ":";
        property(o, kv.second);
    }
    o << "},\"named\":{";
    first = true;
    for (const auto& kv : obj.namedChildren()) {
        if (!first) o << ",";
        first = false;
        str(o, kv.first);
        o << ":";
        object(o, *kv.second);
    }
    o << "},\"children\":[";
    first = true;
    for (const auto& c : obj.anonymousChildren()) {
        if (!first) o << ",";
        first = false;
        object(o, *c);
    }
    o << "]}";
}

}  // namespace

// Parses `path` with the reference's parser; writes the JSON dump (NUL-terminated) into out (capacity cap).
// Returns the number of bytes needed (including the NUL), or -1 on a parse error (message in out).
extern "C" __attribute__((visibility("default"))) long tpm_dump_scene(const char* path, char* out, long cap) {
    std::string text;
    long rc = 0;
    try {
        tpm::SceneLoader loader;
        tpm::Scene scene = loader.loadFromFile(path);
        std::ostringstream o;
        object(o, scene);
        text = o.str();
        rc = (long)text.size() + 1;
    } catch (const std::exception& e) {
        text = e.what();
        rc = -1;
    }
    if (out && cap > 0) {
        const size_t n = std::min<size_t>(text.size(), (size_t)cap - 1);
        std::memcpy(out, text.data(), n);
        out[n] = 0;
    }
    return rc;
}
