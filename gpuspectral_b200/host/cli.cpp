// gsb_render — headless command-line front end: Mitsuba XML (or a built-in scene) in, PFM / PNG out.
// Stands where the reference's main.cpp:15-30 + Window::run loop stand, without the window.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <string>
#include <vector>

#include "path_tracer.h"
#include "scene.h"

using namespace gsbhost;

static void usage() {
    std::fprintf(stderr,
                 "usage: gsb_render (<scene.xml | scene.gltf | scene.glb> | --builtin cornell|gallery|field) [--size WxH] [--spp N]\n"
                 "                  [--max-depth D] [--filter none|box|tent] [--device I] [--hdr out.pfm] [--ldr out.png]\n"
                 "       gsb_render --diff a.pfm reference.pfm      (relMSE, RMSE, means, max |a - b|; no GPU needed)\n");
}

int main(int argc, char** argv) {
    std::string xml, builtin, hdrOut, ldrOut;
    uint32_t W = 0, H = 0, spp = 16;
    int device = 0, maxDepth = -1, filter = 0;
    for (int i = 1; i < argc; i++) {
        std::string a = argv[i];
        auto next = [&]() -> const char* { return i + 1 < argc ? argv[++i] : ""; };
        if (a == "--diff") {
            // image comparison: the figures converged renders are judged by (north_star check 3)
            const std::string fa = next(), fb = next();
            try {
                std::vector<float> ia, ib;
                uint32_t wa, ha, wb, hb;
                readPFM(fa, ia, wa, ha);
                readPFM(fb, ib, wb, hb);
                if (wa != wb || ha != hb) { std::fprintf(stderr, "error: image sizes differ (%ux%u vs %ux%u)\n", wa, ha, wb, hb); return 1; }
                const ImageDiff d = diffImages(ia.data(), ib.data(), wa, ha);
                std::printf("%ux%u relMSE %.6g RMSE %.6g mean %.6g vs %.6g max|a-b| %.6g\n", wa, ha, d.relMse, d.rmse, d.meanA, d.meanB, d.maxAbs);
                return 0;
            } catch (const std::exception& e) {
                std::fprintf(stderr, "error: %s\n", e.what());
                return 1;
            }
        }
        if (a == "--builtin") builtin = next();
        else if (a == "--filter") { const std::string f = next(); filter = f == "box" ? 1 : (f == "tent" ? 2 : 0); }
        else if (a == "--size") { if (std::sscanf(next(), "%ux%u", &W, &H) != 2) { usage(); return 2; } }
        else if (a == "--spp") spp = (uint32_t)std::atoi(next());
        else if (a == "--max-depth") maxDepth = std::atoi(next());
        else if (a == "--device") device = std::atoi(next());
        else if (a == "--hdr") hdrOut = next();
        else if (a == "--ldr") ldrOut = next();
        else if (a == "-h" || a == "--help") { usage(); return 0; }
        else xml = a;
    }
    if (xml.empty() && builtin.empty()) { usage(); return 2; }
    try {
        Scene scene = !xml.empty() ? ((xml.size() > 4 && (xml.rfind(".gltf") == xml.size() - 5 || xml.rfind(".glb") == xml.size() - 4)) ? loadGltfScene(xml) : loadMitsubaScene(xml))
                      : builtin == "gallery" ? makeBsdfGallery()
                      : builtin == "field" ? makeInstancedField(1234, 16, 7, 13)
                                           : makeCornellBox();
        for (auto& w : scene.warnings) std::fprintf(stderr, "warning: %s\n", w.c_str());
        if (W == 0 || H == 0) { W = scene.filmWidth ? scene.filmWidth : 512; H = scene.filmHeight ? scene.filmHeight : 512; }
        PathTracer tracer(device);
        tracer.setup(W, H);
        if (maxDepth >= 0 || filter) {
            gsb_params p;
            gsb_default_params(&p);
            if (maxDepth >= 0) p.max_depth = maxDepth;
            p.pixel_filter = filter;
            tracer.setParams(p);
        }
        tracer.render(scene, spp);
        tracer.synchronize();
        gsb_counters c;
        gsb_get_counters(tracer.context(), &c);
        double ms[5];
        gsb_last_render_ms(tracer.context(), ms);
        std::printf("%ux%u, %u spp: %.1f ms, %.1f Mpaths/s, %.1f Mrays/s\n", W, H, spp, ms[0], c.paths / ms[0] * 1e-3,
                    (c.extension_rays + c.shadow_rays) / ms[0] * 1e-3);
        if (!hdrOut.empty()) { std::vector<float> img; tracer.readHDR(img); writePFM(hdrOut, img.data(), W, H); }
        if (!ldrOut.empty()) { std::vector<uint8_t> img; tracer.readLDR(img); writePNG(ldrOut, img.data(), W, H); }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "error: %s\n", e.what());
        return 1;
    }
    return 0;
}
