// bsdf.cuh — the reference's BSDF library and light sampling as device functions.
//
// One function per GLSL function of assets/shaders/rayhit.rchit (line ranges beside each).
// All arithmetic is float32 in the shader's operation order; RNG draws are made in the
// order GLSL evaluates them (constructor arguments left to right).  Quirks are kept on
// purpose: Beckmann-sampled / GGX-evaluated rough lobes, sqrt(2)-scaled alpha arriving from
// the loader, the 0.01 clamp that only roughPlasticBSDFEval has, delta lobes returning pdf 1.
#pragma once

#include "common.cuh"

namespace gsb {

struct BSDFOutput {
    float3 bsdf;
    float pdf;
    bool isDelta;
};

struct Onb { float3 tangent, binormal, normal; };

// pt_common.glsl:128-151
GSB_D Onb onbCreate(float3 n) {
    Onb onb;
    onb.normal = normalize(n);
    if (fabsf(onb.normal.x) > fabsf(onb.normal.z)) onb.binormal = f3(-onb.normal.y, onb.normal.x, 0.0f);
    else onb.binormal = f3(0.0f, -onb.normal.z, onb.normal.y);
    onb.binormal = normalize(onb.binormal);
    onb.tangent = cross(onb.binormal, onb.normal);
    return onb;
}
GSB_D float3 onbUntransform(const Onb& o, float3 v) { return o.tangent * v.x + o.binormal * v.y + o.normal * v.z; }
GSB_D float3 onbTransform(const Onb& o, float3 v) { return f3(dot(v, o.tangent), dot(v, o.binormal), dot(v, o.normal)); }

// rayhit.rchit:89-105
GSB_D float2 sampleConcentric(uint32_t& rng) {
    const float s0 = randUniform(rng);
    const float s1 = randUniform(rng);
    const float ux = 2.0f * s0 - 1.0f, uy = 2.0f * s1 - 1.0f;
    if (ux == 0.0f && uy == 0.0f) return make_float2(0.0f, 0.0f);
    float r, th;
    if (fabsf(ux) > fabsf(uy)) {
        r = ux;
        th = (GSB_PI / 4.0f) * (uy / ux);
    } else {
        r = uy;
        th = GSB_PI / 2.0f - (GSB_PI / 4.0f) * (ux / uy);
    }
    return make_float2(r * cosf(th), r * sinf(th));
}

// rayhit.rchit:107-115
GSB_D float3 randCosineHemisphere(uint32_t& rng) {
    const float2 u = sampleConcentric(rng);
    const float z = sqrtf(fmaxf(0.0f, 1.0f - u.x * u.x - u.y * u.y));
    return f3(u.x, u.y, z);
}
GSB_D float cosineHemispherePdf(float3 w) { return fmaxf(fabsf(w.z) / GSB_PI, 0.000001f); }

// rayhit.rchit:155-166
GSB_D float3 sampleHalf(uint32_t& rng, float alpha) {
    const float ux = randUniform(rng);
    const float uy = randUniform(rng);
    const float phi = 2.0f * GSB_PI * ux;
    float logSample = logf(1.0f - uy);
    if (isinf(logSample)) logSample = 0.0f;
    const float tan2 = -alpha * alpha * logSample;
    const float cost = 1.0f / sqrtf(1.0f + tan2);
    const float sint = sqrtf(fmaxf(0.0f, 1.0f - cost * cost));
    return f3(cosf(phi) * sint, sinf(phi) * sint, cost);
}

// rayhit.rchit:177-204
GSB_D float beckmannD(float3 wh, float alpha) {
    const float cos2 = wh.z * wh.z;
    const float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    const float a = expf(-tan2 / (alpha * alpha));
    const float b = GSB_PI * alpha * alpha * cos2 * cos2;
    return a / b;
}
GSB_D float ggxD(float3 wh, float alpha) {
    const float cos2 = wh.z * wh.z;
    const float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    if (isinf(tan2)) return 0.0f;
    const float b = (1.0f + tan2 / (alpha * alpha));
    const float a = GSB_PI * alpha * alpha * cos2 * cos2 * b * b;
    return 1.0f / a;
}
GSB_D float ggxLambda(float3 wh, float alpha) {
    const float cos2 = wh.z * wh.z;
    const float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    if (isinf(tan2)) return 0.0f;
    const float a = -1.0f + sqrtf(1.0f + alpha * alpha * tan2);
    return 0.5f * a;
}
GSB_D float ggxMask(float3 wo, float3 wi, float alpha) { return 1.0f / (1.0f + ggxLambda(wo, alpha) + ggxLambda(wi, alpha)); }

// rayhit.rchit:206-210 (nf = ng = 1 at both call sites)
GSB_D float powerHeuristic(float fPdf, float gPdf) {
    const float f = 1.0f * fPdf;
    const float g = 1.0f * gPdf;
    return (f * f) / (f * f + g * g);
}

// rayhit.rchit:218-247
GSB_D float fresnel4(float no, float cosTho, float nt, float cosTht) {
    const float a = nt * cosTho - no * cosTht;
    const float ad = nt * cosTho + no * cosTht;
    const float b = no * cosTho - nt * cosTht;
    const float bd = no * cosTho + nt * cosTht;
    const float A = a * a / (ad * ad);
    const float B = b * b / (bd * bd);
    return 0.5f * (A + B);
}
GSB_D float fresnelW(float3 wo, float no, float nt) {
    const float sinTho = sqrtf(fmaxf(wo.x * wo.x + wo.y * wo.y, 0.0f));
    const float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return 1.0f;
    const float cosTht = sqrtf(sqrtTerm);
    const float cosTho = fabsf(wo.z);
    return fresnel4(no, cosTho, nt, cosTht);
}
GSB_D float fresnelC(float cosTho, float no, float nt) {
    const float sinTho = sqrtf(fmaxf(1.0f - cosTho * cosTho, 0.0f));
    const float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return 1.0f;
    const float cosTht = sqrtf(sqrtTerm);
    return fresnel4(no, cosTho, nt, cosTht);
}

// rayhit.rchit:269-288
GSB_D float3 fresnelDielectricConductor(float3 Eta, float3 Etak, float CosTheta) {
    const float CosTheta2 = CosTheta * CosTheta;
    const float SinTheta2 = 1.0f - CosTheta2;
    const float3 Eta2 = Eta * Eta;
    const float3 Etak2 = Etak * Etak;
    const float3 t0 = Eta2 - Etak2 - SinTheta2;
    const float3 a2plusb2 = vsqrt(t0 * t0 + 4.0f * Eta2 * Etak2);
    const float3 t1 = a2plusb2 + CosTheta2;
    const float3 a = vsqrt(0.5f * (a2plusb2 + t0));
    const float3 t2 = 2.0f * a * CosTheta;
    const float3 Rs = (t1 - t2) / (t1 + t2);
    const float3 t3 = CosTheta2 * a2plusb2 + SinTheta2 * SinTheta2;
    const float3 t4 = t2 * SinTheta2;
    const float3 Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}

// rayhit.rchit:290-299
GSB_D bool refractRay(float3 wo, float3 n, float no, float nt, float3& wt) {
    const float sinTho = sqrtf(fmaxf(wo.x * wo.x + wo.y * wo.y, 0.0f));
    const float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return false;
    const float cosTht = sqrtf(sqrtTerm);
    wt = (no / nt) * (-wo) + ((no / nt) * dot(wo, n) - cosTht) * n;
    return true;
}

// rayhit.rchit:301-330
GSB_D float coupledDiffuseTerm(float R0, float cosTho, float cosThi) {
    const float k = 21.0f / (20.0f * GSB_PI * (1.0f - R0));
    const float a = 1.0f - cosTho;
    const float b = 1.0f - cosThi;
    const float a5 = a * a * a * a * a;
    const float b5 = b * b * b * b * b;
    return k * (1.0f - a5) * (1.0f - b5);
}
GSB_D float fresnelBlendDiffuseTerm(float R0, float cosTho, float cosThi) {
    const float k = 28.0f / (23.0f * GSB_PI);
    const float a = 1.0f - 0.5f * cosTho;
    const float b = 1.0f - 0.5f * cosThi;
    const float a5 = a * a * a * a * a;
    const float b5 = b * b * b * b * b;
    return k * (1.0f - R0) * (1.0f - a5) * (1.0f - b5);
}
GSB_D float internalScatterEscapeFraction(float R0, float no, float nt) {
    const float Re = (GSB_PI * 20.0f * R0 + 1.0f) / 21.0f;
    const float eta = no / nt;
    return 1.0f - eta * eta * (1.0f - Re);
}
GSB_D float schlickFresnel(float R0, float cosTho) {
    const float a = 1.0f - cosTho;
    const float a5 = a * a * a * a * a;
    return R0 + a5 * (1.0f - R0);
}

GSB_D float3 ld3(const float* p) { return f3(p[0], p[1], p[2]); }
GSB_D float3 mirrorZ(float3 wo) { return f3(-wo.x, -wo.y, wo.z); }

// ---- the eight models: rayhit.rchit:341-617 ----------------------------------------------------
// Diffuse :341-358
GSB_D void diffuseSample(const gsb_diffuse_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    wi = randCosineHemisphere(rng);
    res.bsdf = ld3(b.reflectance) / GSB_PI;
    res.pdf = cosineHemispherePdf(wi);
    res.isDelta = false;
}
GSB_D void diffuseEval(const gsb_diffuse_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    res.bsdf = ld3(b.reflectance) / GSB_PI;
    res.pdf = cosineHemispherePdf(wi);
    res.isDelta = false;
}

// SmoothDielectric :362-404
GSB_D void smoothDielectricSample(const gsb_smooth_dielectric_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const bool entering = wo.z > 0.0f;
    const float no = entering ? b.iorOut : b.iorIn;
    const float nt = entering ? b.iorIn : b.iorOut;
    const float cosTho = wo.z;
    float3 wt;
    const float3 up = f3(0.0f, 0.0f, 1.0f);
    if (!refractRay(wo, faceforward(up, -wo, up), no, nt, wt)) {
        wi = mirrorZ(wo);
        res.bsdf = 1.0f * f3(1.0f / fabsf(cosTho));
        res.isDelta = true;
        res.pdf = 1.0f;
        return;
    }
    const float Fr = fresnel4(no, fabsf(cosTho), nt, fabsf(wt.z));
    const float u = randUniform(rng);
    if (u < Fr) {
        wi = mirrorZ(wo);
        res.bsdf = Fr * f3(1.0f / fabsf(cosTho));
        res.isDelta = true;
        res.pdf = Fr;
    } else {
        wi = wt;
        res.bsdf = f3(((no * no) / (nt * nt)) * (1.0f - Fr) / fabsf(wt.z));
        res.isDelta = true;
        res.pdf = 1.0f - Fr;
    }
}
GSB_D void deltaEval(BSDFOutput& res) {  // :400-404, :420-426
    res.bsdf = f3(0.0f);
    res.pdf = 1.0f;
    res.isDelta = true;
}

// SmoothConductor :406-418
GSB_D void smoothConductorSample(const gsb_smooth_conductor_bsdf& b, float3 wo, float3& wi, BSDFOutput& res) {
    const float no = b.iorOut;
    const float nt = b.iorIn;
    const float Fr = nt == 0.0f ? 1.0f : fresnelW(wo, no, nt);
    wi = mirrorZ(wo);
    res.bsdf = Fr * f3(1.0f / fabsf(wo.z));
    res.isDelta = true;
    res.pdf = 1.0f;
}

// SmoothFloor :428-458
GSB_D void smoothFloorSample(const gsb_smooth_floor_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const float Fr = schlickFresnel(b.R0, fabsf(wo.z));
    const float u = randUniform(rng);
    if (u < Fr) {
        wi = mirrorZ(wo);
        res.bsdf = ld3(b.diffuse) * coupledDiffuseTerm(b.R0, fabsf(wo.z), fabsf(wi.z)) + Fr * f3(1.0f / fabsf(wo.z));
        res.pdf = Fr;
        res.isDelta = true;
    } else {
        wi = randCosineHemisphere(rng);
        res.bsdf = ld3(b.diffuse) * coupledDiffuseTerm(b.R0, fabsf(wo.z), fabsf(wi.z));
        res.pdf = (1.0f - Fr) * cosineHemispherePdf(wi);
        res.isDelta = false;
    }
}
GSB_D void smoothFloorEval(const gsb_smooth_floor_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const float Fr = schlickFresnel(b.R0, fabsf(wo.z));
    res.bsdf = ld3(b.diffuse) * coupledDiffuseTerm(b.R0, fabsf(wo.z), fabsf(wi.z));
    res.pdf = (1.0f - Fr) * cosineHemispherePdf(wi);
    res.isDelta = false;
}

// SmoothPlastic :461-506
GSB_D void smoothPlasticSample(const gsb_smooth_plastic_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const float u = randUniform(rng);
    const float no = b.iorOut;
    const float nt = b.iorIn;
    const float Fri = fresnelC(fabsf(wo.z), no, nt);
    if (u < Fri) {
        wi = mirrorZ(wo);
        res.bsdf = Fri * f3(1.0f / fabsf(wo.z));
        res.pdf = Fri;
        res.isDelta = true;
    } else {
        wi = randCosineHemisphere(rng);
        const float Fro = fresnelC(fabsf(wi.z), no, nt);
        const float Ri = internalScatterEscapeFraction(b.R0, no, nt);
        const float eta = no / nt;
        const float3 diffuse = ld3(b.diffuse);
        res.bsdf = diffuse * eta * eta * (1.0f - Fri) * (1.0f - Fro) / (GSB_PI * (1.0f - diffuse * Ri));
        res.pdf = (1.0f - Fri) * cosineHemispherePdf(wi);
        res.isDelta = false;
    }
}
GSB_D void smoothPlasticEval(const gsb_smooth_plastic_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const float no = b.iorOut;
    const float nt = b.iorIn;
    const float Fri = fresnelC(fabsf(wo.z), no, nt);
    const float Fro = fresnelC(fabsf(wi.z), no, nt);
    const float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    const float eta = no / nt;
    const float3 diffuse = ld3(b.diffuse);
    res.bsdf = diffuse * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (GSB_PI * (1.0f - diffuse * Ri));
    res.pdf = (1.0f - Fri) * cosineHemispherePdf(wi);
    res.isDelta = false;
}

// RoughConductor :508-530
GSB_D void roughConductorSample(const gsb_rough_conductor_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const float3 Fr = fresnelDielectricConductor(ld3(b.eta), ld3(b.k), fabsf(wo.z));
    float3 wh = sampleHalf(rng, b.alpha);
    if (wh.z <= 0.0f) wh = wh * -1.0f;
    wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    res.bsdf = ld3(b.reflectance) * Fr * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wi.z) * fabsf(wo.z));
    res.pdf = beckmannD(wh, b.alpha) * fabsf(wh.z) / (4.0f * fabsf(dot(wo, wh)));
    res.isDelta = false;
}
GSB_D void roughConductorEval(const gsb_rough_conductor_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const float3 Fr = fresnelDielectricConductor(ld3(b.eta), ld3(b.k), fabsf(wo.z));
    const float3 wh = normalize(wo + wi);
    res.bsdf = Fr * ld3(b.reflectance) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wi.z) * fabsf(wo.z));
    res.pdf = beckmannD(wh, b.alpha) * fabsf(wh.z) / (4.0f * fabsf(dot(wo, wh)));
    res.isDelta = false;
}

// RoughPlastic :532-582
GSB_D void roughPlasticSample(const gsb_rough_plastic_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const float u = randUniform(rng);
    const float no = b.iorOut;
    const float nt = b.iorIn;
    const float eta = no / nt;
    if (u < 0.5f) {
        float3 wh = sampleHalf(rng, b.alpha);
        if (wh.z <= 0.0f) wh = wh * -1.0f;
        wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    } else {
        wi = randCosineHemisphere(rng);
    }
    const float3 wh = normalize(wi + wo);
    const float Fri = fresnelC(fabsf(dot(wh, wo)), no, nt);
    const float Fro = fresnelC(fabsf(dot(wh, wi)), no, nt);
    const float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    const float3 kD = ld3(b.diffuse);
    const float3 specular = f3(Fri) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wo.z) * fabsf(wi.z));
    const float3 d = kD * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (GSB_PI * (1.0f - kD * Ri));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * fabsf(wh.z) / (4.0f * fabsf(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}
GSB_D void roughPlasticEval(const gsb_rough_plastic_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const float no = b.iorOut;
    const float nt = b.iorIn;
    const float3 wh = normalize(wi + wo);
    const float Fri = fresnelC(fabsf(dot(wh, wo)), no, nt);
    const float Fro = fresnelC(fabsf(dot(wh, wi)), no, nt);
    const float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    const float eta = no / nt;
    const float3 kD = ld3(b.diffuse);
    const float3 specular = f3(Fri) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wo.z) * fabsf(wi.z));
    const float3 d = kD * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (GSB_PI * (1.0f - kD * Ri));
    res.pdf = 0.5f * fmaxf(beckmannD(wh, b.alpha) * fabsf(wh.z), 0.01f) / (4.0f * fabsf(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}

// RoughFloor :583-617
GSB_D void roughFloorSample(const gsb_rough_floor_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const float u = randUniform(rng);
    if (u < 0.5f) {
        float3 wh = sampleHalf(rng, b.alpha);
        if (wh.z <= 0.0f) wh = wh * -1.0f;
        wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    } else {
        wi = randCosineHemisphere(rng);
    }
    const float3 wh = normalize(wi + wo);
    const float Fr = schlickFresnel(b.R0, fabsf(dot(wo, wh)));
    const float3 d = ld3(b.diffuse) * fresnelBlendDiffuseTerm(b.R0, fabsf(wo.z), fabsf(wi.z));
    const float3 specular = f3(Fr) * ggxD(wh, b.alpha) / (4.0f * fabsf(dot(wo, wh)) * fmaxf(fabsf(wo.z), fabsf(wi.z)));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * fabsf(wh.z) / (4.0f * fabsf(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}
GSB_D void roughFloorEval(const gsb_rough_floor_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const float3 wh = normalize(wi + wo);
    const float Fr = schlickFresnel(b.R0, fabsf(dot(wo, wh)));
    const float3 d = ld3(b.diffuse) * fresnelBlendDiffuseTerm(b.R0, fabsf(wo.z), fabsf(wi.z));
    const float ggx = ggxD(wh, b.alpha);
    const float3 specular = f3(Fr) * ggx / (4.0f * fabsf(dot(wo, wh)) * fmaxf(fabsf(wo.z), fabsf(wi.z)));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * fabsf(wh.z) / (4.0f * fabsf(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}

// RoughDielectric — extension, no counterpart in rayhit.rchit (include/gpuspectral_b200.h: GSB_BSDF_ROUGH_DIELECTRIC).
// GGX microfacet reflection + refraction after Walter, Marschner, Li & Torrance, "Microfacet Models for Refraction
// through Rough Surfaces" (EGSR 2007), in this shader's conventions: f excludes the cosine, wo points back along the
// path, light arrives along wi; the transmitted lobe carries the (no/nt)^2 radiance scaling of the smooth dielectric
// (:362-404) and converges to it for alpha -> 0.  Draw order: two uniforms for the half vector, one for the lobe.
GSB_D float3 sampleHalfGGX(uint32_t& rng, float alpha) {
    const float ux = randUniform(rng);
    const float uy = randUniform(rng);
    const float phi = 2.0f * GSB_PI * ux;
    const float tan2 = alpha * alpha * uy / (1.0f - uy);   // uy = 1 -> inf -> grazing half vector, D = 0
    const float cost = 1.0f / sqrtf(1.0f + tan2);
    const float sint = sqrtf(fmaxf(0.0f, 1.0f - cost * cost));
    return f3(cosf(phi) * sint, sinf(phi) * sint, cost);
}
GSB_D void roughDielectricInvalid(float3 wo, float3& wi, BSDFOutput& res) {
    wi = mirrorZ(wo);
    res.bsdf = f3(0.0f);
    res.pdf = 0.0f;
    res.isDelta = false;
}
GSB_D void roughDielectricSample(const gsb_rough_dielectric_bsdf& b, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const bool entering = wo.z > 0.0f;
    const float no = entering ? b.iorOut : b.iorIn;
    const float nt = entering ? b.iorIn : b.iorOut;
    float3 wh = sampleHalfGGX(rng, b.alpha);
    if (!entering) wh = wh * -1.0f;                        // half vector on wo's side
    const float u = randUniform(rng);
    const float cosI = dot(wo, wh);
    const float D = ggxD(wh, b.alpha);
    if (!(cosI > 0.0f) || !(D > 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
    const float Fr = fresnelC(cosI, no, nt);
    res.isDelta = false;
    if (u < Fr) {
        wi = normalize(-wo + 2.0f * cosI * wh);
        if (!(wi.z * wo.z > 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
        res.bsdf = f3(Fr * D * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wi.z) * fabsf(wo.z)));
        res.pdf = Fr * D * fabsf(wh.z) / (4.0f * cosI);
    } else {
        const float eta = no / nt;
        const float cosT2 = 1.0f - eta * eta * (1.0f - cosI * cosI);   // > 0 here: Fr = 1 under total internal reflection
        wi = normalize(-eta * wo + (eta * cosI - sqrtf(fmaxf(cosT2, 0.0f))) * wh);
        const float cosO = dot(wi, wh);
        if (!(wi.z * wo.z < 0.0f) || !(cosO < 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
        const float denom = no * cosI + nt * cosO;
        const float jac = fabsf(cosO) / (denom * denom);
        res.bsdf = f3((1.0f - Fr) * D * ggxMask(wo, wi, b.alpha) * cosI * no * no * jac / (fabsf(wi.z) * fabsf(wo.z)));
        res.pdf = (1.0f - Fr) * D * fabsf(wh.z) * nt * nt * jac;
    }
}
GSB_D void roughDielectricEval(const gsb_rough_dielectric_bsdf& b, float3 wo, float3 wi, BSDFOutput& res) {
    const bool entering = wo.z > 0.0f;
    const float no = entering ? b.iorOut : b.iorIn;
    const float nt = entering ? b.iorIn : b.iorOut;
    res.isDelta = false;
    res.bsdf = f3(0.0f);
    res.pdf = 0.0f;
    const bool reflect = wi.z * wo.z > 0.0f;
    float3 wh = reflect ? normalize(wo + wi) : normalize(no * wo + nt * wi);
    if (dot(wo, wh) < 0.0f) wh = wh * -1.0f;
    const float cosI = dot(wo, wh), cosO = dot(wi, wh);
    const float D = ggxD(wh, b.alpha);
    if (!(cosI > 0.0f) || !(D > 0.0f) || wi.z == 0.0f || wo.z == 0.0f) return;
    if (entering ? wh.z <= 0.0f : wh.z >= 0.0f) return;     // the sampler only produces half vectors on wo's side
    const float Fr = fresnelC(cosI, no, nt);
    if (reflect) {
        res.bsdf = f3(Fr * D * ggxMask(wo, wi, b.alpha) / (4.0f * fabsf(wi.z) * fabsf(wo.z)));
        res.pdf = Fr * D * fabsf(wh.z) / (4.0f * cosI);
    } else {
        if (!(cosO < 0.0f)) return;
        const float denom = no * cosI + nt * cosO;
        if (denom == 0.0f) return;
        const float jac = fabsf(cosO) / (denom * denom);
        res.bsdf = f3((1.0f - Fr) * D * ggxMask(wo, wi, b.alpha) * cosI * no * no * jac / (fabsf(wi.z) * fabsf(wo.z)));
        res.pdf = (1.0f - Fr) * D * fabsf(wh.z) * nt * nt * jac;
    }
}

// :620-627
GSB_D bool isTransmissionBSDF(uint32_t type) { return type == GSB_BSDF_SMOOTH_DIELECTRIC || type == GSB_BSDF_ROUGH_DIELECTRIC; }

// Compile-time dispatch used by the material-sorted shade kernel (TYPE is uniform per block);
// the runtime switch of :630-654 survives only in the diagnostic entry points.
// ---- reflectance textures (extension; gsb_texture in the public header) -------------------------------------------------
// checkerboard: Mitsuba's rule on the transformed coordinate; bitmap: bilinear over texel centres, repeat, v = 1 at row 0
GSB_D float3 textureEval(const TextureDev& t, float u, float v) {
    const float uu = u * t.uscale + t.uoffset, vv = v * t.vscale + t.voffset;
    if (t.type == GSB_TEXTURE_CHECKERBOARD) {
        const int ix = (int)floorf(2.0f * uu), iy = (int)floorf(2.0f * vv);
        return ((ix + iy) & 1) == 0 ? xyz(t.color0) : xyz(t.color1);
    }
    const float w = (float)t.width, h = (float)t.height;
    const float x = (uu - floorf(uu)) * w - 0.5f, y = (1.0f - (vv - floorf(vv))) * h - 0.5f;
    const float fx = floorf(x), fy = floorf(y);
    const float ax = x - fx, ay = y - fy;
    int x0 = (int)fx, y0 = (int)fy;
    const int W = (int)t.width, H = (int)t.height;
    x0 = ((x0 % W) + W) % W; y0 = ((y0 % H) + H) % H;
    const int x1 = (x0 + 1) % W, y1 = (y0 + 1) % H;
    const float3 c00 = xyz(__ldg(t.texels + (size_t)y0 * W + x0)), c10 = xyz(__ldg(t.texels + (size_t)y0 * W + x1));
    const float3 c01 = xyz(__ldg(t.texels + (size_t)y1 * W + x0)), c11 = xyz(__ldg(t.texels + (size_t)y1 * W + x1));
    return (c00 * (1.0f - ax) + c10 * ax) * (1.0f - ay) + (c01 * (1.0f - ax) + c11 * ax) * ay;
}

template <int TYPE>
GSB_D void sampleBSDFT(const SceneDev& sc, uint32_t index, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    if (TYPE == GSB_BSDF_DIFFUSE) diffuseSample(((const gsb_diffuse_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_DIELECTRIC) smoothDielectricSample(((const gsb_smooth_dielectric_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_CONDUCTOR) smoothConductorSample(((const gsb_smooth_conductor_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_PLASTIC) smoothPlasticSample(((const gsb_smooth_plastic_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_CONDUCTOR) roughConductorSample(((const gsb_rough_conductor_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_FLOOR) smoothFloorSample(((const gsb_smooth_floor_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_FLOOR) roughFloorSample(((const gsb_rough_floor_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_PLASTIC) roughPlasticSample(((const gsb_rough_plastic_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_DIELECTRIC) roughDielectricSample(((const gsb_rough_dielectric_bsdf*)sc.bsdf[TYPE])[index], rng, wo, wi, res);
}
template <int TYPE>
GSB_D void evalBSDFT(const SceneDev& sc, uint32_t index, float3 wo, float3 wi, BSDFOutput& res) {
    if (TYPE == GSB_BSDF_DIFFUSE) diffuseEval(((const gsb_diffuse_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_DIELECTRIC) deltaEval(res);
    else if (TYPE == GSB_BSDF_SMOOTH_CONDUCTOR) deltaEval(res);
    else if (TYPE == GSB_BSDF_SMOOTH_PLASTIC) smoothPlasticEval(((const gsb_smooth_plastic_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_CONDUCTOR) roughConductorEval(((const gsb_rough_conductor_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_SMOOTH_FLOOR) smoothFloorEval(((const gsb_smooth_floor_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_FLOOR) roughFloorEval(((const gsb_rough_floor_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_PLASTIC) roughPlasticEval(((const gsb_rough_plastic_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
    else if (TYPE == GSB_BSDF_ROUGH_DIELECTRIC) roughDielectricEval(((const gsb_rough_dielectric_bsdf*)sc.bsdf[TYPE])[index], wo, wi, res);
}

// runtime dispatch :630-654 (diagnostics only)
GSB_D void sampleBSDF(const SceneDev& sc, uint32_t handle, uint32_t& rng, float3 wo, float3& wi, BSDFOutput& res) {
    const uint32_t i = handle & 0xffffu;
    switch (handle >> 16) {
        case 0: sampleBSDFT<0>(sc, i, rng, wo, wi, res); break;
        case 1: sampleBSDFT<1>(sc, i, rng, wo, wi, res); break;
        case 2: sampleBSDFT<2>(sc, i, rng, wo, wi, res); break;
        case 3: sampleBSDFT<3>(sc, i, rng, wo, wi, res); break;
        case 4: sampleBSDFT<4>(sc, i, rng, wo, wi, res); break;
        case 5: sampleBSDFT<5>(sc, i, rng, wo, wi, res); break;
        case 6: sampleBSDFT<6>(sc, i, rng, wo, wi, res); break;
        case 7: sampleBSDFT<7>(sc, i, rng, wo, wi, res); break;
        case 8: sampleBSDFT<8>(sc, i, rng, wo, wi, res); break;
    }
}
GSB_D void evalBSDF(const SceneDev& sc, uint32_t handle, float3 wo, float3 wi, BSDFOutput& res) {
    const uint32_t i = handle & 0xffffu;
    switch (handle >> 16) {
        case 0: evalBSDFT<0>(sc, i, wo, wi, res); break;
        case 1: evalBSDFT<1>(sc, i, wo, wi, res); break;
        case 2: evalBSDFT<2>(sc, i, wo, wi, res); break;
        case 3: evalBSDFT<3>(sc, i, wo, wi, res); break;
        case 4: evalBSDFT<4>(sc, i, wo, wi, res); break;
        case 5: evalBSDFT<5>(sc, i, wo, wi, res); break;
        case 6: evalBSDFT<6>(sc, i, wo, wi, res); break;
        case 7: evalBSDFT<7>(sc, i, wo, wi, res); break;
        case 8: evalBSDFT<8>(sc, i, wo, wi, res); break;
    }
}

// ---- light sampling: rayhit.rchit:117-153 ------------------------------------------------------
struct LightOutput {
    float3 position;
    float3 emission;
    float pdf;
};

// sampleLight + sampleTrangleLight.  One raw draw picks the light, two uniform draws the point.
// ---- environment emitter (extension; gsb_environment in the public header) ----------------------------------------------
GSB_D float3 envRotate(const float* R, float3 d) {
    return f3(R[0] * d.x + R[1] * d.y + R[2] * d.z, R[3] * d.x + R[4] * d.y + R[5] * d.z, R[6] * d.x + R[7] * d.y + R[8] * d.z);
}
GSB_D float3 envRotateInv(const float* R, float3 d) {
    return f3(R[0] * d.x + R[3] * d.y + R[6] * d.z, R[1] * d.x + R[4] * d.y + R[7] * d.z, R[2] * d.x + R[5] * d.y + R[8] * d.z);
}
// radiance arriving from world direction d
GSB_D float3 envRadiance(const EnvDev& e, float3 d) {
    const float3 sc = f3(e.scale[0], e.scale[1], e.scale[2]);
    if (e.type == GSB_ENVIRONMENT_CONSTANT) return sc;
    const float3 l = envRotateInv(e.rot, d);
    const float th = acosf(fminf(fmaxf(l.y, -1.0f), 1.0f));
    const float ph = atan2f(l.x, -l.z);
    const float u = ph * (0.5f / GSB_PI) + 0.5f, v = th * (1.0f / GSB_PI);
    const int x = min((int)(u * (float)e.width), (int)e.width - 1), y = min((int)(v * (float)e.height), (int)e.height - 1);
    return xyz(__ldg(e.texels + (size_t)max(y, 0) * e.width + max(x, 0))) * sc;
}
// largest i in [0, n) with cdf[i] <= x
GSB_D int envSearch(const float* cdf, int n, float x) {
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cdf + mid) <= x) lo = mid; else hi = mid;
    }
    return lo;
}
// direction towards the environment for two uniform draws, its radiance and solid-angle density
GSB_D float3 envSample(const EnvDev& e, float e1, float e2, float3& radiance, float& pdf) {
    if (e.type == GSB_ENVIRONMENT_CONSTANT) {
        const float z = 1.0f - 2.0f * e1, r = sqrtf(fmaxf(0.0f, 1.0f - z * z)), ph = 2.0f * GSB_PI * e2;
        radiance = f3(e.scale[0], e.scale[1], e.scale[2]);
        pdf = 1.0f / (4.0f * GSB_PI);
        return f3(r * cosf(ph), z, r * sinf(ph));
    }
    const int W = (int)e.width, H = (int)e.height;
    const float a = fminf(e1, 0.99999994f), b = fminf(e2, 0.99999994f);
    const int y = envSearch(e.rowCdf, H, a);
    const float r0 = __ldg(e.rowCdf + y), r1 = __ldg(e.rowCdf + y + 1);
    const float* cc = e.colCdf + (size_t)y * (W + 1);
    const int x = envSearch(cc, W, b);
    const float c0 = __ldg(cc + x), c1 = __ldg(cc + x + 1);
    const float fv = r1 > r0 ? (a - r0) / (r1 - r0) : 0.5f, fu = c1 > c0 ? (b - c0) / (c1 - c0) : 0.5f;
    const float u = ((float)x + fu) / (float)W, v = ((float)y + fv) / (float)H;
    const float th = v * GSB_PI, ph = (u - 0.5f) * (2.0f * GSB_PI);
    const float st = sinf(th);
    const float3 l = f3(st * sinf(ph), cosf(th), -st * cosf(ph));
    radiance = xyz(__ldg(e.texels + (size_t)y * W + x)) * f3(e.scale[0], e.scale[1], e.scale[2]);
    pdf = st > 0.0f ? (r1 - r0) * (c1 - c0) * (float)W * (float)H / (2.0f * GSB_PI * GSB_PI * st) : 0.0f;
    return envRotate(e.rot, l);
}

// numLights == 0 (a modulo by zero in the shader) is defined as: the three draws are still
// consumed, pdf = 0, no emission, `haveLight` false (no shadow ray is cast).
// EXT = false compiles the reference's light set only (scenes without an environment emitter pay nothing for the extension).
template <bool EXT = true>
GSB_D LightOutput sampleLight(const SceneDev& sc, uint32_t& rng, float3 pos, bool& haveLight) {
    LightOutput res;
    // the environment emitter (extension) counts as one more light behind the triangle lights
    const uint32_t numLights = sc.numLights + ((EXT && sc.env.type) ? 1u : 0u);
    if (numLights == 0) {
        randPcg(rng); randPcg(rng); randPcg(rng);
        haveLight = false;
        res.position = pos + f3(0.0f, 1.0f, 0.0f);
        res.emission = f3(0.0f);
        res.pdf = 0.0f;
        return res;
    }
    haveLight = true;
    const uint32_t lightIdx = randPcg(rng) % numLights;
    if (EXT && lightIdx == sc.numLights) {
        const float e1 = randUniform(rng);
        const float e2 = randUniform(rng);
        const float3 d = envSample(sc.env, e1, e2, res.emission, res.pdf);
        res.position = pos + d * sc.env.distance;
        res.pdf *= 1.0f / (float)numLights;
        return res;
    }
    const float4* lp = reinterpret_cast<const float4*>(sc.lights + lightIdx);
    const float4 p0 = __ldg(lp), p1 = __ldg(lp + 1), p2 = __ldg(lp + 2), rad = __ldg(lp + 3);
    const float e1 = randUniform(rng);
    const float e2 = randUniform(rng);
    const float u = 1.0f - sqrtf(e1);
    const float v = e2 * sqrtf(e1);
    const float w = 1.0f - u - v;
    const float3 v0 = xyz(p0), v1 = xyz(p1), v2 = xyz(p2);
    const float A = 0.5f * fabsf(length(cross(v2 - v0, v1 - v0)));
    const float3 normal = normalize(cross(v1 - v0, v2 - v0));
    const float3 lightPos = u * v0 + v * v1 + w * v2;
    const float ldist = length(lightPos - pos);
    const float3 l = normalize(lightPos - pos);
    res.position = lightPos;
    res.emission = xyz(rad) * (dot(-l, normal) > 0.0f ? 1.0f : 0.0f);
    res.pdf = ldist * ldist / (fabsf(dot(-l, normal)) * A);
    res.pdf *= 1.0f / (float)numLights;
    return res;
}

GSB_D bool isvalidf(float x) { return !isnan(x) && !isinf(x); }
GSB_D bool isvalid3(float3 v) { return isvalidf(v.x) && isvalidf(v.y) && isvalidf(v.z); }

}  // namespace gsb
