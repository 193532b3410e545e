import sys, struct
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import numpy as np
import gpuspectral_b200 as g
from oracle import oracle as orc
from conftest import host_scene, random_rays
ctx = g.Context(0)
for btype, rec in [(6, struct.pack("5f", 0.5, 0.3, 0.2, 0.04, 0.2)), (7, struct.pack("7fi", 1.0, 0.58, 0.13, 1.3, 1.0, 0.017, 0.0707, 0))]:
    ctx.set_bsdfs(btype, rec, 1)
    o = orc.OracleScene(); o.set_bsdfs(btype, rec, 1)
    n = 20000
    rng = np.random.default_rng(100); wo = rng.normal(size=(n,3)); wo /= np.linalg.norm(wo,axis=1,keepdims=True); wo = wo.astype(np.float32)
    seed = np.random.default_rng(0).integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    h = btype << 16
    wi_g, f_g, pdf_g, d_g, s_g = ctx.test_bsdf_sample(h, wo, seed)
    wi_o, f_o, pdf_o, d_o, s_o = o.bsdf_sample(h, wo, seed)
    def rel(a,b):
        a=a.astype(np.float64); b=b.astype(np.float64)
        return np.abs(a-b)/np.maximum(np.maximum(np.abs(a),np.abs(b)),1e-30)
    rw = np.abs(wi_g.astype(np.float64)-wi_o).max(axis=1); rf = rel(f_g,f_o).max(axis=1); rp = rel(pdf_g,pdf_o)
    bad = np.where((rw > 2e-6) | (rf > 2e-4) | (rp > 2e-4))[0]
    print("type", btype, "bad", len(bad))
    for i in bad[:12]:
        print(i, "wo", wo[i], "wi", wi_g[i], wi_o[i], "f", f_g[i], f_o[i], "pdf", pdf_g[i], pdf_o[i])
# f64 test
s = host_scene("cornell"); ctx.upload(s); o = orc.OracleScene.from_host_scene(s)
n=60000
rays = random_rays(s, n, seed=11)[n//3:]
got = ctx.trace_batch(rays); ref = o.trace_batch(rays, mode=2); ref0 = o.trace_batch(rays, mode=0)
print("hits gpu", (got["inst"]!=g.NO_HIT).sum(), "f64", (ref["inst"]!=g.NO_HIT).sum(), "f32 brute", (ref0["inst"]!=g.NO_HIT).sum())
d = np.where((got["inst"]!=g.NO_HIT) != (ref["inst"]!=g.NO_HIT))[0]
print(len(d))
for i in d[:10]:
    print(rays[i], got[i], ref[i])
