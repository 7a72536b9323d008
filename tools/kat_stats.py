import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from conftest import golden, materials_from_bytes, probe_from_golden
from test_gpu_parity import ulp_diff, make_renderer
from optixpathtracer_b200 import SampleRenderer, scenes, preset, preset_camera
g=golden("bsdf"); mats=materials_from_bytes(g["materials"])
r=SampleRenderer(scenes.cornell())
out=r.debugBsdf(mats,g["etaI"],g["etaO"],g["N"],g["view"],g["L"],g["albedo"],g["seeds"]); r.close()
ok=g["samplePdf"]>0
print("basis equal", np.array_equal(out["basisU"].view(np.uint32),g["basisU"].view(np.uint32)), np.array_equal(out["basisV"].view(np.uint32),g["basisV"].view(np.uint32)))
print("type eq", np.array_equal(out["sampleType"],g["sampleType"]), "rng eq", np.array_equal(out["sampleRng"],g["sampleRng"]), "pdf>0 eq", np.array_equal(out["samplePdf"]>0,ok))
spec=ok&(g["sampleType"]==2); print("spec", spec.sum(), np.array_equal(out["sampleL"][spec].view(np.uint32),g["sampleL"][spec].view(np.uint32)), np.array_equal(out["samplePdf"][spec].view(np.uint32),g["samplePdf"][spec].view(np.uint32)))
scale=np.abs(g["sampleL"][ok]).max(axis=1,keepdims=True)
derr=np.abs(out["sampleL"][ok]-g["sampleL"][ok])/(scale*2.0**-23)
print("dir err ulp max", derr.max(), "p99", np.percentile(derr,99), "frac bit-equal rows", (derr.max(axis=1)==0).mean())
u=ulp_diff(out["samplePdf"][ok],g["samplePdf"][ok]); print("samplePdf ulp max", u.max(), np.percentile(u,99), (u==0).mean())
rel=np.abs(out["samplePdf"][ok]-g["samplePdf"][ok])/np.abs(g["samplePdf"][ok]); print("samplePdf rel max", rel.max())
u=ulp_diff(out["evalPdf"],g["evalPdf"]); print("evalPdf ulp max", u.max(), np.percentile(u,99), (u==0).mean())
fs=np.maximum(np.abs(g["evalF"]).max(axis=1,keepdims=True),1e-30); fe=np.abs(out["evalF"]-g["evalF"])/(fs*2.0**-23)
fin=np.isfinite(g["evalF"]).all(axis=1)
print("evalF ulp max", np.nanmax(fe[fin]), np.percentile(fe[fin],99), (fe[fin].max(axis=1)==0).mean(), "finite eq", np.array_equal(np.isfinite(out["evalF"]),np.isfinite(g["evalF"])))
g=golden("probe"); probe=probe_from_golden(g)
for ug in (True,False):
    r=make_renderer(scenes.cornell(),probe,(16,16),preset_camera("cornell"),preset("sv4_uniform"))
    o=r.debugProbe(g["seeds"],g["dirs"],use_guide=ug); r.close()
    print("probe guide",ug,"color eq", np.array_equal(o["sampleColor"].view(np.uint32),g["sampleColor"].view(np.uint32)), "evalColor eq", np.array_equal(o["evalColor"].view(np.uint32),g["evalColor"].view(np.uint32)),
      "dir err ulp", (np.abs(o["sampleDir"]-g["sampleDir"])/2.0**-23).max(), "pdf ulp", ulp_diff(o["samplePdf"],g["samplePdf"]).max(), "uv abs/2^-24", (np.abs(o["uv"]-g["uv"])/2.0**-24).max(), "evalPdf ulp", ulp_diff(o["evalPdf"],g["evalPdf"]).max())
