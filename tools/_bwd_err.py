import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from egogaussian_b200 import cabi
from helpers import baseline_case, grad_rel_err
from oracle import ref_dgr
for P, deg, geom in ((1_000_000, 3, "scale_rot"), (2_000_000, 0, "cov")):
    kw, ups = baseline_case(P, deg, 1280, 720, 3001, geom, "sh", "cuda")
    bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
    fwd = cabi.forward(**kw)
    rf = ref_dgr.forward(**kw)
    rg = ref_dgr.backward(rf, **bkw, **ups)
    rg2 = ref_dgr.backward(rf, **bkw, **ups)
    print(P, "ref self noise", {k: "%.1e" % grad_rel_err(rg2[k], v) for k, v in rg.items() if v is not None})
    for ppl in ("2", "4lo", None):
        cabi.debug_option("blend_backward", ppl)
        g = cabi.backward(fwd, **bkw, **ups)
        print(P, ppl or "4", {k: "%.1e" % grad_rel_err(g[k].reshape(v.shape), v) for k, v in rg.items() if v is not None})
        if ppl is None:
            d = (g["means2D"] - rg["means2D"]).abs()
            i = int(d.max(dim=1).values.argmax())
            gv = fwd.geom_view()
            print("  worst means2D row", i, g["means2D"][i].tolist(), rg["means2D"][i].tolist(), "max", float(rg["means2D"].abs().max()),
                  "conic_o", gv["conic_opacity"][i].tolist(), "radius", int(fwd.radii[i]))
    cabi.debug_option("blend_backward", None)
