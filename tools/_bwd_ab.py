import sys, torch, json
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from egogaussian_b200 import cabi
from helpers import baseline_case, grad_rel_err
kw, ups = baseline_case(1_000_000, 3, 1280, 720, 3001, "scale_rot", "sh", "cuda")
bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
fwd = cabi.forward(**kw)
res = {}
for ppl in ("2", "4lo", None):
    cabi.debug_option("blend_backward", ppl)
    cabi.profile_enable(True); 
    for _ in range(3): g = cabi.backward(fwd, **bkw, **ups)
    torch.cuda.synchronize(); cabi.profile_read()
    for _ in range(20): g = cabi.backward(fwd, **bkw, **ups)
    torch.cuda.synchronize()
    prof = cabi.profile_read(); cabi.profile_enable(False)
    res[ppl or "4"] = {k: round(v[0] / 20, 4) for k, v in prof.items() if v[1]}
    print(ppl or "4", res[ppl or "4"])
