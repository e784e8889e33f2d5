"""Scratch: stage times with the instance partition as offset table / chained scan."""
import sys, torch, json
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from egogaussian_b200 import cabi
from helpers import baseline_case
for P in (1_000_000, 4_000_000):
    kw, ups = baseline_case(P, 3, 1280, 720, 3001, "scale_rot", "sh", "cuda")
    fwd = cabi.forward(**kw)
    hint = int(fwd.num_rendered * 1.25)
    for mode in (None, "lookback"):
        cabi.debug_option("partition", mode)
        for _ in range(5):
            fwd = cabi.forward(num_rendered_hint=hint, **kw)
        torch.cuda.synchronize()
        cabi.profile_enable(True); cabi.profile_read()
        N = 30
        for _ in range(N):
            fwd = cabi.forward(num_rendered_hint=hint, **kw)
        torch.cuda.synchronize()
        prof = cabi.profile_read(); cabi.profile_enable(False)
        st = {k: round(v[0] / N, 4) for k, v in prof.items() if v[1]}
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record()
        for _ in range(N):
            fwd = cabi.forward(num_rendered_hint=hint, **kw)
        e1.record(); torch.cuda.synchronize()
        print(P, mode or "table", json.dumps(st), "fwd", round(e0.elapsed_time(e1) / N, 4))
    cabi.debug_option("partition", None)
    del kw, ups, fwd
