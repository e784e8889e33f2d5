import sys, torch, json
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from egogaussian_b200 import cabi
from helpers import baseline_case
kw, ups = baseline_case(1_000_000, 3, 1280, 720, 3001, "scale_rot", "sh", "cuda")
bkw = {k: v for k, v in kw.items() if k not in ("opacities", "image_height", "image_width")}
fwd = cabi.forward(**kw)
hint = int(fwd.num_rendered * 1.25)
for _ in range(5):
    fwd = cabi.forward(num_rendered_hint=hint, **kw); g = cabi.backward(fwd, **bkw, **ups)
torch.cuda.synchronize()
cabi.profile_enable(True); cabi.profile_read()
N = 30
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
e0.record()
for _ in range(N):
    fwd = cabi.forward(num_rendered_hint=hint, **kw); g = cabi.backward(fwd, **bkw, **ups)
e1.record(); torch.cuda.synchronize()
prof = cabi.profile_read(); cabi.profile_enable(False)
st = {k: round(v[0] / N, 4) for k, v in prof.items() if v[1]}
print(json.dumps(st), "sum", round(sum(st.values()), 4), "step", round(e0.elapsed_time(e1) / N, 4))

# the same loop without the stage timers: what their event records cost
e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
for _ in range(3):
    fwd = cabi.forward(num_rendered_hint=hint, **kw); g = cabi.backward(fwd, **bkw, **ups)
e0.record()
for _ in range(N):
    fwd = cabi.forward(num_rendered_hint=hint, **kw); g = cabi.backward(fwd, **bkw, **ups)
e1.record(); torch.cuda.synchronize()
print("step without stage timers", round(e0.elapsed_time(e1) / N, 4))
