"""Scratch: a few forwards of the C3 scene (for ncu captures)."""
import sys, torch
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from egogaussian_b200 import cabi
from helpers import baseline_case
kw, ups = baseline_case(1_000_000, 3, 1280, 720, 3001, "scale_rot", "sh", "cuda")
fwd = cabi.forward(**kw)
hint = int(fwd.num_rendered * 1.25)
for _ in range(6):
    fwd = cabi.forward(num_rendered_hint=hint, **kw)
torch.cuda.synchronize()
