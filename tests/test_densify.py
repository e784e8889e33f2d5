"""Densification: clone / split append and the prune of the split originals.

The golden vectors tests/golden/densify_*.npz come from the reference's OWN GaussianModel code with a real
torch.optim.Adam holding state (tests/golden/make_densify_golden.py). On the CPU they pin the torch
restatement oracle/densify_ref.py; on the GPU the device path (egs_densify_select / egs_compact_plan /
egs_append_rows / egs_split_samples / egs_compact_rows behind egogaussian_b200/densify.py) must give the
same tensors: copied rows, Adam moments, generation and object ids bit for bit, the sampled positions and
shrunk scales of a split within 1e-6 relative (expf / logf and a 3x3 product in another order)."""
from pathlib import Path

import numpy as np
import pytest
import torch

from golden.make_densify_golden import DENSIFY_CASES, EXTENT, MAX_GRAD, NAMES, PERCENT_DENSE, densify_inputs

GOLDEN = Path(__file__).resolve().parent / "golden"


def _state(x, dev):
    params = {n: x[n].to(dev) for n in NAMES}
    moments = {n: (x["m_" + n].to(dev), x["v_" + n].to(dev)) for n in NAMES}
    extras = {"generation": x["generation"].to(dev), "is_object": x["is_object"].to(dev)}
    return params, moments, extras


def _check(tag, gold, params, moments, extras, exact=True):
    for n in NAMES:
        want = torch.from_numpy(gold[f"{tag}_{n}"])
        got = params[n].cpu()
        assert got.shape == want.shape, (tag, n, got.shape, want.shape)
        if exact or n not in ("xyz", "scaling"):
            assert torch.equal(got, want), (tag, n)
        else:
            assert torch.allclose(got, want, rtol=1e-6, atol=1e-6), (tag, n, float((got - want).abs().max()))
        assert torch.equal(moments[n][0].cpu(), torch.from_numpy(gold[f"{tag}_m_{n}"])), (tag, "m", n)
        assert torch.equal(moments[n][1].cpu(), torch.from_numpy(gold[f"{tag}_v_{n}"])), (tag, "v", n)
    assert torch.equal(extras["generation"].cpu(), torch.from_numpy(gold[f"{tag}_generation"])), tag
    assert torch.equal(extras["is_object"].cpu(), torch.from_numpy(gold[f"{tag}_is_object"])), tag


def _run(mod, case, dev, gold):
    x = densify_inputs(case)
    params, moments, extras = _state(x, dev)
    ga, dn = x["grad_accum"].to(dev), x["denom"].to(dev)
    p, m, e, n_clone = mod.densify_and_clone(params, moments, extras, ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE,
                                             curr_gen=case["curr_gen"], which_object=case["which"])
    assert n_clone == gold["clone_xyz"].shape[0] - case["P"] and n_clone > 0
    _check("clone", gold, p, m, e)
    noise = torch.from_numpy(gold["noise"]).to(dev)
    p, m, e, n_split = mod.densify_and_split(p, m, e, ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE, noise=noise,
                                             curr_gen=case["curr_gen"], split_prev_gen=case["split_prev_gen"],
                                             which_object=case["which"])
    assert 2 * n_split == noise.shape[0] and n_split > 0
    _check("split", gold, p, m, e, exact=False)


@pytest.mark.parametrize("case", DENSIFY_CASES, ids=[c["name"] for c in DENSIFY_CASES])
def test_restatement_against_reference_golden(case):
    from oracle import densify_ref
    _run(densify_ref, case, torch.device("cpu"), np.load(GOLDEN / f"{case['name']}.npz"))


@pytest.mark.gpu
@pytest.mark.parametrize("case", DENSIFY_CASES, ids=[c["name"] for c in DENSIFY_CASES])
def test_device_path_against_reference_golden(case, built, cuda_device):
    from egogaussian_b200 import densify
    _run(densify, case, cuda_device, np.load(GOLDEN / f"{case['name']}.npz"))


@pytest.mark.gpu
def test_densify_at_scale_against_restatement(built, cuda_device):
    """300k Gaussians, SH degree 3: the device path against the restatement run on the same GPU, and an
    empty selection (nothing to clone) handled without a kernel fault."""
    from egogaussian_b200 import densify
    from oracle import densify_ref
    dev = cuda_device
    case = dict(name="densify_big", P=300_000, M=16, curr_gen=5, which=None, split_prev_gen=True)
    x = densify_inputs(case)
    params, moments, extras = _state(x, dev)
    ga, dn = x["grad_accum"].to(dev), x["denom"].to(dev)
    a = densify.densify_and_clone(params, moments, extras, ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE, curr_gen=5)
    b = densify_ref.densify_and_clone(params, moments, extras, ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE, curr_gen=5)
    assert a[3] == b[3] > 1000
    for n in NAMES:
        assert torch.equal(a[0][n], b[0][n]) and torch.equal(a[1][n][0], b[1][n][0])
    # the number of split rows is known only after the selection: take it from the device path's mask
    from egogaussian_b200 import cabi
    sel = cabi.densify_select(grad_accum=ga.reshape(-1), denom=dn.reshape(-1), scaling_raw=a[0]["scaling"],
                              grad_threshold=MAX_GRAD, size_threshold=PERCENT_DENSE * EXTENT, split=True)
    n_sel = int(sel.sum())
    noise = torch.randn(2 * n_sel, 3, device=dev, generator=torch.Generator(device=dev).manual_seed(1))
    c = densify.densify_and_split(*a[:3], ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE, noise=noise, curr_gen=5)
    d = densify_ref.densify_and_split(*b[:3], ga, dn, MAX_GRAD, EXTENT, PERCENT_DENSE, noise=noise, curr_gen=5)
    assert c[3] == d[3] == n_sel
    for n in NAMES:
        if n in ("xyz", "scaling"):
            assert torch.allclose(c[0][n], d[0][n], rtol=1e-6, atol=1e-6), n
        else:
            assert torch.equal(c[0][n], d[0][n]), n
        assert torch.equal(c[1][n][1], d[1][n][1]), n
    assert torch.equal(c[2]["generation"], d[2]["generation"]) and torch.equal(c[2]["is_object"], d[2]["is_object"])
    # nothing selected: the tensors come back unchanged
    e = densify.densify_and_clone(params, moments, extras, ga, dn, 1e9, EXTENT, PERCENT_DENSE)
    assert e[3] == 0 and torch.equal(e[0]["xyz"], params["xyz"])
