"""TEST INFRASTRUCTURE ONLY -- never imported by the product (egogaussian_b200/).

CPU restatement of GaussianModel.save_ply / load_ply (scene/gaussian_model.py:340-481) in numpy.

PARITY UNPINNED: the reference writes and reads the file through the third-party package `plyfile`
(imported at gaussian_model.py:19; unpinned in the reference's environment.yml), which is absent from
/root/reference and from this image, and the reference ships no .ply fixture. What is restated here is
the reference's own record layout (property order :340-359, channel-major SH :381-382, float storage of
the integer columns :390-394, the defaults of absent columns :444-461) on top of plyfile's published
output for a structured array of 'f4' fields: the ASCII header "ply / format binary_little_endian 1.0 /
element vertex N / property float <name> ... / end_header" followed by the packed records.
"""
import numpy as np


def attribute_names(n_dc, n_rest, extras=True):   # construct_list_of_attributes, :340-359
    l = ['x', 'y', 'z', 'nx', 'ny', 'nz']
    l += ['f_dc_%d' % i for i in range(n_dc)] + ['f_rest_%d' % i for i in range(n_rest)]
    l += ['opacity'] + ['scale_%d' % i for i in range(3)] + ['rot_%d' % i for i in range(4)]
    return l + (['label', 'generation', 'is_object'] if extras else [])


def save_ply_bytes(xyz, features_dc, features_rest, opacity, scaling, rotation, label=None, generation=None,
                   is_object=None):
    """The bytes save_ply writes (:375-397); features_* are [P,K,3] like the model's tensors."""
    P = xyz.shape[0]
    f_dc = np.ascontiguousarray(features_dc.transpose(0, 2, 1).reshape(P, 3 * features_dc.shape[1]))   # transpose(1, 2).flatten(1)
    f_rest = np.ascontiguousarray(features_rest.transpose(0, 2, 1).reshape(P, 3 * features_rest.shape[1]))
    cols = [xyz, np.zeros_like(xyz), f_dc, f_rest, opacity, scaling, rotation]
    extras = label is not None
    if extras:
        cols += [label, generation, is_object]
    attributes = np.concatenate([np.asarray(c, dtype=np.float64) if c.dtype.kind in "iu" else c
                                 for c in cols], axis=1)
    names = attribute_names(f_dc.shape[1], f_rest.shape[1], extras)
    elements = np.empty(P, dtype=[(n, '<f4') for n in names])
    for k, n in enumerate(names):
        elements[n] = attributes[:, k]
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % P
    header += "".join("property float %s\n" % n for n in names) + "end_header\n"
    return header.encode("ascii") + elements.tobytes()


def load_ply_bytes(blob, max_sh_degree, is_object=False, force_bg=False):
    """load_ply (:399-481) from the bytes of a file: dict of numpy arrays in the model's shapes."""
    end = blob.index(b"end_header\n") + len(b"end_header\n")
    lines = blob[:end].decode("ascii").split("\n")
    P = int([l for l in lines if l.startswith("element vertex")][0].split()[2])
    names = [l.split()[2] for l in lines if l.startswith("property")]
    data = np.frombuffer(blob[end:], dtype=[(n, '<f4') for n in names], count=P)
    xyz = np.stack([data["x"], data["y"], data["z"]], axis=1)
    opacity = data["opacity"][..., None]
    features_dc = np.stack([data["f_dc_0"], data["f_dc_1"], data["f_dc_2"]], axis=1)[:, :, None]   # [P,3,1]
    extra = sorted([n for n in names if n.startswith("f_rest_")], key=lambda x: int(x.split('_')[-1]))
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    features_extra = np.zeros((P, len(extra)))
    for idx, n in enumerate(extra):
        features_extra[:, idx] = data[n]
    features_extra = features_extra.reshape(P, 3, (max_sh_degree + 1) ** 2 - 1)
    scales = np.stack([data[n] for n in sorted([n for n in names if n.startswith("scale_")],
                                               key=lambda x: int(x.split('_')[-1]))], axis=1)
    rots = np.stack([data[n] for n in sorted([n for n in names if n.startswith("rot")],
                                             key=lambda x: int(x.split('_')[-1]))], axis=1)
    label = data["label"][..., None] if "label" in names else np.full((P, 1), 0.01)
    generation = data["generation"].astype(int)[..., None] if "generation" in names else np.zeros((P, 1), int)
    if "is_object" in names:
        obj = data["is_object"].astype(int)[..., None]
    else:
        obj = np.ones((P, 1), int) if is_object else np.zeros((P, 1), int)
    if force_bg:
        obj = np.zeros((P, 1), int)
    return dict(xyz=xyz.astype(np.float32), features_dc=features_dc.transpose(0, 2, 1).astype(np.float32),
                features_rest=features_extra.transpose(0, 2, 1).astype(np.float32), opacity=opacity.astype(np.float32),
                scaling=scales.astype(np.float32), rotation=rots.astype(np.float32), label=label.astype(np.float32),
                generation=generation.astype(np.int32), is_object=obj.astype(np.int32))
