"""Seeded synthetic scenes and cameras (SURVEY.md App. C / BASELINE.json configs).

Camera conventions follow the reference: `getWorld2View2` and `getProjectionMatrix`
(utils/graphics_utils.py:38-71) and the transposed, row-vector matrices a `Camera` exposes
(scene/cameras.py:67-70): viewmatrix = W2C^T, projmatrix = viewmatrix @ P^T, campos = inverse
view's translation row. Everything is generated on the CPU from a `torch.Generator` so the same
seed gives the same scene on every machine; callers move tensors to the GPU afterwards.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, Optional

import torch


@dataclass
class Camera:
    width: int
    height: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # [4,4] transposed world->view
    projmatrix: torch.Tensor   # [4,4] transposed full projection
    campos: torch.Tensor       # [3]


@dataclass
class Scene:
    means3D: torch.Tensor            # [P,3]
    scales: torch.Tensor             # [P,3]
    rotations: torch.Tensor          # [P,4] (w,x,y,z), normalised
    opacities: torch.Tensor          # [P,1]
    shs: torch.Tensor                # [P,M,3]
    sh_degree: int
    colors_precomp: torch.Tensor     # [P,3] alternative colour source
    is_object: torch.Tensor          # [P] int32, 10 % ones
    bg: torch.Tensor                 # [3]
    camera: Camera
    extras: Dict[str, torch.Tensor] = field(default_factory=dict)

    @property
    def P(self) -> int:
        return int(self.means3D.shape[0])

    def cov3D(self) -> torch.Tensor:
        """Packed covariance [xx,xy,xz,yy,yz,zz] as scene/gaussian_model.py:29-33 builds it."""
        return build_covariance(self.scales, self.rotations)

    def to(self, device) -> "Scene":
        def mv(t):
            return t.to(device) if isinstance(t, torch.Tensor) else t
        cam = Camera(self.camera.width, self.camera.height, self.camera.tanfovx, self.camera.tanfovy,
                     mv(self.camera.viewmatrix), mv(self.camera.projmatrix), mv(self.camera.campos))
        return Scene(mv(self.means3D), mv(self.scales), mv(self.rotations), mv(self.opacities), mv(self.shs),
                     self.sh_degree, mv(self.colors_precomp), mv(self.is_object), mv(self.bg), cam,
                     {k: mv(v) for k, v in self.extras.items()})


def quat_to_rotmat(q: torch.Tensor) -> torch.Tensor:
    """Rotation matrices from normalised (w,x,y,z) quaternions (utils/general_utils.py:124-149)."""
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.reshape(-1, 3, 3)


def build_covariance(scales: torch.Tensor, rotations: torch.Tensor, modifier: float = 1.0) -> torch.Tensor:
    L = quat_to_rotmat(rotations) * (modifier * scales)[:, None, :]   # R @ diag(s)
    S = L @ L.transpose(1, 2)
    return torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], dim=1).contiguous()


def make_camera(width: int, height: int, fovx_deg: float = 60.0, yaw_deg: float = 0.0,
                x_shift: float = 0.0, znear: float = 0.01, zfar: float = 100.0) -> Camera:
    fx = width / (2.0 * math.tan(math.radians(fovx_deg) / 2))
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(height / (2.0 * fx))
    tanx, tany = math.tan(fovx / 2), math.tan(fovy / 2)
    yaw = math.radians(yaw_deg)
    R = torch.tensor([[math.cos(yaw), 0.0, math.sin(yaw)], [0.0, 1.0, 0.0],
                      [-math.sin(yaw), 0.0, math.cos(yaw)]], dtype=torch.float64)
    T = torch.tensor([x_shift, 0.0, 0.0], dtype=torch.float64)
    w2c = torch.eye(4, dtype=torch.float64)
    w2c[:3, :3] = R.t()          # getWorld2View2: Rt[:3,:3] = R.transpose()
    w2c[:3, 3] = T
    top, right = tany * znear, tanx * znear
    Pm = torch.zeros(4, 4, dtype=torch.float64)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    view = w2c.t().contiguous()
    full = (view.unsqueeze(0).bmm(Pm.t().unsqueeze(0))).squeeze(0)
    campos = view.inverse()[3, :3]
    return Camera(width, height, tanx, tany, view.float().contiguous(), full.float().contiguous(),
                  campos.float().contiguous())


def make_scene(P: int, width: int, height: int, sh_degree: int = 3, seed: int = 0, view: Optional[int] = None,
               bg=(0.0, 0.0, 0.0), max_sh_degree: Optional[int] = None) -> Scene:
    """The App. C recipe. `view` = k selects one of the 8 C4 cameras (yaw (k-3.5)*2 deg, shift)."""
    g = torch.Generator("cpu").manual_seed(int(seed))
    cam0 = make_camera(width, height)
    if view is None:
        cam = cam0
    else:
        cam = make_camera(width, height, yaw_deg=(view - 3.5) * 2.0, x_shift=(view - 3.5) * 0.05)

    def U(n, lo, hi):
        return torch.rand(n, generator=g) * (hi - lo) + lo

    def N(*shape):
        return torch.randn(*shape, generator=g)

    z = U(P, 1.0, 8.0)
    behind = torch.rand(P, generator=g) < 0.02
    z = torch.where(behind, U(P, -1.0, 0.2), z)
    x = U(P, -1.1, 1.1) * z * cam0.tanfovx
    y = U(P, -1.1, 1.1) * z * cam0.tanfovy
    means = torch.stack([x, y, z], dim=1).contiguous()
    s0 = 0.02 * (1e5 / max(P, 1)) ** (1.0 / 3.0)
    scales = torch.exp(math.log(s0) + 0.6 * N(P, 3)).contiguous()
    quat = N(P, 4)
    quat = (quat / quat.norm(dim=1, keepdim=True)).contiguous()
    opac = torch.sigmoid(1.5 * N(P))[:, None].contiguous()
    M = ((max_sh_degree if max_sh_degree is not None else sh_degree) + 1) ** 2
    shs = torch.empty(P, M, 3)
    shs[:, 0, :] = 0.8 * N(P, 3)
    if M > 1:
        shs[:, 1:, :] = 0.1 * N(P, M - 1, 3)
    colors = torch.rand(P, 3, generator=g).contiguous()
    is_object = (torch.rand(P, generator=g) < 0.10).to(torch.int32)
    return Scene(means.float(), scales.float(), quat.float(), opac.float(), shs.float().contiguous(), sh_degree,
                 colors.float(), is_object, torch.tensor(bg, dtype=torch.float32), cam)


def upstream_weights(width: int, height: int, seed: int = 0):
    """Fixed random dL/d(color, depth, alpha) images so all three upstream grads are non-trivial."""
    g = torch.Generator("cpu").manual_seed(int(seed) + 7919)
    n = float(width * height)
    return (torch.randn(3, height, width, generator=g) / n, torch.randn(1, height, width, generator=g) / n,
            torch.randn(1, height, width, generator=g) / n)


def object_pose() -> torch.Tensor:
    """The C4 rigid pose R = Rz(5 deg) Ry(-3 deg), t = (0.05, -0.02, 0.1) as 12 floats (R row-major, t)."""
    a, b = math.radians(5.0), math.radians(-3.0)
    Rz = torch.tensor([[math.cos(a), -math.sin(a), 0.0], [math.sin(a), math.cos(a), 0.0], [0.0, 0.0, 1.0]])
    Ry = torch.tensor([[math.cos(b), 0.0, math.sin(b)], [0.0, 1.0, 0.0], [-math.sin(b), 0.0, math.cos(b)]])
    R = Rz @ Ry
    return torch.cat([R.reshape(-1), torch.tensor([0.05, -0.02, 0.1])]).float().contiguous()


CONFIGS = {
    # name: (P, sh_degree, width, height) -- BASELINE.json `configs`
    "c1": (10_000, 0, 256, 256),
    "c2": (100_000, 3, 1920, 1080),
    "c3": (1_000_000, 3, 1280, 720),
    "tiny": (2_000, 3, 128, 96),
}
