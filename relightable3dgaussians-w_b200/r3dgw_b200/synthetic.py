"""Seeded synthetic scenes shaped like the reference's workloads.

Everything is generated on the CPU in fp32 from `torch.Generator().manual_seed(20261104 + cfg_id)` and
only then moved to the device, so the CPU oracle, the reference extension and the B200 kernels all
see bit-identical inputs.

Camera convention restated from the reference: scene/cameras.py:77-80 (world_view_transform is the
TRANSPOSED world->view matrix, full_proj_transform = world_view_transform @ projection^T) and
utils/graphics_utils.py:60-80 (getProjectionMatrix, znear 0.01, zfar 100).
"""
import math
from dataclasses import dataclass, field
from typing import Optional

import torch

BASE_SEED = 20261104

# name -> (cfg_id, P, W, H, sigma_px)   sigma_px calibrated once and frozen
CONFIGS = {
    "A": dict(cfg_id=0, P=200_000, W=800, H=600, sigma_px=2.0),
    "B": dict(cfg_id=1, P=1_000_000, W=1920, H=1080, sigma_px=2.0),
    "E": dict(cfg_id=4, P=6_000_000, W=3840, H=2160, sigma_px=6.0),
    "tiny": dict(cfg_id=10, P=2_000, W=160, H=96, sigma_px=2.0),
    "small": dict(cfg_id=11, P=20_000, W=320, H=240, sigma_px=2.0),
}


@dataclass
class Camera:
    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    world_view_transform: torch.Tensor   # [4,4], transposed world->view
    full_proj_transform: torch.Tensor    # [4,4]
    camera_center: torch.Tensor          # [3]

    @property
    def tanfovx(self):
        return math.tan(self.FoVx * 0.5)

    @property
    def tanfovy(self):
        return math.tan(self.FoVy * 0.5)

    def to(self, device):
        return Camera(self.image_width, self.image_height, self.FoVx, self.FoVy,
                      self.world_view_transform.to(device), self.full_proj_transform.to(device),
                      self.camera_center.to(device))


def projection_matrix(znear, zfar, fovx, fovy):
    """OpenGL-style perspective with z in [0,1] and w = +z (graphics_utils.py:60-80)."""
    ty, tx = math.tan(fovy / 2), math.tan(fovx / 2)
    top, right = ty * znear, tx * znear
    Pm = torch.zeros(4, 4, dtype=torch.float32)
    Pm[0, 0] = 2.0 * znear / (2 * right)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(W, H, fovx_deg=60.0, R: Optional[torch.Tensor] = None, T: Optional[torch.Tensor] = None):
    """R is the camera-to-world rotation and T the world->view translation, as in scene/cameras.py."""
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(math.tan(fovx / 2) * H / W)
    R = torch.eye(3, dtype=torch.float64) if R is None else R.double()
    T = torch.zeros(3, dtype=torch.float64) if T is None else T.double()
    Rt = torch.zeros(4, 4, dtype=torch.float64)
    Rt[:3, :3] = R.t()
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    w2v = Rt.float()
    wvt = w2v.t().contiguous()
    proj = projection_matrix(0.01, 100.0, fovx, fovy).t().contiguous()
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    return Camera(W, H, fovx, fovy, wvt, full, center)


def arc_cameras(W, H, n_views, centroid_z=8.0, step_deg=10.0):
    """V cameras on a `step_deg` arc around the scene centroid (0,0,centroid_z), all looking at it."""
    cams = []
    for v in range(n_views):
        ang = math.radians(step_deg * (v - (n_views - 1) / 2.0))
        # camera position on a circle of radius centroid_z around the centroid, in the xz plane
        pos = torch.tensor([centroid_z * math.sin(ang), 0.0, centroid_z - centroid_z * math.cos(ang)], dtype=torch.float64)
        fwd = torch.tensor([0.0, 0.0, centroid_z], dtype=torch.float64) - pos
        fwd = fwd / fwd.norm()
        up = torch.tensor([0.0, 1.0, 0.0], dtype=torch.float64)
        right = torch.linalg.cross(up, fwd)
        right = right / right.norm()
        up2 = torch.linalg.cross(fwd, right)
        R_c2w = torch.stack([right, up2, fwd], dim=1)     # columns = camera axes in world
        T = -(R_c2w.t() @ pos)
        cams.append(make_camera(W, H, R=R_c2w, T=T))
    return cams


@dataclass
class Scene:
    name: str
    camera: Camera
    means3D: torch.Tensor     # [P,3]
    scales: torch.Tensor      # [P,3] post-activation
    rotations: torch.Tensor   # [P,4] normalised (r,x,y,z)
    opacities: torch.Tensor   # [P,1]
    is_sky: torch.Tensor      # [P] bool
    albedo: torch.Tensor      # [Pfg,3]
    roughness: torch.Tensor   # [Pfg,1]
    metalness: torch.Tensor   # [Pfg,1]
    env_sh: torch.Tensor      # [25,3]
    sky_sh: torch.Tensor      # [4,3]
    colors: torch.Tensor      # [P,3] generic U1 colours
    bg: torch.Tensor          # [3]
    extra: dict = field(default_factory=dict)

    @property
    def P(self):
        return int(self.means3D.shape[0])

    def to(self, device):
        kw = {}
        for k, v in self.__dict__.items():
            if isinstance(v, torch.Tensor):
                kw[k] = v.to(device)
            elif isinstance(v, Camera):
                kw[k] = v.to(device)
            else:
                kw[k] = v
        return Scene(**kw)


def make_env_sh(gen, n=1):
    """[n,25,3]: base[0] ~ U(2.5,3.5) per channel, band l>=1 ~ N(0, 0.4/(l+1)) (keeps irradiance > 0)."""
    out = torch.zeros(n, 25, 3)
    out[:, 0, :] = 2.5 + torch.rand(n, 3, generator=gen)
    for l in range(1, 5):
        out[:, l * l:(l + 1) * (l + 1), :] = torch.randn(n, 2 * l + 1, 3, generator=gen) * (0.4 / (l + 1))
    return out


def make_scene(name="tiny", P=None, W=None, H=None, sigma_px=None, seed_offset=None, sky_fraction=0.05) -> Scene:
    cfg = dict(CONFIGS.get(name, CONFIGS["tiny"]))
    if P is not None:
        cfg["P"] = P
    if W is not None:
        cfg["W"] = W
    if H is not None:
        cfg["H"] = H
    if sigma_px is not None:
        cfg["sigma_px"] = sigma_px
    P, W, H, sig = cfg["P"], cfg["W"], cfg["H"], cfg["sigma_px"]
    gen = torch.Generator(device="cpu").manual_seed(BASE_SEED + (cfg["cfg_id"] if seed_offset is None else seed_offset))
    cam = make_camera(W, H)
    tfx, tfy = cam.tanfovx, cam.tanfovy
    focal_x = W / (2.0 * tfx)

    u = torch.rand(P, generator=gen)
    n_sky = int(round(P * sky_fraction))
    is_sky = torch.zeros(P, dtype=torch.bool)
    if n_sky > 0:
        is_sky[torch.randperm(P, generator=gen)[:n_sky]] = True

    # positions: 85 % in frustum, 10 % lateral outliers, 5 % behind / near plane
    z = torch.exp(math.log(2.0) + torch.rand(P, generator=gen) * (math.log(40.0) - math.log(2.0)))
    lx = (torch.rand(P, generator=gen) * 2 - 1) * 1.1
    ly = (torch.rand(P, generator=gen) * 2 - 1) * 1.1
    outl = (u >= 0.85) & (u < 0.95)
    lx = torch.where(outl, (torch.rand(P, generator=gen) * 2 - 1) * 2.0, lx)
    ly = torch.where(outl, (torch.rand(P, generator=gen) * 2 - 1) * 2.0, ly)
    behind = u >= 0.95
    z = torch.where(behind, -5.0 + torch.rand(P, generator=gen) * 5.2, z)
    x = z.abs().clamp_min(0.3) * tfx * lx
    y = z.abs().clamp_min(0.3) * tfy * ly
    means = torch.stack([x, y, z], dim=1)

    # sky Gaussians on the upper hemisphere of radius 50 (y is down in COLMAP coordinates)
    if n_sky > 0:
        th = torch.rand(n_sky, generator=gen) * (math.pi / 2)      # scene/gaussian_model.py:96-103 angles; y = -cos(theta) <= 0 is up
        ph = (torch.rand(n_sky, generator=gen) * 2 - 1) * math.pi
        sx = torch.sin(th) * torch.sin(ph)
        sy = -torch.cos(th)
        sz = torch.sin(th) * torch.cos(ph)
        means[is_sky] = 50.0 * torch.stack([sx, sy, sz], dim=1)

    depth = means[:, 2].abs().clamp_min(0.3)
    base = (depth / focal_x).unsqueeze(1) * torch.exp(math.log(sig) + 0.7 * torch.randn(P, 3, generator=gen))
    aniso = torch.exp(0.5 * torch.randn(P, 3, generator=gen))
    scales = base * aniso
    small_axis = torch.randint(0, 3, (P,), generator=gen)
    scales[torch.arange(P), small_axis] *= 0.2
    if n_sky > 0:
        scales[is_sky] = scales[is_sky] * 4.0   # sky splats are large

    q = torch.randn(P, 4, generator=gen)
    rotations = q / q.norm(dim=1, keepdim=True)
    opacities = torch.sigmoid(2.0 * torch.randn(P, 1, generator=gen))

    Pfg = int((~is_sky).sum())
    albedo = torch.sigmoid(torch.randn(Pfg, 3, generator=gen))
    roughness = torch.sigmoid(torch.randn(Pfg, 1, generator=gen))
    metalness = torch.sigmoid(torch.randn(Pfg, 1, generator=gen))
    env_sh = make_env_sh(gen, 1)[0]
    sky_sh = 0.3 * torch.randn(4, 3, generator=gen)
    colors = torch.rand(P, 3, generator=gen)
    bg = torch.zeros(3)
    return Scene(name=name, camera=cam, means3D=means.contiguous(), scales=scales.contiguous(),
                 rotations=rotations.contiguous(), opacities=opacities.contiguous(), is_sky=is_sky,
                 albedo=albedo, roughness=roughness, metalness=metalness, env_sh=env_sh, sky_sh=sky_sh,
                 colors=colors, bg=bg, extra={"sigma_px": sig, "seed": BASE_SEED + cfg["cfg_id"]})
