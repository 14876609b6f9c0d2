"""Seeded synthetic scenes and cameras shaped like VCR-GauS training inputs.

Everything is generated on the CPU with ``torch.Generator().manual_seed(seed)`` so that every implementation (this
library, the CPU oracle, the reference CUDA build) sees identical bits.  No datasets, no checkpoints.

Conventions follow the reference:
  * camera matrices exactly like scene/cameras.py:62-73 + tools/graphics_utils.py:30-80
    (viewmatrix = w2c^T, projmatrix = viewmatrix @ proj^T, znear 0.01, zfar 100);
  * intrinsics like tools/graphics_utils.py:89-101;  per-pixel ray directions like get_all_px_dir (:143-155);
  * normals_precomp like gaussian_renderer/__init__.py:95-101 with GaussianModel.get_normal (scene/gaussian_model.py:168-192).
"""
import math
from dataclasses import dataclass
from typing import Optional

import torch


def _quat_to_rot(q):
    """build_rotation (tools/general_utils.py): q = (r, x, y, z), normalised -> R [N,3,3]."""
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.zeros((q.shape[0], 3, 3), dtype=q.dtype)
    R[:, 0, 0] = 1 - 2 * (y * y + z * z)
    R[:, 0, 1] = 2 * (x * y - r * z)
    R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z)
    R[:, 1, 1] = 1 - 2 * (x * x + z * z)
    R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y)
    R[:, 2, 1] = 2 * (y * z + r * x)
    R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


@dataclass
class Scene:
    means3D: torch.Tensor      # [P,3]
    scales: torch.Tensor       # [P,3]  (already exp-activated)
    rotations: torch.Tensor    # [P,4]  (already normalised)
    opacities: torch.Tensor    # [P,1]  (already sigmoid-activated)
    shs: torch.Tensor          # [P,16,3]
    semantics: torch.Tensor    # [P,2]
    extent: float

    @property
    def P(self):
        return self.means3D.shape[0]

    def to(self, device):
        return Scene(*(t.to(device) if isinstance(t, torch.Tensor) else t for t in
                       (self.means3D, self.scales, self.rotations, self.opacities, self.shs, self.semantics, self.extent)))


def make_scene(P: int, seed: int = 0, extent: float = 1.0, scale0: Optional[float] = None, sh_coeffs: int = 16) -> Scene:
    """70 % of the centres on random planes / spheres inside the cube [-extent, extent]^3 (surface-like, as VCR-GauS
    flattens Gaussians), 30 % uniform; log-normal scales with one axis x0.05 (flat disks); random unit quaternions;
    opacity = sigmoid(N(0, 2^2)); SH DC ~ U(-1.5, 1.5), rest ~ N(0, 0.1^2)."""
    g = torch.Generator().manual_seed(seed)
    n_surf = int(0.7 * P)
    n_uni = P - n_surf
    n_prims = 6
    prim = torch.randint(0, n_prims, (n_surf,), generator=g)
    pts = torch.empty((n_surf, 3))
    for k in range(n_prims):
        sel = (prim == k).nonzero(as_tuple=True)[0]
        if sel.numel() == 0:
            continue
        if k % 2 == 0:  # plane through a random point with a random normal
            c = (torch.rand(3, generator=g) - 0.5) * extent
            nrm = torch.randn(3, generator=g)
            nrm = nrm / nrm.norm()
            a = torch.linalg.cross(nrm, torch.tensor([0.3, 0.5, 0.8]))
            a = a / a.norm()
            b = torch.linalg.cross(nrm, a)
            uv = (torch.rand((sel.numel(), 2), generator=g) - 0.5) * 2 * extent
            pts[sel] = c + uv[:, :1] * a + uv[:, 1:] * b
        else:  # sphere
            c = (torch.rand(3, generator=g) - 0.5) * extent
            rad = (0.2 + 0.4 * torch.rand(1, generator=g)) * extent
            d = torch.randn((sel.numel(), 3), generator=g)
            d = d / d.norm(dim=1, keepdim=True)
            pts[sel] = c + rad * d
    pts = pts.clamp(-extent, extent)
    uni = (torch.rand((n_uni, 3), generator=g) - 0.5) * 2 * extent
    means = torch.cat([pts, uni], 0)
    means = means[torch.randperm(P, generator=g)]
    if scale0 is None:
        # mean spacing of P points on ~6 surfaces of area ~ (2*extent)^2
        scale0 = 0.9 * extent * math.sqrt(24.0 / max(P, 1))
    log_s = math.log(scale0) + 0.5 * torch.randn((P, 3), generator=g)
    flat_axis = torch.randint(0, 3, (P,), generator=g)
    scales = torch.exp(log_s)
    scales[torch.arange(P), flat_axis] *= 0.05
    q = torch.randn((P, 4), generator=g)
    q = q / q.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(2.0 * torch.randn((P, 1), generator=g))
    shs = 0.1 * torch.randn((P, sh_coeffs, 3), generator=g)
    shs[:, 0, :] = (torch.rand((P, 3), generator=g) - 0.5) * 3.0
    sem = torch.rand((P, 2), generator=g)
    return Scene(means.float().contiguous(), scales.float().contiguous(), q.float().contiguous(), opac.float().contiguous(),
                 shs.float().contiguous(), sem.float().contiguous(), float(extent))


@dataclass
class Camera:
    image_width: int
    image_height: int
    FoVx: float
    FoVy: float
    R: torch.Tensor                     # c2w rotation (the reference stores R = w2c^T, scene/cameras.py)
    T: torch.Tensor                     # w2c translation
    world_view_transform: torch.Tensor  # [4,4] = w2c^T
    full_proj_transform: torch.Tensor   # [4,4]
    camera_center: torch.Tensor         # [3]
    intr: torch.Tensor                  # [3,3]

    @property
    def tanfovx(self):
        return math.tan(self.FoVx * 0.5)

    @property
    def tanfovy(self):
        return math.tan(self.FoVy * 0.5)


def _projection(znear, zfar, fovX, fovY):
    tY, tX = math.tan(fovY / 2), math.tan(fovX / 2)
    top, right = tY * znear, tX * znear
    bottom, left = -top, -right
    Pm = torch.zeros(4, 4)
    Pm[0, 0] = 2.0 * znear / (right - left)
    Pm[1, 1] = 2.0 * znear / (top - bottom)
    Pm[0, 2] = (right + left) / (right - left)
    Pm[1, 2] = (top + bottom) / (top - bottom)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    return Pm


def make_camera(width: int, height: int, idx: int = 0, n_views: int = 8, extent: float = 1.0, radius_mul: float = 2.5,
                fill: float = 0.8, seed: int = 0) -> Camera:
    """Look-at pose number `idx` on a ring of radius radius_mul*extent around the origin; the field of view is chosen
    so that the cube fills ~`fill` of the frame."""
    ang = 2 * math.pi * idx / max(n_views, 1) + 0.1 * seed
    rad = radius_mul * extent
    eye = torch.tensor([rad * math.cos(ang), 0.35 * rad * math.sin(2 * ang + 0.3), rad * math.sin(ang)], dtype=torch.float64)
    fwd = -eye / eye.norm()
    up0 = torch.tensor([0.0, -1.0, 0.0], dtype=torch.float64)  # image y points down
    right = torch.linalg.cross(up0, fwd)  # noqa
    right = right / right.norm()
    down = torch.linalg.cross(fwd, right)
    c2w = torch.stack([right, down, fwd], dim=1)  # columns = camera axes in world
    w2c_R = c2w.t()
    T = -(w2c_R @ eye)
    Rt = torch.eye(4, dtype=torch.float64)
    Rt[:3, :3] = w2c_R
    Rt[:3, 3] = T
    half = math.sqrt(3.0) * extent * 0.75 / fill
    fovy = 2 * math.atan(half / (rad))
    fovx = 2 * math.atan(math.tan(fovy / 2) * width / height)
    wvt = Rt.float().t().contiguous()
    proj = _projection(0.01, 100.0, fovx, fovy).t()
    full = (wvt.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    center = wvt.inverse()[3, :3].contiguous()
    intr = torch.eye(3, dtype=torch.float32)
    intr[0, 0] = width / (2 * math.tan(fovx / 2))
    intr[1, 1] = height / (2 * math.tan(fovy / 2))
    intr[0, 2] = width / 2
    intr[1, 2] = height / 2
    return Camera(width, height, fovx, fovy, c2w.float(), T.float(), wvt, full, center, intr)


def pixel_dirs(cam: Camera) -> torch.Tensor:
    """get_all_px_dir (tools/graphics_utils.py:143-155): unit ray directions, [3,H,W]."""
    H, W = cam.image_height, cam.image_width
    xs = (torch.arange(W, dtype=torch.float32) + 0.5) / (W - 1)
    ys = (torch.arange(H, dtype=torch.float32) + 0.5) / (H - 1)
    vy, vx = torch.meshgrid(ys, xs, indexing="ij")
    ndc = torch.stack([vx, vy, torch.ones_like(vx)], dim=-1)
    inv_scale = torch.tensor([[W - 1, H - 1]], dtype=torch.float32)
    cam_xy = ndc[..., :2] * inv_scale * ndc[..., 2:3]
    cam_xyz = torch.cat([cam_xy, ndc[..., 2:3]], dim=-1) @ torch.inverse(cam.intr.t())
    d = torch.nn.functional.normalize(cam_xyz, dim=-1)
    return d.permute(2, 0, 1).contiguous()


def camera_space_normals(scene: Scene, cam: Camera) -> torch.Tensor:
    """normals_precomp of gaussian_renderer/__init__.py:95-101: min-scale axis of R(q), flipped away from the camera,
    rotated into camera space."""
    rots = _quat_to_rot(scene.rotations)
    axis = torch.argmin(scene.scales, dim=-1)
    normal = rots.gather(2, axis[:, None, None].expand(-1, 3, -1)).squeeze(-1)
    view_dir = scene.means3D - cam.camera_center
    normal = normal * ((((view_dir * normal).sum(dim=-1) > 0) * 1 - 0.5) * 2)[..., None]
    R_w2c = cam.R.t().to(torch.float32)
    return (normal @ R_w2c.transpose(0, 1)).contiguous()


def raster_inputs(scene: Scene, cam: Camera, bg=None, use_dirs=True, use_sem=False, sh_degree=3, seed=0):
    """All tensors one rasterizer call needs, on the CPU (move with .to(device))."""
    if bg is None:
        bg = torch.rand(3, generator=torch.Generator().manual_seed(1000 + seed))
    return dict(
        means3D=scene.means3D, shs=scene.shs, opacities=scene.opacities, scales=scene.scales, rotations=scene.rotations,
        normals=camera_space_normals(scene, cam), semantics=scene.semantics if use_sem else None,
        dirs=pixel_dirs(cam) if use_dirs else None, bg=bg.float(), viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, campos=cam.camera_center, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy,
        W=cam.image_width, H=cam.image_height, sh_degree=sh_degree)
