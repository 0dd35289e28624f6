"""SubTerrain -- heightfield container of the tracking env (reference: PARC/util/terrain_util.py:18-156).

Holds `hf [X, Y]` fp32 (row-major, x-major), `min_point [2]`, `dxdy [2]`, `dims [2]` i64 exactly
like the reference; the per-step lookups (`get_hf_val_from_points`, the ray-height observation)
run in the CUDA kernels, which receive the grid through `cstruct()`.
"""
import copy

import numpy as np
import torch

from . import _lib
from .file_io import MSTerrainData


class SubTerrain:
    def __init__(self, terrain_name="terrain", x_dim=256, y_dim=256, dx=1.0, dy=1.0, min_x=-1.0, min_y=-1.0,
                 device="cuda:0"):
        self.terrain_name = terrain_name
        self.hf = torch.zeros(size=(x_dim, y_dim), dtype=torch.float32, device=device)
        self.dims = torch.tensor([x_dim, y_dim], dtype=torch.int64, device=device)
        self.min_point = torch.tensor([min_x, min_y], dtype=torch.float32, device=device)
        self.dxdy = torch.tensor([dx, dy], dtype=torch.float32, device=device)
        self.hf_mask = torch.zeros(size=(x_dim, y_dim), dtype=torch.bool, device=device)
        self.hf_maxmin = torch.zeros(size=(x_dim, y_dim, 2), dtype=torch.float32, device=device)
        self.hf_maxmin[..., 0] = 1.0
        self.hf_maxmin[..., 1] = -1.0

    @classmethod
    def from_ms_terrain_data(cls, terrain_data: MSTerrainData, device):
        hf = terrain_data.hf
        mp = terrain_data.min_point
        inst = cls(x_dim=hf.shape[0], y_dim=hf.shape[1], dx=terrain_data.dx, dy=terrain_data.dx,
                   min_x=float(mp[0]), min_y=float(mp[1]), device=device)
        inst.hf = torch.as_tensor(hf).clone().to(dtype=torch.float32, device=device)
        if terrain_data.hf_maxmin is not None:
            inst.hf_maxmin = torch.as_tensor(terrain_data.hf_maxmin).to(dtype=torch.float32, device=device)
        return inst

    def to_ms_terrain_data(self):
        return MSTerrainData(hf=self.hf.cpu().numpy(), hf_maxmin=self.hf_maxmin.cpu().numpy(),
                             min_point=self.min_point.cpu().numpy(), dx=self.dxdy[0].item())

    def set_device(self, device):
        for name in ("hf", "dims", "min_point", "dxdy", "hf_mask", "hf_maxmin"):
            setattr(self, name, getattr(self, name).to(device=device))

    def torch_copy(self):
        new = copy.copy(self)
        for name in ("hf", "dims", "min_point", "dxdy", "hf_mask", "hf_maxmin"):
            setattr(new, name, getattr(self, name).clone())
        return new

    def get_real_size(self):
        return self.dims * self.dxdy

    def get_max_point(self):
        return self.min_point + self.get_real_size() - self.dxdy

    def get_point(self, ij):
        return self.min_point + ij * self.dxdy

    def pad(self, padding_size, height=0.0):
        """Reference: terrain_util.py:243-262 (host setup, once per clip)."""
        p = padding_size
        self.hf = torch.nn.functional.pad(self.hf, [p, p, p, p], value=height)
        new_mask = torch.zeros((self.hf_mask.shape[0] + 2 * p, self.hf_mask.shape[1] + 2 * p), dtype=torch.bool,
                               device=self.hf.device)
        if p > 0:
            new_mask[p:-p, p:-p] = self.hf_mask
        else:
            new_mask[:] = self.hf_mask
        self.hf_mask = new_mask
        max_val = torch.max(self.hf_maxmin[..., 0]).item()
        min_val = torch.min(self.hf_maxmin[..., 1]).item()
        hi = torch.nn.functional.pad(self.hf_maxmin[..., 0], [p, p, p, p], value=max_val)
        lo = torch.nn.functional.pad(self.hf_maxmin[..., 1], [p, p, p, p], value=min_val)
        self.hf_maxmin = torch.stack([hi, lo], dim=-1)
        self.min_point = self.min_point - self.dxdy * p
        self.dims = self.dims + 2 * p

    # ---- C-ABI descriptor -------------------------------------------------------------------
    def cstruct(self):
        hf = self.hf
        assert hf.dtype == torch.float32 and hf.is_contiguous(), "hf must be contiguous fp32"
        t = _lib.TerrainStruct()
        t.hf = _lib.ptr(hf)
        t.dim_x, t.dim_y = int(hf.shape[0]), int(hf.shape[1])
        mp = self.min_point.detach().cpu()
        dd = self.dxdy.detach().cpu()
        t.min_x, t.min_y = float(mp[0]), float(mp[1])
        t.dx, t.dy = float(dd[0]), float(dd[1])
        return t

    # ---- kernel-backed lookups ----------------------------------------------------------------
    def get_grid_index(self, point: torch.Tensor):
        """clamp(round_half_even((p - min) / dxdy), 0, dims - 1) -> i64 [...,2] (:143-152)."""
        lead = point.shape[:-1]
        pts = point.reshape(-1, 2).contiguous().float()
        n = pts.shape[0]
        # each point is its own "env" with a single zero ray: reuses the ray-height kernel
        root = torch.cat([pts, torch.zeros((n, 1), dtype=torch.float32, device=pts.device)], dim=-1).contiguous()
        heading = torch.zeros((n,), dtype=torch.float32, device=pts.device)
        ray = torch.zeros((1, 2), dtype=torch.float32, device=pts.device)
        out = torch.empty((n, 1), dtype=torch.float32, device=pts.device)
        idx = torch.empty((n, 1, 2), dtype=torch.int64, device=pts.device)
        ts = self.cstruct()
        _lib.check(_lib.lib().parc_b200_ray_heights(ts, _lib.ptr(root), _lib.ptr(heading), _lib.ptr(ray), 1, n,
                                                  -3.0e38, 3.0e38, _lib.ptr(out), 1, _lib.ptr(idx),
                                                  _lib.stream_ptr(pts.device)))
        return idx.reshape(*lead, 2)

    def get_hf_val_from_points(self, xy_points):
        """Nearest-cell height of each xy point (:154-156)."""
        lead = xy_points.shape[:-1]
        pts = xy_points.reshape(-1, 2).contiguous().float()
        n = pts.shape[0]
        root = torch.cat([pts, torch.zeros((n, 1), dtype=torch.float32, device=pts.device)], dim=-1).contiguous()
        heading = torch.zeros((n,), dtype=torch.float32, device=pts.device)
        ray = torch.zeros((1, 2), dtype=torch.float32, device=pts.device)
        out = torch.empty((n, 1), dtype=torch.float32, device=pts.device)
        ts = self.cstruct()
        _lib.check(_lib.lib().parc_b200_ray_heights(ts, _lib.ptr(root), _lib.ptr(heading), _lib.ptr(ray), 1, n,
                                                  -3.0e38, 3.0e38, _lib.ptr(out), 1, None,
                                                  _lib.stream_ptr(pts.device)))
        return out.reshape(lead)


class _PointsHfSdfFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, hf, min_box_center, dx, dy, base_z, inverted):
        from . import _lib
        B, N = points.shape[0], points.shape[1]
        p = points.detach().to(torch.float32).contiguous()
        sd = torch.empty(B, N, dtype=torch.float32, device=p.device)
        grad = torch.empty(B, N, 3, dtype=torch.float32, device=p.device) if ctx.needs_input_grad[0] else None
        _lib.check(_lib.lib().parc_b200_points_hf_sdf(_lib.ptr(p), B, N, _lib.ptr(hf), hf.shape[1], hf.shape[2], _lib.ptr(min_box_center),
                                                      dx, dy, base_z, int(bool(inverted)), _lib.ptr(sd), _lib.ptr(grad),
                                                      _lib.stream_ptr(p.device)))
        if grad is not None:
            ctx.save_for_backward(grad)
        return sd

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return g.unsqueeze(-1) * grad, None, None, None, None, None, None


def points_hf_sdf(points, hf, hf_min_box_center, hf_dxdy, base_z=-10.0, inverted=True, radius=None):
    """terrain_util.points_hf_sdf (PARC/util/terrain_util.py:1736-1794): signed distance of points [B,N,3] to the
    box-per-cell heightfield hf [B,X,Y] whose (0,0) cell centre is hf_min_box_center [B,2]; differentiable in
    `points`. `radius` (:1782-1785, geom_util.sdRoundBox = sdBox - radius): every column's distance shrinks by the same
    amount, so the minimum over the columns shifts by exactly `radius` (the sign flips with `inverted`)."""
    assert points.dim() == 3 and hf.dim() == 3 and hf_min_box_center.dim() == 2 and hf_dxdy.dim() == 1
    dev = points.device
    hf = hf.to(device=dev, dtype=torch.float32).contiguous()
    mc = hf_min_box_center.to(device=dev, dtype=torch.float32).contiguous()
    sd = _PointsHfSdfFn.apply(points, hf, mc, float(hf_dxdy[0]), float(hf_dxdy[1]), float(base_z), bool(inverted))
    if radius is not None:
        assert isinstance(radius, float) and radius > 0.0          # :1783-1784
        sd = sd + radius if inverted else sd - radius
    return sd


# ---- motion-aware terrain post-processing -----------------------------------------
def _mask_pass(motion_frames, terrain: SubTerrain, char_model, char_body_points):
    """FK + parc_b200_hf_mask_inds: -> (inds [T,Np,2] i64, counts [T] i32, min_body_heights [X,Y], mask [X,Y] u8)"""
    root_pos, root_rot, joint_rot = motion_frames.root_pos, motion_frames.root_rot, motion_frames.joint_rot
    assert root_pos.dim() == 2 and root_rot.dim() == 2 and joint_rot.dim() == 3, "one motion: [T,3], [T,4], [T,J-1,4]"
    dev, f32 = terrain.hf.device, torch.float32
    body_pos, body_rot = char_model.forward_kinematics(root_pos.to(dev), root_rot.to(dev), joint_rot.to(dev))      # :1865
    body_pos, body_rot = body_pos.contiguous().float(), body_rot.contiguous().float()
    T, J = int(root_pos.shape[0]), char_model.get_num_joints()
    assert len(char_body_points) == J
    pts = torch.cat([p.to(device=dev, dtype=f32) for p in char_body_points], dim=0).contiguous()
    starts = [0]
    for p in char_body_points:
        starts.append(starts[-1] + int(p.shape[0]))
    Np = starts[-1]
    point_start = torch.tensor(starts, dtype=torch.int32, device=dev)
    X, Y = terrain.hf.shape
    inds = torch.empty((T, Np, 2), dtype=torch.int64, device=dev)
    counts = torch.empty((T,), dtype=torch.int32, device=dev)
    mins = torch.empty((X, Y), dtype=f32, device=dev)
    mask = torch.empty((X, Y), dtype=torch.uint8, device=dev)
    p = _lib.ptr
    _lib.check(_lib.lib().parc_b200_hf_mask_inds(char_model._cstruct(), terrain.cstruct(), p(body_pos), p(body_rot), T, p(pts), p(point_start),
                                                 Np, p(inds), p(counts), p(mins), p(mask), _lib.stream_ptr(dev)))
    return inds, counts, mins, mask


def compute_hf_mask_inds(motion_frames, terrain: SubTerrain, char_model, char_body_points):
    """terrain_util.compute_hf_mask_inds (PARC/util/terrain_util.py:1851-1897): for each frame the unique grid cells
    under the character's body sample points (sorted as torch.unique(dim=0) sorts them), and the lowest sample height
    seen over every cell. One launch with a block per frame instead of the frames x bodies x points Python loop; the
    only host synchronisation is the read of the per-frame counts that cuts the list."""
    terrain.hf_mask[...] = False                                                                                   # :1869
    inds, counts, mins, _ = _mask_pass(motion_frames, terrain, char_model, char_body_points)
    counts = counts.cpu().tolist()
    return [inds[t, :c] for t, c in enumerate(counts)], mins


def compute_hf_mask_from_inds(terrain: SubTerrain, mask_grid_inds):
    """:1899-1906"""
    hf_mask = torch.zeros_like(terrain.hf_mask)
    if len(mask_grid_inds) > 0:
        allc = torch.cat([x.reshape(-1, 2) for x in mask_grid_inds], dim=0)
        hf_mask[allc[:, 0], allc[:, 1]] = True
    return hf_mask


def compute_hf_mask(motion_frames, terrain: SubTerrain, char_model, char_body_points):
    """:1908-1915 -- the union comes straight out of the kernel"""
    return _mask_pass(motion_frames, terrain, char_model, char_body_points)[3].bool()


def compute_hf_extra_vals(motion_frames, terrain: SubTerrain, char_model, char_body_points, z_buf=3.0, jump_buf=0.8):
    """:1918-1947 -> (mask_grid_inds, copy of the terrain whose hf_maxmin tells the augmenter how far each cell may move)"""
    ret_terrain = terrain.torch_copy()
    ret_terrain.hf_mask[...] = False
    inds, counts, mins, mask = _mask_pass(motion_frames, ret_terrain, char_model, char_body_points)
    max_h = torch.max(motion_frames.root_pos[:, 2]).item()
    min_h = torch.min(ret_terrain.hf).item()
    mm = torch.empty_like(ret_terrain.hf_maxmin)
    p = _lib.ptr
    _lib.check(_lib.lib().parc_b200_hf_extra_vals(ret_terrain.cstruct(), p(mask), p(mins), float(max_h), float(min_h), float(z_buf),
                                                  float(jump_buf), p(mm), _lib.stream_ptr(mm.device)))
    ret_terrain.hf_maxmin = mm
    counts = counts.cpu().tolist()
    return [inds[t, :c] for t, c in enumerate(counts)], ret_terrain
