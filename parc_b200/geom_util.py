"""Body surface sample points of a character (reference: PARC/util/geom_util.py:745-872).

Host-side setup: runs once per character on tiny tensors and produces a *kernel input* (the point
table of `parc_b200_motion_loss`), exactly as in the reference. The sphere sampling is
`trimesh.creation.icosphere(subdivisions=0)` there, i.e. the 12 icosahedron vertices; trimesh is not
a dependency here, so the vertices are written out (their order does not affect any loss term: sums
and per-body minima)."""
import math

import torch

from .kin_char_model import GeomType, KinCharModel


def get_sphere_point_surface_samples(radius, device, num_subdivisions=0):
    """geom_util.py:743-751 (subdivisions=0 only)."""
    assert num_subdivisions == 0, "only the raw icosahedron (subdivisions=0) is supported"
    p = (1.0 + 5.0 ** 0.5) / 2.0
    v = torch.tensor([(-1, p, 0), (1, p, 0), (-1, -p, 0), (1, -p, 0), (0, -1, p), (0, 1, p), (0, -1, -p), (0, 1, -p),
                      (p, 0, -1), (p, 0, 1), (-p, 0, -1), (-p, 0, 1)], dtype=torch.float64)
    v = v / v.norm(dim=-1, keepdim=True) * radius
    return v.to(device=device, dtype=torch.float32)


def get_box_point_surface_samples(box_halfdims, device, num_slices=2, dim_x=6, dim_y=3):
    """geom_util.py:726-741: dim_x * dim_y grid on `num_slices` z levels."""
    z = torch.linspace(0.0, 1.0, num_slices, device=device) * box_halfdims[2] * 2.0 - box_halfdims[2]
    x = torch.linspace(0.0, 1.0, dim_x, device=device) * box_halfdims[0] * 2.0 - box_halfdims[0]
    y = torch.linspace(0.0, 1.0, dim_y, device=device) * box_halfdims[1] * 2.0 - box_halfdims[1]
    x, y = torch.meshgrid(x, y, indexing="ij")
    n = dim_x * dim_y
    pts = torch.cat([x.reshape(-1)[None, :, None].expand(num_slices, n, 1), y.reshape(-1)[None, :, None].expand(num_slices, n, 1),
                     z[:, None, None].expand(num_slices, n, 1)], dim=-1)
    return pts.reshape(-1, 3)


def get_capsule_point_surface_samples(capsule_length, capsule_radius, device, num_cylinder_slices=3, num_circle_points=4,
                                      num_sphere_subdivisons=0, ignore_hemispheres=True):
    """geom_util.py:753-788: rings of `num_circle_points` on `num_cylinder_slices` levels of the cylinder part."""
    pts = []
    if not ignore_hemispheres:
        sp = get_sphere_point_surface_samples(capsule_radius, device, num_sphere_subdivisons)
        up = sp[sp[..., 2] > 1e-5].clone()
        up[..., 2] += capsule_length / 2.0
        lo = sp[sp[..., 2] < -1e-5].clone()
        lo[..., 2] -= capsule_length / 2.0
        pts += [up, lo]
    z = torch.linspace(0.0, 1.0, num_cylinder_slices, device=device) * capsule_length - capsule_length / 2.0
    theta = torch.linspace(0, 2 * math.pi, num_circle_points + 1, device=device)[:-1]
    x, y = capsule_radius * torch.cos(theta), capsule_radius * torch.sin(theta)
    mid = torch.cat([x[:, None, None].expand(num_circle_points, num_cylinder_slices, 1),
                     y[:, None, None].expand(num_circle_points, num_cylinder_slices, 1),
                     z[None, :, None].expand(num_circle_points, num_cylinder_slices, 1)], dim=-1).reshape(-1, 3)
    pts.append(mid)
    return torch.cat(pts, dim=0)


def _quat_rotate(q, v):
    qv, w = q[..., :3], q[..., 3:4]
    t = 2.0 * torch.cross(qv.expand_as(v), v, dim=-1)
    return v + w * t + torch.cross(qv.expand_as(v), t, dim=-1)


def get_char_point_samples(char_model: KinCharModel, sphere_num_subdivisions=0, box_num_slices=2, box_dim_x=3, box_dim_y=6,
                           capsule_num_circle_points=4, capsule_num_sphere_subdivisons=0, capsule_num_cylinder_slices=4):
    """geom_util.py:790-872 -> list of [P_b, 3] tensors, one per body, in the body's frame."""
    device = char_model._device
    out = []
    for b in range(char_model.get_num_joints()):
        cur = []
        geoms = char_model.get_geoms(b)
        for geom in geoms:
            if geom._shape_type == GeomType.SPHERE:
                cur.append(get_sphere_point_surface_samples(geom._dims.item(), device, sphere_num_subdivisions) + geom._offset)
            elif geom._shape_type == GeomType.BOX:
                cur.append(get_box_point_surface_samples(geom._dims, device, box_num_slices, box_dim_x, box_dim_y) + geom._offset)
            elif geom._shape_type == GeomType.CAPSULE:
                fromto = geom._dims
                offset = geom._offset + fromto / 2.0
                z_axis = torch.tensor([0.0, 0.0, 1.0], device=device)
                axis = torch.cross(z_axis, fromto, dim=-1)
                axis = z_axis if torch.norm(axis) < 1e-5 else axis / torch.norm(axis)
                angle = torch.acos(torch.dot(axis, fromto))                       # as written in the reference (:835)
                n = axis / axis.norm().clamp(min=1e-9)
                q = torch.cat([n * torch.sin(angle / 2), torch.cos(angle / 2)[None]])
                q = q / q.norm()
                cp = get_capsule_point_surface_samples(torch.norm(fromto).item(), geom._radius, device,
                                                       num_cylinder_slices=capsule_num_cylinder_slices,
                                                       num_circle_points=capsule_num_circle_points,
                                                       num_sphere_subdivisons=capsule_num_sphere_subdivisons)
                cur.append(_quat_rotate(q[None], cp) + offset)
            else:   # CYLINDER / MESH: a single origin point, as in the reference (:855-860)
                cur.append(torch.zeros(1, 3, dtype=torch.float32, device=device))
        if len(geoms) == 0:
            cur.append(torch.zeros(1, 3, dtype=torch.float32, device=device))
        out.append(torch.cat(cur, dim=0))
    return out
