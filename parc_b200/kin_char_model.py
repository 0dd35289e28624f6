"""KinCharModel -- kinematic tree of the character, backed by the sm_100a kernels.

Host-side mirror of the reference class (PARC/anim/kin_char_model.py:149-701): same constructor,
`init`, `load_char_file`, getters and the three hot methods `forward_kinematics` (:617-649),
`dof_to_rot` (:586-599), `rot_to_dof` (:601-615) with identical argument meaning. The per-joint
Python loops of the reference are replaced by one CUDA launch each (parc_b200/csrc/kin_kernels.cu).
The MJCF parser here is an independent implementation of the same conventions: bodies in
depth-first document order, three hinges on a body = one SPHERICAL joint (exp-map dofs), one hinge
= HINGE, none = FIXED, MJCF quaternions are (w,x,y,z) and stored as (x,y,z,w), joint ranges in
degrees (:821-916).
"""
import copy
import enum
import json
import math
import os
import xml.etree.ElementTree as ET
from typing import List, Optional

import numpy as np
import torch

from . import _lib


class JointType(enum.Enum):
    ROOT = 0
    HINGE = 1
    SPHERICAL = 2
    FIXED = 3


class GeomType(enum.Enum):
    BOX = 0
    SPHERE = 1
    CAPSULE = 2
    CYLINDER = 3
    MESH = 4


_DOF_DIM = {JointType.ROOT: 0, JointType.HINGE: 1, JointType.SPHERICAL: 3, JointType.FIXED: 0}


class Joint:
    """One joint of the tree (reference: kin_char_model.py:21-104)."""

    def __init__(self, name, joint_type, axis, limits=None):
        self.name = name
        self.joint_type = joint_type
        self.axis = axis
        self.dof_idx = -1
        self.limits = limits

    def get_copy(self, new_device):
        j = Joint(self.name, self.joint_type,
                  None if self.axis is None else self.axis.clone().to(device=new_device),
                  None if self.limits is None else self.limits.clone().to(device=new_device))
        j.dof_idx = self.dof_idx
        return j

    def get_dof_dim(self):
        return _DOF_DIM[self.joint_type]

    def get_joint_dof(self, dof):
        return dof[..., self.dof_idx:self.dof_idx + self.get_dof_dim()]

    def set_joint_dof(self, j_dof, out_dof):
        out_dof[..., self.dof_idx:self.dof_idx + self.get_dof_dim()] = j_dof


class Geom:
    """Collision shape attached to a body (reference: kin_char_model.py:113-147)."""

    def __init__(self, shape_type, offset, dims, device, quat=None, radius=None, mesh_name=None, name=None):
        self._shape_type = shape_type
        self._offset = offset if isinstance(offset, torch.Tensor) else torch.tensor(np.asarray(offset), dtype=torch.float32, device=device)
        self._dims = dims if isinstance(dims, torch.Tensor) else torch.tensor(np.asarray(dims), dtype=torch.float32, device=device)
        self._radius = radius
        self._mesh_name = mesh_name
        self._quat = quat
        self._name = name

    def get_copy(self, new_device):
        return Geom(self._shape_type, self._offset.clone().to(new_device), self._dims.clone().to(new_device), new_device,
                    quat=self._quat, radius=self._radius, mesh_name=self._mesh_name, name=self._name)


def _floats(text, default=None):
    if text is None:
        return default
    return np.array([float(t) for t in text.split()], dtype=np.float64)


def _mjcf_quat_xyzw(text):
    q = _floats(text, np.array([1.0, 0.0, 0.0, 0.0]))
    return np.array([q[1], q[2], q[3], q[0]])


class _FKFunction(torch.autograd.Function):
    """forward_kinematics with an analytic vector-Jacobian product (the reference differentiates
    through the torch ops of kin_char_model.py:617-649, e.g. motion_optimization.py:463)."""

    @staticmethod
    def forward(ctx, model, root_pos, root_rot, joint_rot):
        lead = root_pos.shape[:-1]
        J = model.get_num_joints()
        rp = root_pos.reshape(-1, 3).contiguous().float()
        rr = root_rot.reshape(-1, 4).contiguous().float()
        jr = joint_rot.reshape(-1, max(J - 1, 0), 4).contiguous().float()
        n = rp.shape[0]
        body_pos = torch.empty((n, J, 3), dtype=torch.float32, device=rp.device)
        body_rot = torch.empty((n, J, 4), dtype=torch.float32, device=rp.device)
        _lib.check(_lib.lib().parc_b200_forward_kinematics(
            model._cstruct(), _lib.ptr(rp), _lib.ptr(rr), _lib.ptr(jr) if J > 1 else None, n,
            _lib.ptr(body_pos), _lib.ptr(body_rot), _lib.stream_ptr(rp.device)))
        ctx.model = model
        ctx.save_for_backward(rp, rr, jr)
        ctx.lead = lead
        return body_pos.reshape(*lead, J, 3), body_rot.reshape(*lead, J, 4)

    @staticmethod
    def backward(ctx, g_pos, g_rot):
        rp, rr, jr = ctx.saved_tensors
        model = ctx.model
        J = model.get_num_joints()
        n = rp.shape[0]
        gp = None if g_pos is None else g_pos.reshape(n, J, 3).contiguous().float()
        gr = None if g_rot is None else g_rot.reshape(n, J, 4).contiguous().float()
        g_rp = torch.empty_like(rp)
        g_rr = torch.empty_like(rr)
        g_jr = torch.empty_like(jr)
        _lib.check(_lib.lib().parc_b200_forward_kinematics_bwd(
            model._cstruct(), _lib.ptr(rp), _lib.ptr(rr), _lib.ptr(jr) if J > 1 else None, n,
            _lib.ptr(gp), _lib.ptr(gr), _lib.ptr(g_rp), _lib.ptr(g_rr), _lib.ptr(g_jr) if J > 1 else None,
            _lib.stream_ptr(rp.device)))
        lead = ctx.lead
        return None, g_rp.reshape(*lead, 3), g_rr.reshape(*lead, 4), g_jr.reshape(*lead, max(J - 1, 0), 4)


class KinCharModel:
    def __init__(self, device):
        self._device = torch.device(device)
        self._cm = None

    # ------------------------------------------------------------------ construction
    def init(self, body_names, parent_indices, local_translation, local_rotation, joints, geoms=None,
             contact_body_names=None):
        num_bodies = len(body_names)
        assert len(parent_indices) == num_bodies
        assert len(local_translation) == num_bodies
        assert len(local_rotation) == num_bodies
        assert len(joints) == num_bodies

        self._body_names = body_names
        if not isinstance(parent_indices, torch.Tensor):
            parent_indices = torch.tensor(parent_indices, device=self._device, dtype=torch.long)
        self._parent_indices = parent_indices
        if not isinstance(local_translation, torch.Tensor):
            local_translation = torch.tensor(np.array(local_translation), device=self._device, dtype=torch.float32)
        self._local_translation = local_translation
        self._original_local_translation = self._local_translation.clone()
        if not isinstance(local_rotation, torch.Tensor):
            local_rotation = torch.tensor(np.array(local_rotation), device=self._device, dtype=torch.float32)
        self._local_rotation = local_rotation

        self._joints = joints
        dof_idx = 0
        for j in joints:                      # _label_dof_indices, :936-943
            if j is not None:
                j.dof_idx = dof_idx
                dof_idx += j.get_dof_dim()
        self._dof_size = dof_idx
        self._name_body_map = {n: i for i, n in enumerate(body_names)}
        self._lower_dof_limits, self._upper_dof_limits = self._gather_joint_limits(joints)
        self._geoms = geoms

        if contact_body_names is None:
            contact_body_names = copy.deepcopy(body_names)
            self._contact_body_ids = torch.arange(0, num_bodies, dtype=torch.int64, device=self._device)
        else:
            ids = [self.get_body_id(n) for n in contact_body_names]
            self._contact_body_ids = torch.tensor(ids, dtype=torch.int64, device=self._device)
        self._contact_body_names = contact_body_names
        self._name_contact_body_map = {n: i for i, n in enumerate(contact_body_names)}
        self._cm = None

    def _gather_joint_limits(self, joints):
        lo, hi = [], []
        for j in joints:
            if j.limits is None:
                continue
            if j.joint_type == JointType.HINGE:
                lo.append(j.limits[0].unsqueeze(0))
                hi.append(j.limits[1].unsqueeze(0))
            elif j.joint_type == JointType.SPHERICAL:
                lo.append(j.limits[:, 0])
                hi.append(j.limits[:, 1])
        lo = torch.cat(lo) if lo else lo
        hi = torch.cat(hi) if hi else hi
        return lo, hi

    def get_copy(self, new_device=None):
        if new_device is None:
            new_device = self._device
        m = KinCharModel(new_device)
        geoms = None
        if self._geoms is not None:
            geoms = [[g.get_copy(new_device) for g in body_geoms] for body_geoms in self._geoms]
        m.init(body_names=copy.deepcopy(self._body_names),
               contact_body_names=copy.deepcopy(self._contact_body_names),
               parent_indices=self._parent_indices.clone().to(device=new_device),
               local_translation=self._local_translation.clone().to(device=new_device),
               local_rotation=self._local_rotation.clone().to(device=new_device),
               joints=[j.get_copy(new_device) for j in self._joints],
               geoms=geoms)
        return m

    def load_char_file(self, char_file):
        """MJCF (.xml) or the JSON skeleton spec written by `to_spec` (.json)."""
        if os.path.splitext(str(char_file))[1] == ".json":
            with open(char_file, "r") as f:
                self.load_spec(json.load(f))
            return
        root = ET.parse(char_file).getroot()
        world = root.find("worldbody")
        assert world is not None
        first = world.find("body")
        assert first is not None

        default_joint_type = None                # _parse_default_joint_type, :945-958
        dflt = root.find("default")
        if dflt is not None:
            for d in dflt.findall("default"):
                if d.attrib.get("class") == "body" and d.find("joint") is not None:
                    default_joint_type = d.find("joint").attrib.get("type")
                    break
        default_geom = None
        if dflt is not None:
            g = dflt.find("geom")
            if g is None and dflt.find("default") is not None:
                g = dflt.find("default").find("geom")
            if g is not None:
                default_geom = g.attrib.get("type")
        geom_types = {"sphere": GeomType.SPHERE, "box": GeomType.BOX, "capsule": GeomType.CAPSULE,
                      "cylinder": GeomType.CYLINDER, "mesh": GeomType.MESH}
        default_geom_type = geom_types.get(default_geom, GeomType.SPHERE)

        names, parents, trans, rots, joints, geoms = [], [], [], [], [], []

        def parse_geoms(node, body_name):
            out = []
            for gi, gn in enumerate(node.findall("geom")):
                gtype = geom_types.get(gn.attrib.get("type"), default_geom_type)
                gname = gn.attrib.get("name") or "%s_geom_%d" % (body_name, gi)
                quat = _mjcf_quat_xyzw(gn.attrib.get("quat"))
                if gtype in (GeomType.SPHERE, GeomType.BOX):
                    offset = _floats(gn.attrib.get("pos"), np.zeros(3))
                    size = gn.attrib.get("size")
                    dims = np.array(0.1) if size is None else _floats(size)
                    out.append(Geom(gtype, offset, dims, self._device, quat=quat, name=gname))
                elif gtype == GeomType.CAPSULE:
                    ft = _floats(gn.attrib.get("fromto"))
                    out.append(Geom(gtype, ft[0:3], ft[3:6] - ft[0:3], self._device, quat=quat,
                                    radius=float(gn.attrib.get("size")), name=gname))
                elif gtype == GeomType.CYLINDER:
                    out.append(Geom(gtype, _floats(gn.attrib.get("pos"), np.zeros(3)), _floats(gn.attrib.get("size")),
                                    self._device, quat=quat, name=gname))
                else:
                    out.append(Geom(gtype, _floats(gn.attrib.get("pos"), np.zeros(3)), np.ones(3), self._device,
                                    radius=1.0, mesh_name=gn.attrib.get("mesh"), quat=quat, name=gname))
            return out

        def parse_joint(body_name, nodes):
            def limits_of(n):
                r = n.attrib.get("range")
                assert r is not None, "Need joint limits"
                return _floats(r)

            def check_no_offset(n):
                p = _floats(n.attrib.get("pos"))
                assert p is None or not np.any(p), "Joint offsets are not supported"

            if len(nodes) == 0:
                return Joint(body_name, JointType.FIXED, None)
            if len(nodes) == 3:
                lim = []
                for n in nodes:
                    check_no_offset(n)
                    lim.append(limits_of(n))
                    assert (n.attrib.get("type") or default_joint_type) == "hinge", "Invalid format for a spherical joint"
                lim = torch.from_numpy(np.stack(lim)).to(dtype=torch.float32, device=self._device) * (torch.pi / 180.0)
                name = nodes[0].attrib.get("name")
                return Joint(name[:name.rfind("_")], JointType.SPHERICAL, None, lim)
            assert len(nodes) == 1, "Series joints are not supported."
            jt = nodes[0].attrib.get("type") or default_joint_type
            if jt == "fixed":
                return Joint(body_name, JointType.FIXED, None)
            assert jt == "hinge", "Unsupported joint type: %s" % jt
            check_no_offset(nodes[0])
            axis = torch.tensor(_floats(nodes[0].attrib.get("axis")), device=self._device, dtype=torch.float32)
            lim = torch.from_numpy(limits_of(nodes[0])).to(dtype=torch.float32, device=self._device) * (torch.pi / 180.0)
            return Joint(nodes[0].attrib.get("name"), JointType.HINGE, axis, lim)

        def visit(node, parent):
            name = node.attrib.get("name")
            assert name is not None, "Body element is missing required 'name' attribute in character file."
            idx = len(names)
            names.append(name)
            parents.append(parent)
            trans.append(_floats(node.attrib.get("pos"), np.zeros(3)))
            rots.append(_mjcf_quat_xyzw(node.attrib.get("quat")) if node.attrib.get("quat") is not None
                        else np.array([0.0, 0.0, 0.0, 1.0]))
            geoms.append(parse_geoms(node, name))
            joints.append(Joint("root", JointType.ROOT, None) if idx == 0 else parse_joint(name, node.findall("joint")))
            for child in node.findall("body"):
                visit(child, idx)

        visit(first, -1)

        contact_names = None
        cci = root.find("custom_contact_info")
        if cci is not None:
            contact_names = []
            for e in cci.findall("body"):
                assert e.get("name") in names
                contact_names.append(e.get("name"))
        self._meshes = {}
        self.init(body_names=names, parent_indices=parents, local_translation=trans, local_rotation=rots,
                  joints=joints, geoms=geoms, contact_body_names=contact_names)

    # JSON skeleton spec: lets the GPU box build the repo humanoid without the reference tree.
    def to_spec(self):
        def geom_spec(g):
            return {"type": g._shape_type.name, "offset": g._offset.tolist(), "dims": np.atleast_1d(g._dims.cpu().numpy()).tolist(),
                    "radius": g._radius, "quat": None if g._quat is None else list(map(float, g._quat)), "name": g._name}
        return {
            "body_names": list(self._body_names),
            "parent_indices": [int(p) for p in self._parent_indices.tolist()],
            "local_translation": self._local_translation.double().tolist(),
            "local_rotation": self._local_rotation.double().tolist(),
            "joints": [{"name": j.name, "type": j.joint_type.name,
                        "axis": None if j.axis is None else j.axis.double().tolist(),
                        "limits": None if j.limits is None else j.limits.double().tolist()} for j in self._joints],
            "geoms": None if self._geoms is None else [[geom_spec(g) for g in bg] for bg in self._geoms],
            "contact_body_names": list(self._contact_body_names),
        }

    def load_spec(self, spec):
        dev = self._device
        joints = []
        for js in spec["joints"]:
            axis = None if js["axis"] is None else torch.tensor(js["axis"], dtype=torch.float32, device=dev)
            lim = None if js["limits"] is None else torch.tensor(js["limits"], dtype=torch.float32, device=dev)
            joints.append(Joint(js["name"], JointType[js["type"]], axis, lim))
        geoms = None
        if spec.get("geoms") is not None:
            geoms = [[Geom(GeomType[g["type"]], g["offset"], g["dims"] if len(g["dims"]) != 1 or g["type"] != "SPHERE" else g["dims"],
                           dev, quat=None if g["quat"] is None else np.array(g["quat"]), radius=g["radius"], name=g["name"])
                      for g in bg] for bg in spec["geoms"]]
        cnames = spec.get("contact_body_names")
        if cnames is not None and list(cnames) == list(spec["body_names"]):
            cnames = None
        self.init(body_names=list(spec["body_names"]), parent_indices=spec["parent_indices"],
                  local_translation=np.array(spec["local_translation"]), local_rotation=np.array(spec["local_rotation"]),
                  joints=joints, geoms=geoms, contact_body_names=cnames)

    # ------------------------------------------------------------------ getters (:498-585, 692-701, 1148)
    def get_body_names(self):
        return self._body_names

    def get_num_bodies(self):
        return len(self._body_names)

    def get_num_contact_bodies(self):
        return len(self._contact_body_names)

    def is_contact_body(self, body_name):
        return body_name in self._contact_body_names

    def get_contact_body_id(self, body_name):
        assert body_name in self._contact_body_names
        return self._name_contact_body_map[body_name]

    def get_contact_body_ids(self):
        return self._contact_body_ids

    def get_contact_body_name(self, body_id):
        return self._contact_body_names[body_id]

    def get_joint(self, j) -> Joint:
        assert j > 0
        return self._joints[j]

    def get_parent_id(self, j):
        return self._parent_indices[j]

    def get_dof_size(self):
        return self._dof_size

    def get_joint_dof_idx(self, j):
        return self.get_joint(j).dof_idx

    def get_joint_dof_dim(self, j):
        return self.get_joint(j).get_dof_dim()

    def get_num_joints(self):
        return len(self._joints)

    def get_num_non_root_joints(self):
        return len(self._joints) - 1

    def get_body_name(self, body_id):
        return self._body_names[body_id]

    def get_body_id(self, body_name):
        assert body_name in self._name_body_map
        return self._name_body_map[body_name]

    def get_joint_id(self, body_name):
        assert body_name in self._name_body_map
        return self._name_body_map[body_name] - 1

    def get_geoms(self, body_id) -> List[Geom]:
        return self._geoms[body_id]

    # ------------------------------------------------------------------ C-ABI descriptor
    def _cstruct(self):
        """parc_char_model_t for the current tree (rebuilt if the tree tensors were replaced)."""
        key = (self._local_translation.data_ptr(), self._local_translation._version,
               self._local_rotation.data_ptr(), self._local_rotation._version)
        if self._cm is not None and self._cm_key == key:
            return self._cm
        J = self.get_num_joints()
        assert J <= _lib.MAX_BODIES, "at most %d bodies are supported" % _lib.MAX_BODIES
        assert self._dof_size <= _lib.MAX_DOFS
        cm = _lib.CharModelStruct()
        cm.num_bodies = J
        cm.dof_size = self._dof_size
        parents = self._parent_indices.tolist()
        lt = self._local_translation.detach().cpu().tolist()
        lr = self._local_rotation.detach().cpu().tolist()
        for j in range(J):
            cm.parent[j] = int(parents[j])
            jt = self._joints[j]
            cm.joint_type[j] = jt.joint_type.value
            cm.dof_idx[j] = max(jt.dof_idx, 0)
            for k in range(3):
                cm.local_trans[j][k] = lt[j][k]
            for k in range(4):
                cm.local_rot[j][k] = lr[j][k]
            ax = [0.0, 0.0, 0.0] if jt.axis is None else jt.axis.detach().cpu().tolist()
            for k in range(3):
                cm.axis[j][k] = ax[k]
        self._cm, self._cm_key = cm, key
        return cm

    # ------------------------------------------------------------------ hot methods
    def forward_kinematics(self, root_pos, root_rot, joint_rot):
        """root_pos [...,3], root_rot [...,4], joint_rot [...,J-1,4] -> body_pos [...,J,3], body_rot [...,J,4]."""
        return _FKFunction.apply(self, root_pos, root_rot, joint_rot)

    def dof_to_rot(self, dof):
        """dof [...,D] -> joint_rot [...,J-1,4] (reference :586-599)."""
        J = self.get_num_joints()
        assert dof.shape[-1] == self._dof_size, "dof has wrong trailing dim"
        lead = dof.shape[:-1]
        d = dof.reshape(-1, self._dof_size).contiguous().float()
        n = d.shape[0]
        rot = torch.empty((n, J - 1, 4), dtype=torch.float32, device=d.device)
        _lib.check(_lib.lib().parc_b200_dof_to_rot(self._cstruct(), _lib.ptr(d), n, _lib.ptr(rot), _lib.stream_ptr(d.device)))
        return rot.reshape(*lead, J - 1, 4)

    def rot_to_dof(self, rot):
        """joint_rot [...,J-1,4] -> dof [...,D] (reference :601-615)."""
        J = self.get_num_joints()
        assert rot.shape[-2] == J - 1 and rot.shape[-1] == 4
        lead = rot.shape[:-2]
        r = rot.reshape(-1, J - 1, 4).contiguous().float()
        n = r.shape[0]
        dof = torch.zeros((n, self._dof_size), dtype=torch.float32, device=r.device)
        _lib.check(_lib.lib().parc_b200_rot_to_dof(self._cstruct(), _lib.ptr(r), n, _lib.ptr(dof), _lib.stream_ptr(r.device)))
        return dof.reshape(*lead, self._dof_size)

    def frame_velocities(self, root_pos, root_rot, joint_rot, fps, dt):
        """Finite-difference velocities of ONE clip [nf,...] (motion_lib.py:318-327 and
        compute_frame_dof_vel, kin_char_model.py:651-690) -> root_vel, root_ang_vel, dof_vel."""
        nf = root_pos.shape[0]
        rp = root_pos.contiguous().float()
        rr = root_rot.contiguous().float()
        jr = joint_rot.contiguous().float()
        root_vel = torch.empty((nf, 3), dtype=torch.float32, device=rp.device)
        root_ang_vel = torch.empty((nf, 3), dtype=torch.float32, device=rp.device)
        dof_vel = torch.zeros((nf, self._dof_size), dtype=torch.float32, device=rp.device)
        _lib.check(_lib.lib().parc_b200_frame_velocities(
            self._cstruct(), _lib.ptr(rp), _lib.ptr(rr), _lib.ptr(jr), nf, float(fps), float(dt),
            _lib.ptr(root_vel), _lib.ptr(root_ang_vel), _lib.ptr(dof_vel), _lib.stream_ptr(rp.device)))
        return root_vel, root_ang_vel, dof_vel

    def library_velocities(self, root_pos, root_rot, joint_rot, clip_start, clip_frames, clip_fps, clip_dt):
        """frame_velocities for every clip of a concatenated library [F,...] in ONE launch (clip c owns rows
        [clip_start[c], clip_start[c] + clip_frames[c]) and has fps clip_fps[c], dt clip_dt[c]); bit-identical to
        the per-clip calls (the per-file loop of motion_lib.py:285-372)."""
        F = root_pos.shape[0]
        dev = root_pos.device
        rp, rr, jr = root_pos.contiguous().float(), root_rot.contiguous().float(), joint_rot.contiguous().float()
        cs = clip_start.to(device=dev, dtype=torch.int64).contiguous()
        cf = clip_frames.to(device=dev, dtype=torch.int64).contiguous()
        fps = clip_fps.to(device=dev, dtype=torch.float32).contiguous()
        dt = clip_dt.to(device=dev, dtype=torch.float32).contiguous()
        assert cs.shape == cf.shape == fps.shape == dt.shape and int(cf.sum()) == F
        root_vel = torch.empty((F, 3), dtype=torch.float32, device=dev)
        root_ang_vel = torch.empty((F, 3), dtype=torch.float32, device=dev)
        dof_vel = torch.zeros((F, self._dof_size), dtype=torch.float32, device=dev)
        _lib.check(_lib.lib().parc_b200_library_velocities(
            self._cstruct(), _lib.ptr(rp), _lib.ptr(rr), _lib.ptr(jr), F, int(cs.shape[0]), _lib.ptr(cs), _lib.ptr(cf), _lib.ptr(fps),
            _lib.ptr(dt), _lib.ptr(root_vel), _lib.ptr(root_ang_vel), _lib.ptr(dof_vel), _lib.stream_ptr(dev)))
        return root_vel, root_ang_vel, dof_vel

    def compute_frame_dof_vel(self, joint_rot, dt):
        """[T,J-1,4] -> [T,D] (reference :651-659); only the single-clip shape is kernel-backed."""
        assert joint_rot.dim() == 3, "compute_frame_dof_vel: expected [T, J-1, 4]"
        T = joint_rot.shape[0]
        dummy_p = torch.zeros((T, 3), dtype=torch.float32, device=joint_rot.device)
        dummy_r = torch.zeros((T, 4), dtype=torch.float32, device=joint_rot.device)
        dummy_r[:, 3] = 1.0
        return self.frame_velocities(dummy_p, dummy_r, joint_rot, 1.0 / dt, dt)[2]
