"""MotionLib -- motion clip library evaluated by sm_100a kernels.

Host-side mirror of the reference class (PARC/anim/motion_lib.py:24-516): same constructors
(`MotionLib(...)`, `from_file`, `from_frames`), same public methods and the same attribute surface
(`_frame_root_pos`, `_motion_lengths`, `_motion_weights`, `_terrains`, ...). The flat frame
tensors are additionally repacked on the device into 16-byte-aligned frame records
(include/parc_b200.h, parc_motion_lib_t) which is what the kernels read.
"""
import copy
import enum
import os

import torch
import yaml

from . import _lib
from .file_io import MotionFrames, load_ms_file, HF_MASK_INDS_KEY
from .terrain import SubTerrain


class LoopMode(enum.Enum):
    CLAMP = 0
    WRAP = 1


def extract_pose_data(frame):
    return frame[..., 0:3], frame[..., 3:6], frame[..., 6:]


def resolve_path(path):
    """`$DATA_DIR` substitution of the reference's path_loader (PARC/util/path_loader.py:34-71),
    driven by the PARC_DATA_DIR environment variable instead of a user yaml."""
    path = str(path)
    if "$DATA_DIR" in path:
        data_dir = os.environ.get("PARC_DATA_DIR")
        assert data_dir is not None, "set PARC_DATA_DIR to resolve " + path
        path = path.replace("$DATA_DIR", data_dir)
    return path


def _quat_normalize(q):
    """torch_util.quat_normalize (PARC/util/torch_util.py:469-472); load-time only."""
    q = torch.where(q[..., 3:] < 0, -q, q)
    return q / q.norm(p=2, dim=-1).clamp(min=1e-9).unsqueeze(-1)


class MotionLib:
    def __init__(self, char_model=None, device=None, contact_info=False):
        if char_model is None:
            raise ValueError("kin_char_model must be provided")
        if device is None:
            raise ValueError("device must be provided")
        self._device = torch.device(device)
        self._kin_char_model = char_model
        self._contact_info = contact_info

    # ------------------------------------------------------------------ constructors
    @classmethod
    def from_file(cls, motion_file, char_model, device, contact_info=False):
        inst = cls(char_model, device, contact_info=contact_info)
        inst._load_motion_file(resolve_path(motion_file))
        return inst

    @classmethod
    def from_frames(cls, frames: MotionFrames, char_model, device, loop_mode, fps: float, contact_info=False):
        inst = cls(char_model, device, contact_info=contact_info)
        inst._load_motion_frames(frames, loop_mode, fps)
        return inst

    @classmethod
    def from_clips(cls, clips, char_model, device, contact_info=False, weights=None, names=None, verbose=False):
        """Builds a library from in-memory clips: a list of dicts with keys root_pos [nf,3],
        root_rot [nf,4], joint_rot [nf,J-1,4], contacts [nf,C] or None, fps, loop_mode
        (LoopMode / int / "CLAMP"), terrain (SubTerrain or None). Same math as `from_file`."""
        inst = cls(char_model, device, contact_info=contact_info)
        inst._build(clips, weights, names, verbose=verbose)
        return inst

    # ------------------------------------------------------------------ loading
    def _fetch_motion_files(self, motion_file):
        if os.path.splitext(motion_file)[1] == ".yaml":
            with open(motion_file, "r") as f:
                cfg = yaml.safe_load(f)
            files, weights = [], []
            for entry in cfg["motions"]:
                assert entry["weight"] >= 0
                files.append(resolve_path(entry["file"]))
                weights.append(entry["weight"])
            return files, weights
        return [motion_file], [1.0]

    def _load_motion_file(self, motion_file):
        files, weights = self._fetch_motion_files(motion_file)
        clips, names, mask_inds = [], [], []
        for f in files:
            data = load_ms_file(f, device=self._device)
            md = data.motion_data
            name = os.path.basename(os.path.splitext(f)[0])
            assert name not in names, name + " is a repeat. full path: " + f
            names.append(name)
            terrain = None
            if data.terrain_data is not None:
                terrain = SubTerrain.from_ms_terrain_data(data.terrain_data, device=self._device)
            clips.append(dict(root_pos=md.root_pos, root_rot=md.root_rot, joint_rot=md.joint_rot,
                              contacts=md.body_contacts, fps=md.fps, loop_mode=LoopMode[md.loop_mode].value,
                              terrain=terrain))
            misc = data.misc_data
            mask_inds.append(misc[HF_MASK_INDS_KEY] if (misc is not None and HF_MASK_INDS_KEY in misc) else None)
        self._motion_files = files
        self._build(clips, weights, names)
        self._hf_mask_inds = mask_inds
        print("Loaded {:d} motions with a total length of {:.3f}s.".format(self.num_motions(), self.get_total_length()))

    def _build(self, clips, weights=None, names=None, verbose=False):
        dev = self._device
        kcm = self._kin_char_model
        M = len(clips)
        assert M > 0
        J = kcm.get_num_joints()
        fps_l, dt_l, nf_l, len_l, loop_l = [], [], [], [], []
        rp_l, rr_l, jr_l, c_l, terrains = [], [], [], [], []
        for c in clips:
            root_pos, root_rot, joint_rot = c["root_pos"], c["root_rot"], c["joint_rot"]
            nf = root_pos.shape[0]
            assert nf == root_rot.shape[0] == joint_rot.shape[0]
            fps = c["fps"]
            lm = c["loop_mode"]
            lm = lm.value if isinstance(lm, LoopMode) else (LoopMode[lm].value if isinstance(lm, str) else int(lm))
            fps_l.append(fps); dt_l.append(1.0 / fps); nf_l.append(nf)
            len_l.append(1.0 / fps * (nf - 1))                       # motion_lib.py:313
            loop_l.append(lm)
            rp_l.append(root_pos); rr_l.append(root_rot); jr_l.append(joint_rot)
            if self._contact_info:
                con = c.get("contacts")
                c_l.append(con if con is not None else torch.zeros((nf, J), dtype=torch.float32, device=root_pos.device))
            terrains.append(c.get("terrain"))
        f32 = torch.float32
        # one concatenation + one upload per tensor kind instead of one per clip
        frames = dict(root_pos=torch.cat([x.to(dtype=f32) for x in rp_l], dim=0).to(dev),
                      root_rot=torch.cat([x.to(dtype=f32) for x in rr_l], dim=0).to(dev),
                      joint_rot=torch.cat([x.to(dtype=f32) for x in jr_l], dim=0).to(dev),
                      contacts=torch.cat([x.to(dtype=f32) for x in c_l], dim=0).to(dev) if self._contact_info else None)
        self._build_from_concatenated(frames, nf_l, fps_l, dt_l, len_l, loop_l, weights, names, terrains)

    def _build_from_concatenated(self, frames, nf_l, fps_l, dt_l, len_l, loop_l, weights, names, terrains, hf_mask_inds=None):
        """Library tensors from the clips' frames already concatenated on the device: per-clip tables, the first/last
        frame deltas, and the finite-difference velocities of ALL clips in one launch (motion_lib.py:313-327)."""
        dev = self._device
        M = len(nf_l)
        if weights is None:
            weights = [1.0] * M
        self._motion_weights = torch.tensor(weights, dtype=torch.float32, device=dev)
        self._motion_weights /= self._motion_weights.sum()
        self._motion_fps = torch.tensor(fps_l, dtype=torch.float32, device=dev)
        self._motion_dt = torch.tensor(dt_l, dtype=torch.float32, device=dev)
        self._motion_num_frames = torch.tensor(nf_l, dtype=torch.long, device=dev)
        self._motion_lengths = torch.tensor(len_l, dtype=torch.float32, device=dev)
        self._motion_loop_modes = torch.tensor(loop_l, dtype=torch.int, device=dev)
        self._frame_root_pos = frames["root_pos"]
        self._frame_root_rot = frames["root_rot"]
        self._frame_joint_rot = frames["joint_rot"]
        self._frame_contacts = frames["contacts"] if self._contact_info else None
        start = torch.cumsum(self._motion_num_frames, 0) - self._motion_num_frames
        delta = self._frame_root_pos[start + self._motion_num_frames - 1] - self._frame_root_pos[start]   # motion_lib.py:315-316
        delta[..., -1] = 0.0
        self._motion_root_pos_delta = delta
        self._frame_root_vel, self._frame_root_ang_vel, self._frame_dof_vel = self._kin_char_model.library_velocities(
            self._frame_root_pos, self._frame_root_rot, self._frame_joint_rot, start, self._motion_num_frames, self._motion_fps,
            self._motion_dt)
        self._terrains = terrains
        self._motion_names = names if names is not None else ["motion_%d" % i for i in range(M)]
        self._hf_mask_inds = hf_mask_inds if hf_mask_inds is not None else [None] * M
        self._finish()

    def _finish(self):
        M = self.num_motions()
        self._motion_ids = torch.arange(M, dtype=torch.long, device=self._device)
        shifted = self._motion_num_frames.roll(1)                     # motion_lib.py:394-396
        shifted[0] = 0
        self._motion_start_idx = shifted.cumsum(0)
        self.repack()

    def _load_motion_frames(self, motion_frames, loop_mode, fps):
        """Reference: motion_lib.py:132-247. NOTE the reference passes `fps` where
        compute_frame_dof_vel expects `dt` (:222 vs kin_char_model.py:651); reproduced as is so
        that results are identical (dof_vel = exp_map / fps)."""
        if not isinstance(motion_frames, MotionFrames):
            raise TypeError("motion_frames must be a MotionFrames instance")
        frames = motion_frames.get_copy(self._device)
        if frames.root_pos is None or frames.root_rot is None or frames.joint_rot is None:
            raise ValueError("MotionFrames must contain root_pos, root_rot, and joint_rot tensors")
        root_pos, root_rot, joint_rot = frames.root_pos, frames.root_rot, frames.joint_rot
        if root_pos.dim() == 2:
            root_pos = root_pos.unsqueeze(0)
        if root_rot.dim() == 2:
            root_rot = root_rot.unsqueeze(0)
        if joint_rot.dim() == 3:
            joint_rot = joint_rot.unsqueeze(0)
        if root_pos.dim() != 3 or root_rot.dim() != 3 or joint_rot.dim() != 4:
            raise ValueError("Unexpected MotionFrames tensor shapes")
        M, nf = root_pos.shape[0], root_pos.shape[1]
        if root_rot.shape[0] != M or root_rot.shape[1] != nf:
            raise ValueError("root_rot must have the same batch and frame dimensions as root_pos")
        if joint_rot.shape[0] != M or joint_rot.shape[1] != nf:
            raise ValueError("joint_rot must have the same batch and frame dimensions as root_pos")
        dev = self._device
        kcm = self._kin_char_model
        root_pos = root_pos.to(dtype=torch.float32)
        root_rot = _quat_normalize(root_rot.to(dtype=torch.float32))
        joint_rot = _quat_normalize(joint_rot.to(dtype=torch.float32))

        contacts = None
        if self._contact_info:
            contacts = frames.contacts
            if contacts is None:
                contacts = torch.zeros((nf, kcm.get_num_bodies()), dtype=torch.float32, device=dev)
            if contacts.dim() == 2:
                contacts = contacts.unsqueeze(0)
            if contacts.shape[0] != M or contacts.shape[1] != nf:
                raise ValueError("contacts must have the same batch and frame dimensions as root_pos")
            contacts = contacts.to(device=dev, dtype=torch.float32)

        if isinstance(loop_mode, LoopMode):
            loop_tensor = torch.full((M,), loop_mode.value, dtype=torch.int, device=dev)
        elif torch.is_tensor(loop_mode):
            loop_tensor = (torch.full((M,), int(loop_mode.item()), dtype=torch.int, device=dev)
                           if loop_mode.numel() == 1 else loop_mode.to(device=dev, dtype=torch.int))
        elif isinstance(loop_mode, (list, tuple)):
            loop_tensor = torch.tensor(loop_mode, dtype=torch.int, device=dev)
        else:
            loop_tensor = torch.full((M,), int(loop_mode), dtype=torch.int, device=dev)
        if loop_tensor.shape[0] != M:
            raise ValueError("loop_mode must provide one value per motion")

        rv, rav, dv = [], [], []
        for m in range(M):
            a, b, _ = kcm.frame_velocities(root_pos[m], root_rot[m], joint_rot[m], fps, 1.0 / fps)
            # the reference hands `fps` to compute_frame_dof_vel as if it were dt (:222)
            _, _, d = kcm.frame_velocities(root_pos[m], root_rot[m], joint_rot[m], fps, float(fps))
            rv.append(a); rav.append(b); dv.append(d)

        self._motion_fps = torch.full((M,), fps, dtype=torch.float32, device=dev)
        self._motion_dt = torch.full((M,), 1.0 / fps, dtype=torch.float32, device=dev)
        self._motion_num_frames = torch.full((M,), nf, dtype=torch.long, device=dev)
        self._motion_lengths = torch.full((M,), 1.0 / fps * (nf - 1), dtype=torch.float32, device=dev)
        self._motion_loop_modes = loop_tensor
        delta = root_pos[:, -1] - root_pos[:, 0]
        delta[..., -1] = 0.0
        self._motion_root_pos_delta = delta
        self._frame_root_pos = root_pos.reshape(-1, 3)
        self._frame_root_rot = root_rot.reshape(-1, 4)
        self._frame_joint_rot = joint_rot.reshape(-1, joint_rot.shape[-2], 4)
        self._frame_root_vel = torch.cat(rv, dim=0)
        self._frame_root_ang_vel = torch.cat(rav, dim=0)
        self._frame_dof_vel = torch.cat(dv, dim=0)
        self._frame_contacts = contacts.reshape(-1, contacts.shape[-1]) if contacts is not None else None
        self._motion_weights = torch.ones((M,), dtype=torch.float32, device=dev)
        self._terrains = [None]
        self._motion_names = [""]
        self._hf_mask_inds = [None]
        self._finish()

    # ------------------------------------------------------------------ packed on-disk library
    PACK_MAGIC = b"PARCPACK1\n"

    def save_packed(self, path):
        """One flat file for the whole library instead of one pickle per clip (the reference's loader walks a yaml of
        .pkl files, motion_lib.py:255-372): magic | u64 header length | JSON header | 64-byte-aligned raw arrays. It holds
        the SOURCE data only (poses, contacts, per-clip tables, terrains, hf_mask_inds); velocities and the device
        records are rebuilt by `from_packed` in two launches."""
        import json
        import numpy as np
        arrays = {}

        def put(name, t):
            arrays[name] = np.ascontiguousarray(t.detach().cpu().numpy() if torch.is_tensor(t) else np.asarray(t))

        put("root_pos", self._frame_root_pos); put("root_rot", self._frame_root_rot); put("joint_rot", self._frame_joint_rot)
        if self._frame_contacts is not None:
            put("contacts", self._frame_contacts)
        put("num_frames", self._motion_num_frames); put("loop_modes", self._motion_loop_modes)
        put("fps", self._motion_fps.double()); put("weights", self._motion_weights)
        M = self.num_motions()
        has_t = np.array([t is not None for t in self._terrains] + [False] * (M - len(self._terrains)), dtype=np.uint8)
        put("has_terrain", has_t)
        ts = [t for t in self._terrains if t is not None]
        if ts:
            put("terrain_dims", torch.stack([torch.tensor(t.hf.shape, dtype=torch.int64) for t in ts]))
            put("terrain_min_point", torch.stack([t.min_point.float().cpu() for t in ts]))
            put("terrain_dxdy", torch.stack([t.dxdy.float().cpu() for t in ts]))
            put("terrain_hf", torch.cat([t.hf.float().cpu().reshape(-1) for t in ts]))
            put("terrain_hf_maxmin", torch.cat([t.hf_maxmin.float().cpu().reshape(-1) for t in ts]))
        masks = list(self._hf_mask_inds) + [None] * (M - len(self._hf_mask_inds))
        put("has_mask_inds", np.array([m is not None for m in masks], dtype=np.uint8))
        flat, offs = [], [0]
        for m in masks:
            if m is None:
                continue
            for fr in m:                                     # one [k,2] index list per frame of the clip
                fr = torch.as_tensor(fr, dtype=torch.int64).reshape(-1, 2).cpu()
                flat.append(fr); offs.append(offs[-1] + fr.shape[0])
        if flat:
            put("mask_inds", torch.cat(flat, dim=0)); put("mask_offsets", np.asarray(offs, dtype=np.int64))
        header = {"version": 1, "contact_info": bool(self._contact_info), "names": list(self._motion_names), "arrays": {}}
        off = 0
        for k, a in arrays.items():
            off = (off + 63) & ~63
            header["arrays"][k] = {"dtype": str(a.dtype), "shape": list(a.shape), "offset": off}
            off += a.nbytes
        hb = json.dumps(header).encode()
        with open(path, "wb") as f:
            f.write(self.PACK_MAGIC); f.write(len(hb).to_bytes(8, "little")); f.write(hb)
            base = f.tell()
            base_al = (base + 63) & ~63
            f.write(b"\0" * (base_al - base))
            for k, a in arrays.items():
                f.seek(base_al + header["arrays"][k]["offset"])
                f.write(a.tobytes())
        return path

    @classmethod
    def from_packed(cls, path, char_model, device, contact_info=None):
        """Loads a `save_packed` file: one read, one upload per array, velocities of all clips in one launch, records in
        one more. Equal, tensor for tensor, to the library the file was saved from."""
        import json
        import numpy as np
        raw = np.fromfile(path, dtype=np.uint8)
        magic = cls.PACK_MAGIC
        if raw[:len(magic)].tobytes() != magic:
            raise ValueError("%s is not a packed parc_b200 motion library" % path)
        hl = int.from_bytes(raw[len(magic):len(magic) + 8].tobytes(), "little")
        header = json.loads(raw[len(magic) + 8:len(magic) + 8 + hl].tobytes().decode())
        base = (len(magic) + 8 + hl + 63) & ~63

        def get(name):
            if name not in header["arrays"]:
                return None
            d = header["arrays"][name]
            n = int(np.prod(d["shape"])) if d["shape"] else 1
            a = np.frombuffer(raw, dtype=np.dtype(d["dtype"]), count=n, offset=base + d["offset"]).reshape(d["shape"])
            return torch.from_numpy(a.copy())

        ci = header["contact_info"] if contact_info is None else contact_info
        inst = cls(char_model, device, contact_info=ci)
        dev = inst._device
        nf = get("num_frames")
        M = int(nf.shape[0])
        fps = [float(x) for x in get("fps")]
        fps = [int(x) if float(x).is_integer() else x for x in fps]
        frames = dict(root_pos=get("root_pos").to(dev), root_rot=get("root_rot").to(dev), joint_rot=get("joint_rot").to(dev),
                      contacts=None)
        if ci:
            con = get("contacts")
            frames["contacts"] = (con if con is not None else torch.zeros(frames["root_pos"].shape[0], char_model.get_num_joints())).to(dev)
        terrains = [None] * M
        has_t = get("has_terrain")
        if has_t is not None and bool(has_t.any()):
            dims, mp, dd = get("terrain_dims"), get("terrain_min_point"), get("terrain_dxdy")
            hf_all, mm_all = get("terrain_hf").to(dev), get("terrain_hf_maxmin").to(dev)
            o, k = 0, 0
            for m in range(M):
                if not has_t[m]:
                    continue
                X, Y = int(dims[k, 0]), int(dims[k, 1])
                t = SubTerrain(x_dim=X, y_dim=Y, dx=float(dd[k, 0]), dy=float(dd[k, 1]), min_x=float(mp[k, 0]), min_y=float(mp[k, 1]),
                               device=dev)
                t.min_point, t.dxdy = mp[k].to(dev), dd[k].to(dev)
                t.hf = hf_all[o:o + X * Y].reshape(X, Y).clone()
                t.hf_maxmin = mm_all[2 * o:2 * (o + X * Y)].reshape(X, Y, 2).clone()
                terrains[m] = t
                o += X * Y; k += 1
        masks = [None] * M
        has_m = get("has_mask_inds")
        if has_m is not None and bool(has_m.any()):
            inds, offs = get("mask_inds").to(dev), get("mask_offsets")
            f0 = 0
            for m in range(M):
                if not has_m[m]:
                    continue
                n_m = int(nf[m])
                masks[m] = [inds[int(offs[f0 + t]):int(offs[f0 + t + 1])] for t in range(n_m)]
                f0 += n_m
        nf_l = [int(x) for x in nf]
        inst._build_from_concatenated(frames, nf_l, fps, [1.0 / f for f in fps], [1.0 / f * (n - 1) for f, n in zip(fps, nf_l)],
                                      [int(x) for x in get("loop_modes")], [float(x) for x in get("weights")], header["names"], terrains,
                                      hf_mask_inds=masks)
        inst._motion_weights = get("weights").to(dev)      # already normalised: keep the saved bits
        return inst

    # ------------------------------------------------------------------ device records
    def repack(self):
        """(Re)build the packed frame records from the flat `_frame_*` tensors."""
        kcm = self._kin_char_model
        J, D = kcm.get_num_joints(), kcm.get_dof_size()
        F = int(self._frame_root_pos.shape[0])
        M = self.num_motions()
        C = int(self._frame_contacts.shape[-1]) if self._frame_contacts is not None else 0
        pair_off = (4 * J + 4 + C + 1) & ~1                        # csrc/parc_lanes.cuh: interp_pair_off
        irec = (pair_off + 2 * J + 3) & ~3
        vrec = (6 + D + 3) & ~3
        dev = self._device
        self._blob_interp = torch.empty((F, irec), dtype=torch.float32, device=dev)
        self._blob_vel = torch.empty((F, vrec), dtype=torch.float32, device=dev)
        self._blob_meta = torch.empty((M, 8), dtype=torch.int32, device=dev)
        f32, i64, i32 = torch.float32, torch.int64, torch.int32
        cp = _lib.cptr
        _lib.check(_lib.lib().parc_b200_motionlib_pack(
            kcm._cstruct(), F, M, C,
            cp(self._frame_root_pos.contiguous(), f32), cp(self._frame_root_rot.contiguous(), f32),
            cp(self._frame_root_vel.contiguous(), f32), cp(self._frame_root_ang_vel.contiguous(), f32),
            cp(self._frame_joint_rot.contiguous(), f32), cp(self._frame_dof_vel.contiguous(), f32),
            cp(self._frame_contacts.contiguous(), f32) if C > 0 else None,
            cp(self._motion_num_frames.contiguous(), i64), cp(self._motion_start_idx.contiguous(), i64),
            cp(self._motion_lengths.contiguous(), f32), cp(self._motion_loop_modes.contiguous(), i32),
            cp(self._motion_root_pos_delta.contiguous(), f32),
            _lib.ptr(self._blob_interp), _lib.ptr(self._blob_vel), _lib.ptr(self._blob_meta),
            _lib.stream_ptr(dev)))
        s = _lib.MotionLibStruct()
        s.interp, s.vel, s.meta = _lib.ptr(self._blob_interp), _lib.ptr(self._blob_vel), _lib.ptr(self._blob_meta)
        s.total_frames, s.num_motions, s.irec, s.vrec, s.num_contacts = F, M, irec, vrec, C
        self._cs = s

    def cstruct(self):
        return self._cs

    # ------------------------------------------------------------------ reference API
    def num_motions(self):
        return self._motion_lengths.shape[0]

    def get_total_length(self):
        return torch.sum(self._motion_lengths).item()

    def sample_motions(self, n, motion_weights=None):
        """torch.multinomial draw, as the reference (:56-60); the RNG stream is torch's."""
        if motion_weights is None:
            motion_weights = self._motion_weights
        return torch.multinomial(motion_weights, num_samples=n, replacement=True)

    def sample_time(self, motion_ids, truncate_time=None):
        phase = torch.rand(motion_ids.shape, device=self._device)
        motion_len = self._motion_lengths[motion_ids]
        if truncate_time is not None:
            assert truncate_time >= 0.0
            motion_len -= truncate_time
        return phase * motion_len

    def get_motion_length(self, motion_ids):
        return self._motion_lengths[motion_ids]

    def get_motion_num_frames(self, motion_ids):
        return self._motion_num_frames[motion_ids]

    def get_motion_loop_mode(self, motion_ids):
        return self._motion_loop_modes[motion_ids]

    def get_motion_loop_mode_enum(self, motion_id):
        return LoopMode(self._motion_loop_modes[motion_id].item())

    def get_motion_fps(self, motion_id):
        return self._motion_fps[motion_id].item()

    def get_motion_names(self):
        assert hasattr(self, "_motion_names")
        return self._motion_names

    def _eval(self, motion_ids, motion_times, want_idx=False):
        kcm = self._kin_char_model
        J, D = kcm.get_num_joints(), kcm.get_dof_size()
        shape = motion_ids.shape
        ids = motion_ids.reshape(-1).contiguous().to(torch.int64)
        times = motion_times.reshape(-1).contiguous().to(torch.float32)
        assert ids.shape[0] == times.shape[0]
        n = ids.shape[0]
        dev = ids.device
        C = self._cs.num_contacts
        e = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
        root_pos, root_rot, root_vel, root_ang_vel = e(n, 3), e(n, 4), e(n, 3), e(n, 3)
        joint_rot, dof_vel = e(n, J - 1, 4), e(n, D)
        contacts = e(n, C) if self._contact_info else None
        idx0 = idx1 = blend = None
        if want_idx:
            idx0 = torch.empty((n,), dtype=torch.int64, device=dev)
            idx1 = torch.empty((n,), dtype=torch.int64, device=dev)
            blend = e(n)
        p = _lib.ptr
        _lib.check(_lib.lib().parc_b200_calc_motion_frame(
            kcm._cstruct(), self._cs, p(ids), p(times), n, p(root_pos), p(root_rot), p(root_vel), p(root_ang_vel),
            p(joint_rot), p(dof_vel), p(contacts), p(idx0), p(idx1), p(blend), _lib.stream_ptr(dev)))
        r = lambda t, *tail: t.reshape(*shape, *tail)
        ret = [r(root_pos, 3), r(root_rot, 4), r(root_vel, 3), r(root_ang_vel, 3), r(joint_rot, J - 1, 4), r(dof_vel, D)]
        if self._contact_info:
            ret.append(r(contacts, C))
        if want_idx:
            return tuple(ret), (idx0.reshape(shape), idx1.reshape(shape), blend.reshape(shape))
        return tuple(ret)

    def calc_motion_frame(self, motion_ids, motion_times):
        """-> (root_pos, root_rot, root_vel, root_ang_vel, joint_rot, dof_vel[, contacts]); fresh tensors (:94-126)."""
        return self._eval(motion_ids, motion_times)

    def _calc_frame_blend(self, motion_ids, times):
        """-> (frame_idx0, frame_idx1, blend) exactly as the reference (:425-438)."""
        return self._eval(motion_ids, times, want_idx=True)[1]

    def calc_motion_phase(self, motion_ids, times):
        """Reference :88-92 / calc_phase :519-530 (cold helper: plain tensor expression)."""
        motion_len = self._motion_lengths[motion_ids]
        loop_mode = self._motion_loop_modes[motion_ids]
        phase = times / motion_len
        wrap = loop_mode == LoopMode.WRAP.value
        phase = torch.where(wrap, phase - torch.floor(phase), phase)
        return torch.clip(phase, 0.0, 1.0)

    def joint_rot_to_dof(self, joint_rot):
        return self._kin_char_model.rot_to_dof(joint_rot)

    def calc_motion_frame_dofs(self, motion_ids, motion_times):
        """Reference :459-471: [root_pos, exp_map(root_rot), joint dofs(, contacts)]."""
        out = self.calc_motion_frame(motion_ids, motion_times)
        root_pos, root_rot, joint_rot = out[0], out[1], out[4]
        # exp map of the root rotation through the same kernel: treat it as a 1-joint spherical chain
        root_exp = _root_rot_to_exp_map(root_rot)
        parts = [root_pos, root_exp, self.joint_rot_to_dof(joint_rot)]
        if self._contact_info:
            parts.append(out[6])
        return torch.cat(parts, dim=-1)

    @property
    def _motion_frames(self):
        return torch.cat([self._frame_root_pos, _root_rot_to_exp_map(self._frame_root_rot),
                          self.joint_rot_to_dof(self._frame_joint_rot)], dim=-1)

    def get_frames_for_id(self, id, compute_fk=False):
        nf = self._motion_num_frames[id]
        s = self._motion_start_idx[id]
        sl = slice(s, s + nf)
        root_pos, root_rot, joint_rot = self._frame_root_pos[sl], self._frame_root_rot[sl], self._frame_joint_rot[sl]
        body_pos = body_rot = None
        if compute_fk:
            body_pos, body_rot = self._kin_char_model.forward_kinematics(root_pos, root_rot, joint_rot)
        contacts = self._frame_contacts[sl] if self._contact_info else None
        return MotionFrames(root_pos=root_pos, root_rot=root_rot, joint_rot=joint_rot, body_pos=body_pos,
                            body_rot=body_rot, contacts=contacts)

    def clone(self, device):
        new = copy.copy(self)
        new._device = torch.device(device)
        for attr, value in vars(self).items():
            if isinstance(value, torch.Tensor):
                setattr(new, attr, value.to(device).clone())
        new._kin_char_model = self._kin_char_model.get_copy(device)
        new.repack()
        return new


_SPH_MODEL = {}


def _root_rot_to_exp_map(q):
    """quat_to_exp_map (PARC/util/torch_util.py:371-376) via the rot_to_dof kernel on a 2-body
    chain whose only joint is spherical."""
    from .kin_char_model import KinCharModel, Joint, JointType
    dev = q.device
    key = str(dev)
    if key not in _SPH_MODEL:
        m = KinCharModel(dev)
        m.init(["a", "b"], [-1, 0], [[0, 0, 0], [0, 0, 0]], [[0, 0, 0, 1], [0, 0, 0, 1]],
               [Joint("root", JointType.ROOT, None), Joint("j", JointType.SPHERICAL, None)])
        _SPH_MODEL[key] = m
    return _SPH_MODEL[key].rot_to_dof(q.unsqueeze(-2))
