"""bench.py -- env-steps/s of the fused obs + reward + termination + RSI-reset step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1]): 65,536 envs per GPU, 1,000-clip synthetic library of the repo
humanoid, 32x32 local heightmap observations (obs row = 1,895 floats). A "step" is one pass of
IGEnv._post_physics_step (heights, time, reference motion, observation, reward, termination),
the fail-rate EMA and the reference-state-init reset of the envs that finished. Envs shard by
env index over ranks with no data-path collective ("weak" scaling: 65,536 envs on every GPU);
the only exchange is one all-reduce of reward / termination statistics at the end.

Own arm prints ONE JSON line with value (inputs resident in HBM), e2e (host buffers, H2D/D2H in
the timed region), roofline of the dominant kernel and a cpu_baseline (the CPU oracle port, rank 0,
N=1 only). `--impl reference` times the CPU port alone on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

ENVS_PER_GPU = 65536
NUM_MOTIONS = 1000
CPU_SAMPLE_ENVS = 4096
METRIC = "env-steps/s (obs+reward+RSI) at 65,536 envs/GPU; achieved HBM GB/s vs peak"
UNIT = "env-steps/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", type=str, default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU (default = the BASELINE config)")
    ap.add_argument("--motions", type=int, default=NUM_MOTIONS)
    ap.add_argument("--cpu-envs", type=int, default=CPU_SAMPLE_ENVS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fail-fraction", type=float, default=0.02,
                    help="share of the envs the simulator stand-in pushes off their reference every step (they terminate and are "
                         "reset): 0.02 is the stationary BASELINE configs[1] workload, 0.1 with --envs 262144 is configs[4] (reset-heavy)")
    ap.add_argument("--chunks", type=int, default=6, help="row ranges of the host-fed step (upload of range c+1 overlaps the step of range c)")
    ap.add_argument("--fail-rate-sync", type=int, default=10, help="multi-GPU runs: all-reduce the per-motion fail / pass counts every this many "
                                                                   "steps so that every rank samples RSI motions from the same fail rates (0 = never)")
    ap.add_argument("--no-extra", action="store_true", help="skip the `extra` block (BASELINE configs[2], [3], [4] measured beside the headline)")
    ap.add_argument("--four-tensor-state", action="store_true", help="host-fed step with Isaac Gym's four state tensors instead of one interleaved row per env")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# workload construction (product side: synthetic inputs + CUDA classes, no oracle)
# ------------------------------------------------------------------------------------------------
def build_world(num_motions, seed=1234):
    from parc_b200 import synth
    clips = synth.make_clips(synth.humanoid_spec(), num_motions, seed=seed, wrap_fraction=0.05)
    hf, min_point, dx, dm_offsets = synth.build_terrain_square(clips, padding_cells=1)
    return dict(clips=clips, hf=hf, min_point=min_point, dx=dx, dm_offsets=dm_offsets, ray=synth.grid_ray_points())


def build_gpu_env(world, num_envs, first_env, total_envs, device, interleaved=True):
    """Rank-local env over global env ids [first_env, first_env + num_envs) of a total_envs grid."""
    from parc_b200 import KinCharModel, MotionLib, SubTerrain, synth
    from parc_b200.tracking_env import TrackingEnv, default_env_config
    model = KinCharModel(device)
    model.load_char_file(synth.HUMANOID_JSON)
    pc = [dict(root_pos=c["root_pos"], root_rot=c["root_rot"], joint_rot=c["joint_rot"], contacts=c["contacts"], fps=c["fps"],
               loop_mode=c["loop_mode"], terrain=None) for c in world["clips"]]
    lib = MotionLib.from_clips(pc, model, device, contact_info=True)
    hf = world["hf"]
    terrain = SubTerrain(x_dim=hf.shape[0], y_dim=hf.shape[1], dx=world["dx"], dy=world["dx"],
                         min_x=float(world["min_point"][0]), min_y=float(world["min_point"][1]), device=device)
    terrain.hf = hf.to(device).contiguous()
    offs = synth.env_offsets(total_envs)[first_env:first_env + num_envs]
    env = TrackingEnv(default_env_config(), num_envs, device, model, lib, terrain, world["dm_offsets"], world["ray"],
                      env_offsets=offs, seed=20260 + first_env, interleaved_state=interleaved)
    return env


def init_gpu_state(env, first_env, seed=4321):
    """RSI-reset every env with seeded draws, scatter them over their episodes, and install a
    perturbed 'simulated' state around the reference pose."""
    from parc_b200 import synth
    N, dev = env.get_num_envs(), env._device
    lib = env._mlib
    draws = synth.reset_draws(N, lib.num_motions(), 1, lib._motion_lengths, lib._motion_weights, seed=seed + first_env)
    env.reset(torch.arange(N, device=dev), draws=draws, refresh_obs=False)
    g = torch.Generator().manual_seed(seed + 7 + first_env)
    env._timestep_buf.copy_(torch.randint(0, 100, (N,), generator=g, dtype=torch.int32).to(dev))
    env._time_buf.copy_(env._timestep_buf.float() * np.float32(1.0 / 30.0))
    env.post_physics_step(parts=2, update_fail_rates=False)      # PARC_STEP_REF: reference state at the current time
    nz = synth.sim_noise(N, env._J, env._D, seed=5678 + first_env)
    ref = dict(root_pos=env._ref_root_pos, root_rot=env._ref_root_rot, root_vel=env._ref_root_vel,
               root_ang_vel=env._ref_root_ang_vel, dof_pos=env._ref_dof_pos, dof_vel=env._ref_dof_vel)
    sim = synth.perturb_state(ref, nz, env._kin_char_model.forward_kinematics, env._kin_char_model.dof_to_rot)
    env._char_root_pos.copy_(sim["root_pos"]); env._char_root_rot.copy_(sim["root_rot"])
    env._char_root_vel.copy_(sim["root_vel"]); env._char_root_ang_vel.copy_(sim["root_ang_vel"])
    env._char_dof_pos.copy_(sim["dof_pos"]); env._char_dof_vel.copy_(sim["dof_vel"])
    env._char_rigid_body_pos.copy_(sim["body_pos"]); env._char_contact_forces.copy_(sim["contact_forces"])
    torch.cuda.synchronize(dev)


def algorithmic_bytes(env):
    """N * B_env(P) + B_lib, B_env = 4,940 + 4 P bytes,
    B_lib = min(N * 4,504, 448 F) + 4 X Y."""
    N = env.get_num_envs()
    P = int(env._ray_xy_points.shape[0])
    F = int(env._mlib._frame_root_pos.shape[0])
    X, Y = env._terrain.hf.shape
    b_env = 4940 + 4 * P
    b_lib = min(N * 4504, 448 * F) + 4 * X * Y
    return N * b_env + b_lib, b_env


LAUNCHES_PER_STEP = 6   # device-resident loop: sim_standin, env_step, fail_rate_ema, env_reset_sampled, env_step(reset rows), done-count clear


# ------------------------------------------------------------------------------------------------
# CPU leg: the oracle port (oracle/parc_oracle.py) on the host cores -- checker code used as the
# reported baseline only, never on the product path
# ------------------------------------------------------------------------------------------------
def run_cpu_port(world, num_envs, steps, warmup, threads, fail_fraction=0.02):
    from oracle import parc_oracle as O
    from parc_b200 import synth
    from parc_b200.tracking_env import default_env_config
    torch.set_num_threads(threads)
    spec = synth.humanoid_spec()
    m = O.CharModel(spec)
    lib = O.Library(m, [dict(c, terrain=None) for c in world["clips"]])
    terrain = O.Terrain(world["hf"], world["min_point"], [world["dx"], world["dx"]])
    ec = default_env_config()
    cfg = O.EnvConfig(m, rand_root_pos_offset_scale=ec["rand_root_pos_offset_scale"])
    env = O.TrackingEnv(m, lib, terrain, world["dm_offsets"], cfg, num_envs, synth.env_offsets(num_envs), world["ray"])
    N = num_envs
    draws = synth.reset_draws(N, lib.num_motions(), 1, lib.lengths, lib.motion_weights, seed=4321)
    env.reset(torch.arange(N), *draws)
    g = torch.Generator().manual_seed(4328)
    env.timestep_buf[:] = torch.randint(0, 100, (N,), generator=g, dtype=torch.int32)
    env.time_buf[:] = env.timestep * env.timestep_buf
    env.update_ref_motion()
    ref = dict(root_pos=env.ref_root_pos, root_rot=env.ref_root_rot, root_vel=env.ref_root_vel,
               root_ang_vel=env.ref_root_ang_vel, dof_pos=env.ref_dof_pos, dof_vel=env.ref_dof_vel)
    sim = synth.perturb_state(ref, synth.sim_noise(N, m.J, m.D, seed=5678), m.forward_kinematics, m.dof_to_rot)
    env.char_root_pos[:], env.char_root_rot[:] = sim["root_pos"], sim["root_rot"]
    env.char_root_vel[:], env.char_root_ang_vel[:] = sim["root_vel"], sim["root_ang_vel"]
    env.char_dof_pos[:], env.char_dof_vel[:] = sim["dof_pos"], sim["dof_vel"]
    env.char_rigid_body_pos[:], env.char_contact_forces[:] = sim["body_pos"], sim["contact_forces"]

    noise = synth.standin_noise(N, m.J, m.D, seed=5678, fail_fraction=fail_fraction)
    resets = []

    def one_step(i):
        synth.apply_standin_torch(env, noise, m.J, m.D)      # the same simulator stand-in as the GPU arm
        env.post_physics_step()
        ids = torch.nonzero(env.done_buf != O.DONE_NULL).flatten()
        resets.append(int(ids.numel()))
        if ids.numel() > 0:
            d = synth.reset_draws(ids.numel(), lib.num_motions(), 1, lib.lengths, env.sampling_weights(), seed=1000 + i)
            env.reset(ids, *d)
            env.refresh_obs_after_reset(ids)

    for i in range(warmup):
        one_step(i)
    times = []
    for i in range(steps):
        t0 = time.perf_counter()
        one_step(warmup + i)
        times.append(time.perf_counter() - t0)
    total = float(np.sum(times))
    return {"value": N * steps / total, "median_s": float(np.median(times)), "min_s": float(np.min(times)), "total_s": total,
            "resets_per_step": float(np.mean(resets[warmup:]))}


def physical_cores():
    try:
        import psutil
        return psutil.cpu_count(logical=False) or (os.cpu_count() or 1)
    except Exception:
        return os.cpu_count() or 1


def cpu_baseline_block(world, args, steps, warmup, full_config_sample=True):
    """The CPU arm: torch.set_num_threads swept over {1, physical cores} on a 4,096-env sample of
    the workload, the best median kept; then, memory and time allowing, a short sample at the FULL 65,536 envs per step with
    that thread count, so that the reported value is measured on the same configuration as the GPU arm."""
    logical, phys = os.cpu_count() or 1, physical_cores()
    sweep = []
    for th in sorted({1, phys}):
        st, wu = (max(4, steps // 3), 2) if th == 1 and phys > 1 else (steps, warmup)
        r = run_cpu_port(world, args.cpu_envs, st, wu, th, args.fail_fraction)
        sweep.append({"threads": th, "envs_per_step": args.cpu_envs, "steps": st, "value": args.cpu_envs / r["median_s"], "median_s": r["median_s"],
                      "min_s": r["min_s"], "resets_per_step": r["resets_per_step"]})
    best = max(sweep, key=lambda x: x["value"])
    chosen, same_config = dict(best), False
    if full_config_sample and args.envs >= args.cpu_envs * 2:
        try:
            r = run_cpu_port(world, args.envs, 3, 1, best["threads"], args.fail_fraction)
            chosen = {"threads": best["threads"], "envs_per_step": args.envs, "steps": 3, "value": args.envs / r["median_s"], "median_s": r["median_s"],
                      "min_s": r["min_s"], "resets_per_step": r["resets_per_step"]}
            same_config = args.envs == ENVS_PER_GPU and args.motions == NUM_MOTIONS
        except (RuntimeError, MemoryError) as exc:      # not enough host memory for the 65,536-env torch temporaries
            chosen["full_config_error"] = str(exc)[:120]
    cb = {"value": chosen["value"], "unit": UNIT, "cores": chosen["threads"], "kind": "port", "threads": chosen["threads"],
          "physical_cores": phys, "logical_cores": logical, "envs_per_step": chosen["envs_per_step"], "median_s": chosen["median_s"],
          "min_s": chosen["min_s"], "same_config": same_config, "thread_sweep": sweep,
          "sample": "oracle port (torch CPU eager) of the same %d-clip / 32x32-heightmap workload incl. its resets: %d envs x %d steps at %d threads "
                    "(median %.3f s/step, min %.3f s/step); thread sweep over {1, %d physical cores} at %d envs/step"
                    % (args.motions, chosen["envs_per_step"], chosen["steps"], chosen["threads"], chosen["median_s"], chosen["min_s"], phys, args.cpu_envs)}
    return cb, chosen


def run_reference_arm(args):
    """`--impl reference`: the reference's CPU path for this workload = the oracle port (the
    reference itself is a Python tree under /root/reference that does not travel to the GPU box)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = build_world(args.motions)
    steps, warmup = max(args.steps, 1), max(args.warmup, 1)
    # bounded: ~0.1-0.25 s/step at 4,096 envs -> cap the sample so the run stays within minutes
    steps_cpu, warm_cpu = min(steps, 30), min(warmup, 5)
    cb, ch = cpu_baseline_block(world, args, steps_cpu, warm_cpu)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": ch["steps"],
            "warmup": warm_cpu, "ms_per_step": 1e3 * ch["median_s"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "65,536-env / 1,000-clip / 32x32-heightmap tracking step (same stand-in + step + reset as the b200 "
                                   "arm); the CPU arm's value is measured at %d envs per step" % ch["envs_per_step"],
                       "envs_per_step": ch["envs_per_step"], "clips": args.motions, "ray_points": 1024, "obs_dim": 1895,
                       "same_config": cb["same_config"]},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def device_loop(env, noise, K, W, barrier, max_over_ranks):
    """K control steps with the state resident in HBM, through the repo's API (TrackingEnv.sim_standin + TrackingEnv.step):
    total time by CUDA events over the loop, the dominant kernel by events around each fused-step launch."""
    env._kernel_events = None
    fr = getattr(env, "_fail_rate_sync", None)
    if fr is not None:
        interval, fr.interval = fr.interval, 1       # the warm-up exchanges every step: first-use costs of the all-reduce and of the
    for _ in range(W - 1):                           # closed-form update (kernel loading, NCCL channels) stay out of the timed steps
        env.sim_standin(noise)
        env.step()
    env.sim_standin(noise)
    env.post_physics_step()
    done_per_step = int(env._done_count.item())      # envs that finish (and are reset) in a steady-state step
    env.reset_done()
    if fr is not None:
        fr.interval, fr.steps = interval, 0
    barrier()
    env._kernel_events = []
    t_start, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_start.record()
    host_t0 = time.perf_counter()
    for _ in range(K):
        env.sim_standin(noise)                 # "physics": the character follows its reference with a fixed error
        env.step()                             # fused post-physics pass + fail-rate EMA + device-side RSI reset + obs refresh
    host_enqueue_ms = (time.perf_counter() - host_t0) * 1e3 / K
    t_end.record()
    barrier()
    ev, env._kernel_events = env._kernel_events, None
    ms_total = max_over_ranks(t_start.elapsed_time(t_end))
    kern_ms = max_over_ranks(float(np.mean([a.elapsed_time(b) for a, b in ev])))
    return {"ms_total": ms_total, "kern_ms": kern_ms, "done_per_step": done_per_step, "host_enqueue_ms": host_enqueue_ms}


def roofline_block(env, kern_ms, ms_per_step, traffic=None, warp_inst=None, sm_mhz=None):
    step_bytes, b_env = algorithmic_bytes(env)
    peak, peak_src = measured_peaks()
    achieved = step_bytes / (kern_ms * 1e-3) / 1e9
    out = {"bound": "hbm", "kernel": "env_step_kernel<512, lattice, fast>", "achieved": achieved, "peak": peak, "unit": "GB/s",
           "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": step_bytes,
           "bytes_per_env": b_env, "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms / ms_per_step}
    if warp_inst:
        # the second roofline of this kernel: instruction issue = warp instructions (ncu, profiled launch) / (148 SMs x 4 warp
        # schedulers x SM clock): the time the launch would take if every scheduler issued every cycle
        floor_ms = warp_inst / (148 * 4 * (sm_mhz or 1965.0) * 1e6) * 1e3
        out["issue"] = {"warp_instructions_per_launch": warp_inst, "warp_instructions_per_env": warp_inst / env.get_num_envs(),
                        "floor_ms": floor_ms, "frac": floor_ms / kern_ms}
    return out


def run_gpu_arm(args):
    import torch.distributed as dist
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU fallback on the product path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    distributed = world_size > 1
    if distributed:
        dist.init_process_group(backend="nccl", device_id=dev)
    n_gpus = world_size
    N = args.envs
    K, W = max(args.steps, 1), max(args.warmup, 3)

    world = build_world(args.motions)
    env = build_gpu_env(world, N, rank * N, n_gpus * N, dev, interleaved=not args.four_tensor_state)
    init_gpu_state(env, rank * N)
    if distributed and args.fail_rate_sync > 0:
        env.enable_fail_rate_sync(interval=args.fail_rate_sync)      # the only exchange besides the final statistics
    from parc_b200 import synth
    noise = synth.standin_noise(N, env._J, env._D, seed=5678 + rank * N, fail_fraction=args.fail_fraction).to(dev)   # simulator stand-in

    def barrier():
        if distributed:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def max_over_ranks(ms):
        if not distributed:
            return ms
        tt = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return float(tt.item())

    def final_stats(e):
        # final scalar reward / termination statistics: the only cross-rank exchange of the path
        st = torch.stack([e._reward_buf.sum().double(), e._done_count.sum().double(), e._motion_id_fail_rates.mean().double()])
        if distributed:
            dist.all_reduce(st, op=dist.ReduceOp.SUM)
        return st

    final_stats(env)       # warm-up covers the statistics reduction too (first use loads its kernels / NCCL channels)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    # ---- device-resident timing -------------------------------------------------------------------------
    dl = device_loop(env, noise, K, W, barrier, max_over_ranks)
    env.sim_standin(noise)
    env.post_physics_step()
    stats = final_stats(env)             # done count of a steady-state step, before the reset clears the flags
    env.reset_done()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms_total, kern_ms = dl["ms_total"], dl["kern_ms"]
    value = n_gpus * N * K / (ms_total * 1e-3)

    # ---- end to end: host buffers in, host results out, every step -----------------------------------
    host_state = env.make_host_state()
    for k, v in env.sim_state_tensors().items():
        host_state[k].copy_(v)
    host_res = env.make_host_results()
    torch.cuda.synchronize(dev)
    h2d = sum(v.numel() * v.element_size() for v in host_state.values())
    d2h = sum(v.numel() * v.element_size() for v in host_res.values())
    Ke = K

    def time_e2e(one_step):
        for _ in range(3):
            one_step()
            torch.cuda.current_stream(dev).synchronize()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(Ke):
            one_step()
            torch.cuda.current_stream(dev).synchronize()   # the caller needs the results before the next physics step
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1))

    def e2e_serial():
        env.upload_sim_state(host_state)       # this step's simulator state from pinned host memory, all of it first
        env.sim_standin(noise)                 # (its values are then re-centred on the reference: stationary workload)
        env.post_physics_step()
        env.download_results(host_res)         # reward + done back to the host
        env.reset_done()

    def e2e_chunked():
        # the public host-buffer call: parc_b200_env_step_host uploads range c+1 while range c is stepped
        env.step_host(host_state, host_res, num_chunks=args.chunks, standin_noise=noise)

    def h2d_only():
        env.upload_sim_state(host_state)       # the PCIe floor of the host-fed step: the upload alone

    h2d_ms = time_e2e(h2d_only)
    e2e_serial_ms = time_e2e(e2e_serial)
    e2e_ms = time_e2e(e2e_chunked)
    e2e_value = n_gpus * N * Ke / (e2e_ms * 1e-3)

    # ---- the other BASELINE configs, measured by the same invocation (`extra`) -------------------------
    extra = None
    if not args.no_extra:
        extra = {}
        del env, noise, host_state, host_res
        torch.cuda.empty_cache()
        Kx, Wx = min(K, 20), 3

        def side_config(envs_per_rank, fail_fraction, label):
            e = build_gpu_env(world, envs_per_rank, rank * envs_per_rank, n_gpus * envs_per_rank, dev)
            init_gpu_state(e, rank * envs_per_rank)
            nz = synth.standin_noise(envs_per_rank, e._J, e._D, seed=5678 + rank * envs_per_rank, fail_fraction=fail_fraction).to(dev)
            smp = ClockSampler(local_rank)
            if rank == 0:
                smp.start()
            r = device_loop(e, nz, Kx, Wx, barrier, max_over_ranks)
            ck = smp.stop() if rank == 0 else None
            out = {"workload": label, "n_gpus": n_gpus, "envs_per_gpu": envs_per_rank, "envs_total": n_gpus * envs_per_rank, "steps": Kx, "warmup": Wx,
                   "value": n_gpus * envs_per_rank * Kx / (r["ms_total"] * 1e-3), "unit": UNIT, "ms_per_step": r["ms_total"] / Kx,
                   "resets_per_step_per_gpu": r["done_per_step"], "standin_fail_fraction": fail_fraction,
                   "roofline": roofline_block(e, r["kern_ms"], r["ms_total"] / Kx), "clocks": ck}
            del e, nz
            torch.cuda.empty_cache()
            return out

        # configs[4]: reset-heavy sweep, 262,144 envs in total over the invoked ranks, ~10 % terminate and are RSI-reset per step
        extra["config5_reset_heavy"] = side_config(262144 // n_gpus, 0.1, "BASELINE configs[4]: 262,144 envs over %d GPU(s), 10 %% early "
                                                   "termination + RSI reset per step" % n_gpus)
        # configs[2]: one GPU's share of 1,048,576 envs over 8 GPUs (the full config when invoked with --gpus 8)
        extra["config3_share"] = side_config(131072, args.fail_fraction, "BASELINE configs[2]: 131,072 envs per GPU (= 1,048,576 over 8 GPUs%s)"
                                             % ("" if n_gpus == 8 else "; this run has %d" % n_gpus))
        # configs[3]: kinematic optimizer batch, rank 0 only (it does not shard in the reference either: independent motions)
        if rank == 0:
            import bench_optimizer
            extra["config4_optimizer"] = bench_optimizer.run(1024, 300, steps=5, warmup=3, cpu_motions=0, dev=dev, peak_gbs=measured_peaks()[0])
        barrier()

    # ---- CPU baseline (rank 0, N=1 only) ------------------------------------------------------------
    cb = None
    if rank == 0 and n_gpus == 1 and not args.no_cpu_baseline:
        cb, _ = cpu_baseline_block(world, args, 20, 3)

    if rank == 0:
        traffic, traffic_src, warp_inst = None, None, None
        tpath = os.path.join(ROOT, "profiles", "env_step_traffic.json")
        if os.path.exists(tpath):
            with open(tpath) as f:
                tj = json.load(f)
            traffic, traffic_src = tj.get("dram_bytes_per_launch"), tj.get("source")
            warp_inst = tj.get("warp_instructions_per_launch") if N == ENVS_PER_GPU else None
        # re-create what the roofline needs (the env object is gone when extras ran): sizes only
        class _Sizes:
            pass
        rf_env = _Sizes()
        rf_env.get_num_envs = lambda: N
        rf_env._ray_xy_points = world["ray"]
        rf_env._mlib = _Sizes()
        rf_env._mlib._frame_root_pos = torch.empty(sum(int(c["root_pos"].shape[0]) for c in world["clips"]), 0)
        rf_env._terrain = _Sizes()
        rf_env._terrain.hf = world["hf"]
        roof = roofline_block(rf_env, kern_ms, ms_total / K, traffic if N == ENVS_PER_GPU else None, warp_inst, (clocks or {}).get("sm_mhz"))
        roof["traffic_source"] = traffic_src
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": ms_total / K, "host_enqueue_ms_per_step": dl["host_enqueue_ms"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "65,536 envs/GPU, 1,000-clip synthetic library of the repo humanoid, 32x32 local heightmap "
                                   "observations (BASELINE configs[1]); step = simulator stand-in (char := ref + fixed noise, FK) + post-physics "
                                   "pass + fail-rate EMA + device-side RSI reset + obs refresh of the reset envs, through TrackingEnv.step()",
                       "envs_per_gpu": N, "clips": args.motions, "frames": int(rf_env._mlib._frame_root_pos.shape[0]),
                       "ray_points": int(world["ray"].shape[0]), "resets_per_step": dl["done_per_step"],
                       "standin_fail_fraction": args.fail_fraction, "obs_dim": 871 + int(world["ray"].shape[0]),
                       "terrain_cells": list(world["hf"].shape), "parallelism": "env-sharded x%d, no data-path collective" % n_gpus,
                       "fail_rate_sync": ("all-reduce of 2 x %d int64 outcome counts every %d steps (NCCL)" % (args.motions, args.fail_rate_sync))
                                         if (distributed and args.fail_rate_sync > 0) else "none (one process)",
                       "l2": "per-step working set %.2f GB > 126 MB L2 (no flush needed)" % (roof["algorithmic_bytes_per_launch"] / 1e9)},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": e2e_ms / Ke, "chunks": args.chunks, "ms_per_step_unpipelined": e2e_serial_ms / Ke,
                    "h2d_ms": h2d_ms / Ke, "h2d_gbs": h2d / (h2d_ms / Ke * 1e-3) / 1e9,
                    "state_layout": "four tensors" if args.four_tensor_state else "one interleaved row per env (one copy per range)",
                    "note": "TrackingEnv.step_host -> parc_b200_env_step_host: every step uploads the simulator state from pinned host memory "
                            "in row ranges (range c+1 in flight while range c is stepped), fail-rate EMA, D2H of reward+done, RSI reset, host "
                            "sync; the observation stays on the device for the policy, as in the reference. h2d_ms = the upload alone (PCIe floor)"},
            "gpu_launches": LAUNCHES_PER_STEP * K,
            "roofline": roof,
            "cpu_baseline": cb,
            "extra": extra,
            "stats": {"reward_sum": float(stats[0]), "done_last_step": float(stats[1]), "mean_fail_rate": float(stats[2]) / n_gpus},
        }
        print(json.dumps(line))
    if distributed:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
