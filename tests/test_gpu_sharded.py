"""GPU, world size 2: the PRODUCT's env-sharded step (one process per rank, contiguous env blocks, no data-path collective)
with the cross-rank fail-rate combination of SURVEY.md section 8e (sharding.FailRateSync: one all-reduce of per-motion
fail / pass counts). NCCL when the box has two GPUs, gloo with both ranks on cuda:0 otherwise (the driver's one-GPU box)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import helpers as H

pytestmark = pytest.mark.gpu
N_TOTAL, M = 96, 5


def _scenario():
    return H.Scenario(num_envs=N_TOTAL, num_motions=M, ray="grid", seed=21, wrap_fraction=0.2, max_frames=80)


def _shard_env(sc, full, first, count, dev):
    sub = H.Scenario.__new__(H.Scenario)
    sub.__dict__.update(sc.__dict__)
    sub.N, sub.env_offsets = count, sc.env_offsets[first:first + count]
    penv = sub.product_env(dev)
    cut = H.Scenario.__new__(H.Scenario)         # an oracle view of the rows, only to reuse copy_oracle_state_to_product
    cut.__dict__.update(sub.__dict__)
    o = cut.oracle_env()
    sl = slice(first, first + count)
    for name in ("motion_ids", "motion_terrain_ids", "motion_time_offsets", "timestep_buf", "time_buf", "char_root_pos", "char_root_rot",
                 "char_root_vel", "char_root_ang_vel", "char_dof_pos", "char_dof_vel", "char_rigid_body_pos", "char_contact_forces"):
        getattr(o, name)[:] = getattr(full, name)[sl]
    H.copy_oracle_state_to_product(o, penv)
    return penv


def _worker(rank, world, port, out_dir, two_gpus):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dev = "cuda:%d" % (rank if two_gpus else 0)
    torch.cuda.set_device(torch.device(dev))
    dist.init_process_group("nccl" if two_gpus else "gloo", rank=rank, world_size=world)
    try:
        from parc_b200 import sharding
        sc = _scenario()
        full = sc.oracle_env()
        H.randomize_oracle_state(full, seed=8, fail_fraction=0.3)
        first, count = sharding.shard_range(N_TOTAL, rank, world)
        penv = _shard_env(sc, full, first, count, dev)
        penv._motion_id_fail_rates.copy_(torch.linspace(0.2, 0.9, M))
        penv.enable_fail_rate_sync(interval=1)
        penv.post_physics_step()
        torch.cuda.synchronize()
        np.savez(os.path.join(out_dir, "rank%d.npz" % rank), first=first, count=count, obs=penv._obs_buf.cpu().numpy(),
                 reward=penv._reward_buf.cpu().numpy(), done=penv._done_buf.cpu().numpy(), rates=penv._motion_id_fail_rates.cpu().numpy(),
                 cdf=penv._cdf.cpu().numpy(), code=penv._done_code.cpu().numpy(), mids=penv._motion_ids.cpu().numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_ranks_step_their_shards_and_agree_on_the_fail_rates(tmp_path):
    from parc_b200 import sharding
    two_gpus = torch.cuda.device_count() >= 2
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path), two_gpus), nprocs=2, join=True)
    r = [np.load(str(tmp_path / ("rank%d.npz" % k))) for k in range(2)]
    # per-env outputs: each shard's rows equal the single-process run's rows (the oracle on all 96 envs)
    sc = _scenario()
    full = sc.oracle_env()
    H.randomize_oracle_state(full, seed=8, fail_fraction=0.3)
    full.post_physics_step()
    W = float(full.ref_root_pos.abs().max())
    for k in range(2):
        sl = slice(int(r[k]["first"]), int(r[k]["first"]) + int(r[k]["count"]))
        assert np.array_equal(r[k]["done"], full.done_buf[sl].numpy()), "done flags of shard %d" % k
        H.assert_close_world(torch.from_numpy(r[k]["obs"]), full.obs_buf[sl], W, what="obs of shard %d" % k)
        H.assert_close(torch.from_numpy(r[k]["reward"]), full.reward_buf[sl], atol=2e-5, what="reward of shard %d" % k)
    # fail rates: identical on both ranks, and equal to "combine the counts of the shards" applied to the starting rates
    assert np.array_equal(r[0]["rates"], r[1]["rates"]) and np.array_equal(r[0]["cdf"], r[1]["cdf"])
    counts = sum(sharding.motion_outcome_counts(torch.from_numpy(x["code"]), torch.from_numpy(x["mids"]), M) for x in r)
    assert int(counts.sum()) > 5
    want = sharding.apply_outcome_counts(torch.linspace(0.2, 0.9, M), counts, 0.01)
    np.testing.assert_allclose(r[0]["rates"], want.numpy(), rtol=1e-6, atol=0)
    assert not np.allclose(r[0]["rates"], np.linspace(0.2, 0.9, M))


def test_outcome_counts_kernel_equals_the_torch_counts():
    """parc_b200_outcome_counts (what FailRateSync accumulates on the device every step) against sharding.motion_outcome_counts
    on the same finished envs, accumulated over two steps."""
    from parc_b200 import _lib, sharding
    sc = H.Scenario(num_envs=1003, num_motions=7, ray="grid", seed=4, wrap_fraction=0.2, max_frames=90)
    oenv = sc.oracle_env()
    H.randomize_oracle_state(oenv, seed=6, fail_fraction=0.3)
    penv = sc.product_env("cuda:0")
    H.copy_oracle_state_to_product(oenv, penv)
    Mn = penv._mlib.num_motions()
    acc = torch.zeros(2, Mn, dtype=torch.int64, device="cuda:0")
    want = torch.zeros_like(acc)
    for _ in range(2):
        penv.post_physics_step(update_fail_rates=False)
        want += sharding.motion_outcome_counts(penv._done_code, penv._motion_ids, Mn)
        _lib.check(_lib.lib().parc_b200_outcome_counts(penv._bufs, Mn, _lib.ptr(acc), penv._stream()))
        penv.reset_done()
    assert int(want.sum()) > 0 and int(want[0].sum()) > 0 and int(want[1].sum()) >= 0
    assert torch.equal(acc, want)
