"""Env sharding across the GPUs of one box.

Every per-step quantity of the tracking path depends on one env's state, its clip and the read-only
terrain, so envs are split into contiguous index blocks, one per rank, and the motion library and
the tiled heightfield are replicated. Nothing is exchanged on the hot path; the only collective is
a small sum of episode statistics (and, when asked for, of the per-motion fail / success counts
that drive the fail-rate EMA, dm_env.py:648-660) at a logging interval.

The reference has no multi-process mode; its global quantities are the running means of
ReturnTracker (return_tracker.py:73-90) and `_motion_id_fail_rates`. Those are what reduce here.
"""
import torch
import torch.distributed as dist

STAT_KEYS = ("reward_sum", "done_count", "fail_count", "env_steps")


def shard_range(total_envs: int, rank: int, world_size: int):
    """Contiguous block of global env ids owned by `rank`: (first, count). Blocks differ by at most
    one env and cover [0, total_envs) exactly."""
    assert total_envs >= 0 and world_size >= 1 and 0 <= rank < world_size
    base, rem = divmod(total_envs, world_size)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def owner_of(env_id: int, total_envs: int, world_size: int) -> int:
    """Rank that owns a global env id (inverse of shard_range)."""
    base, rem = divmod(total_envs, world_size)
    split = rem * (base + 1)
    if env_id < split:
        return env_id // (base + 1)
    return rem + (env_id - split) // max(base, 1)


def local_stats(reward_buf: torch.Tensor, done_buf: torch.Tensor) -> torch.Tensor:
    """[reward_sum, done_count, fail_count, env_steps] of one shard after a step, as float64 on the
    buffers' device (DoneFlags.FAIL == 1, base_env.py:13-17)."""
    return torch.stack([reward_buf.sum(dtype=torch.float64), (done_buf != 0).sum().double(), (done_buf == 1).sum().double(),
                        torch.tensor(float(reward_buf.shape[0]), dtype=torch.float64, device=reward_buf.device)])


def reduce_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the statistics vector over ranks (NCCL on GPUs, gloo in the CPU tests). No-op without a
    process group -- a single process owns every env."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def motion_outcome_counts(done_code: torch.Tensor, motion_ids: torch.Tensor, num_motions: int) -> torch.Tensor:
    """[2, M] int64: per motion, how many envs of this shard finished as FAIL (row 0) and as
    TIME/SUCC/motion end (row 1) in the last step. `done_code` is the step kernel's uint8 code
    (0 running, 1 fail, 2 other). Summed over ranks this is the cross-shard input of the fail-rate
    EMA; the order of updates inside one step cannot follow global env ids across processes, so the
    combined update applies all passes then all fails of a step (documented in DESIGN.md)."""
    out = torch.zeros(2, num_motions, dtype=torch.int64, device=motion_ids.device)
    # scatter-add of 0/1 flags: no boolean indexing, hence no device -> host synchronisation on the step path
    out[0].scatter_add_(0, motion_ids, (done_code == 1).to(torch.int64))
    out[1].scatter_add_(0, motion_ids, (done_code == 2).to(torch.int64))
    return out


def apply_outcome_counts(fail_rates: torch.Tensor, counts: torch.Tensor, ema_weight: float = 0.01) -> torch.Tensor:
    """Fail-rate EMA from reduced counts: rate <- rate * (1-w)^(passes+fails), then `fails` times
    rate <- rate + w applied after the decay of the later updates, i.e. the closed form of the
    reference's sequential rule with all passes first. In place; returns fail_rates."""
    keep = float(torch.tensor(1.0 - ema_weight, dtype=torch.float32))     # the fp32 value of 1 - w, as a host scalar (no device copy, no sync)
    fails, passes = counts[0].to(fail_rates.device), counts[1].to(fail_rates.device)
    r = fail_rates * torch.pow(keep, passes.float())
    # k fails in a row: r*keep^k + w*(1 + keep + ... + keep^(k-1)) = r*keep^k + (1 - keep^k)
    kk = torch.pow(keep, fails.float())
    fail_rates.copy_(r * kk + (1.0 - kk))
    return fail_rates


class FailRateSync:
    """Cross-rank fail rates for env-sharded runs. Each rank keeps applying the reference's sequential
    EMA to its own envs every step (its RSI sampling needs fresh rates without waiting for anybody); every `interval` steps
    the per-motion fail / pass counts accumulated since the last exchange are summed over the ranks with ONE all-reduce of
    2 M int64 (NCCL over NVLink on the GPUs) and every rank replaces its rates by
    apply_outcome_counts(rates at the last exchange, summed counts) -- "run the reference once per shard and combine the
    counts". All ranks then hold the same rates again."""

    def __init__(self, fail_rates: torch.Tensor, interval: int = 1, ema_weight: float = 0.01, group=None):
        self.interval, self.w, self.group = max(int(interval), 1), float(ema_weight), group
        self.base = fail_rates.clone()
        self.acc = torch.zeros(2, fail_rates.shape[0], dtype=torch.int64, device=fail_rates.device)
        self.steps = 0
        # the first collective of a process group sets up its communicator (tens of milliseconds with NCCL): do it here, on the
        # zero table, not inside somebody's step loop. Every rank constructs its FailRateSync, so the call is collective.
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(self.acc, op=dist.ReduceOp.SUM, group=self.group)

    def after_step(self, fail_rates, done_code, motion_ids, accumulate=None):
        """Call once per control step, after the local EMA. `accumulate(acc)`: adds this step's counts to `acc` on the device
        (TrackingEnv passes parc_b200_outcome_counts: one launch, nothing on the host); without it the counts come from
        motion_outcome_counts. Returns True when the rates were replaced by the combined ones."""
        if accumulate is not None:
            accumulate(self.acc)
        else:
            self.acc += motion_outcome_counts(done_code, motion_ids, fail_rates.shape[0])
        self.steps += 1
        if self.steps % self.interval:
            return False
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(self.acc, op=dist.ReduceOp.SUM, group=self.group)
        apply_outcome_counts(self.base, self.acc, self.w)
        fail_rates.copy_(self.base)
        self.acc.zero_()
        return True
