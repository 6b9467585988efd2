"""Batch sharding over ranks and the per-rank statistics gather.

The path shards by independent OCPs: rank r of `world` takes a contiguous slice of the global batch, there is no
data-path collective; the only exchange is one all_gather of a small statistics vector per timed region.  The same
functions serve bench.py (NCCL, one rank per GPU) and the CPU tests (gloo, world_size 2)."""
import numpy as np

STAT_FIELDS = ("ms_total", "e2e_s", "flops", "solved", "qp_iters", "sqp_iters", "sum_T", "min_T", "max_T", "n_ocp")


def shard_range(rank, world, n):
    """Contiguous slice [lo, hi) of n problems owned by `rank`; the remainder goes to the low ranks."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def local_stats(ms_total, e2e_s, flops, status, qp_iters, sqp_iters, sol_T):
    sol_T = np.asarray(sol_T, dtype=np.float64)
    return np.array([ms_total, e2e_s, flops, float((np.asarray(status) == 1).sum()), float(np.asarray(qp_iters).sum()),
                     float(np.asarray(sqp_iters).sum()), float(sol_T.sum()), float(sol_T.min()) if sol_T.size else np.inf,
                     float(sol_T.max()) if sol_T.size else -np.inf, float(sol_T.size)], dtype=np.float64)


def gather_stats(local, dist=None, device=None):
    """all_gather of the per-rank vector -> (world, len(STAT_FIELDS)) array on every rank."""
    import torch
    t = torch.as_tensor(np.asarray(local, dtype=np.float64), device=device)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return t.cpu().numpy()[None]
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return torch.stack(out).cpu().numpy()


def reduce_stats(allr):
    """Whole-job figures from the gathered matrix: times are max over ranks, counts are sums."""
    allr = np.asarray(allr)
    n = float(allr[:, 9].sum())
    return {"ms_total_max": float(allr[:, 0].max()), "e2e_s_max": float(allr[:, 1].max()), "flops_rank0": float(allr[0, 2]),
            "solved": int(allr[:, 3].sum()), "qp_iters": float(allr[:, 4].sum()), "sqp_iters": float(allr[:, 5].sum()),
            "mean_T": float(allr[:, 6].sum() / n) if n else float("nan"), "min_T": float(allr[:, 7].min()), "max_T": float(allr[:, 8].max()),
            "n_ocp": n}
