"""Host-side logic of the multi-GPU path on CPU: two `gloo` ranks shard the 2F+1 finite-difference
columns with omo_shard_range, "evaluate" their slice (a stand-in function here -- no GPU, the product has no CPU
path), all-gather the padded slices the way bench.py / INTEGRATION.md do, and assemble the gradient with
omo_grad_from_columns. Checks that the gathered column order and the quotient are those of the single-rank path."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from optimo_b200.core import grad_from_columns, shard_range


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _column_value(x, free_idx, h, col):
    """Stand-in objective of perturbation column `col` (column 0 = x, 2k+1 = x + h e_k, 2k+2 = x - h e_k)."""
    y = x.copy()
    if col > 0:
        k = free_idx[(col - 1) >> 1]
        y[k] += -h if ((col - 1) & 1) else h
    return float(np.sum(np.sin(y) * np.arange(1, y.size + 1)))


def _worker(rank, world, port, F, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(3)
    x = rng.standard_normal(3 * F)
    free_idx = np.arange(0, 3 * F, 3)
    h = 1e-4
    n_cols = 2 * F + 1
    begin, count = shard_range(n_cols, rank, world)
    pad = max(shard_range(n_cols, r, world)[1] for r in range(world))
    mine = torch.zeros(pad, dtype=torch.float64)
    for c in range(count):
        mine[c] = _column_value(x, free_idx, h, begin + c)
    gathered = [torch.zeros(pad, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(gathered, mine)
    cols = np.concatenate([gathered[r].numpy()[: shard_range(n_cols, r, world)[1]] for r in range(world)])
    f0, g = grad_from_columns(cols, h)
    # max-over-ranks timing reduction used by bench.py
    t = torch.tensor([1.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), cols=cols, f0=f0, g=g, tmax=t.item())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_column_sharding_matches_single_rank(tmp_path):
    F, world = 23, 2          # 47 columns: uneven split 24 / 23
    mp.spawn(_worker, args=(world, _free_port(), F, str(tmp_path)), nprocs=world, join=True)
    rng = np.random.default_rng(3)
    x = rng.standard_normal(3 * F)
    free_idx = np.arange(0, 3 * F, 3)
    ref_cols = np.array([_column_value(x, free_idx, 1e-4, c) for c in range(2 * F + 1)])
    f0, g = grad_from_columns(ref_cols, 1e-4)
    for r in range(world):
        d = np.load(tmp_path / f"rank{r}.npz")
        assert np.array_equal(d["cols"], ref_cols)
        assert d["f0"] == f0 and np.array_equal(d["g"], g)
        assert d["tmax"] == float(world)
    analytic = np.cos(x[free_idx]) * (free_idx + 1)
    assert np.max(np.abs(g - analytic)) < 1e-5
