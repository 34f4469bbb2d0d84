"""Multi-GPU layer: fields are independent given the global hyper-parameters
(CelestePy/models.py:75-79 loops over fields; each Field owns its images and
epsilon, :110-160), so ranks shard FIELDS and no pixel ever crosses NVLink.
The only exchange is one small float64 all-reduce per sweep carrying the global
log-likelihood, the per-band noise-photon sums and the sufficient statistics
of the shared empirical colour prior.

One process per GPU (torchrun); backend "nccl" on GPUs, "gloo" in CPU tests.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist

N_BANDS = 5
COLOR_DIM = 5          # FluxColorMoG works in (4 colours + 1 magnitude), util/dists/flux_prior.py:4-40


def init_process_group(backend: str | None = None):
    """Join the job torchrun launched (RANK/WORLD_SIZE/MASTER_* in the env)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1 or dist.is_initialized():
        return
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        dist.init_process_group(backend)


def rank_world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_fields(n_fields: int, rank: int | None = None, world: int | None = None):
    """Round-robin field ownership: field f lives on rank f % world."""
    if rank is None or world is None:
        rank, world = rank_world()
    return list(range(rank, n_fields, world))


class SweepStats(object):
    """Layout of the per-sweep all-reduce vector (float64):
         [0]                    sum of per-image log-likelihoods
         [1 : 6]                noise-photon sums per band (models.py:156-160)
         [6], [7]               number of stars, galaxies
         then for each of the two colour priors (star, galaxy) with K comps:
         sum_s r_sk (K) | sum_s r_sk c_s (K x 5) | sum_s r_sk c_s c_s^T (K x 15, upper triangle)
    """

    def __init__(self, n_prior_comp: int = 8, device=None):
        self.K = int(n_prior_comp)
        self.block = self.K * (1 + COLOR_DIM + COLOR_DIM * (COLOR_DIM + 1) // 2)
        self.size = 8 + 2 * self.block
        self.vec = torch.zeros(self.size, dtype=torch.float64, device=device)

    def zero_(self):
        self.vec.zero_()
        return self

    def add_loglik(self, ll):
        """ll: tensor of per-image log-likelihoods living on self.vec's device."""
        self.vec[0] += ll.sum()

    def add_noise(self, per_band):
        self.vec[1:6] += torch.as_tensor(per_band, dtype=torch.float64, device=self.vec.device)

    def add_counts(self, n_star, n_gal):
        self.vec[6] += n_star
        self.vec[7] += n_gal

    def add_prior_stats(self, which: int, resp, colors):
        """which: 0 star prior, 1 galaxy prior; resp [S, K] responsibilities,
        colors [S, 5] colour-space coordinates (host arrays; a few hundred rows)."""
        resp = np.asarray(resp, dtype=np.float64)
        colors = np.asarray(colors, dtype=np.float64)
        iu = np.triu_indices(COLOR_DIM)
        s0 = resp.sum(axis=0)
        s1 = resp.T @ colors
        outer = colors[:, iu[0]] * colors[:, iu[1]]
        s2 = resp.T @ outer
        flat = np.concatenate([s0, s1.ravel(), s2.ravel()])
        o = 8 + which * self.block
        self.vec[o:o + self.block] += torch.from_numpy(flat).to(self.vec.device)

    def add_prior_stats_device(self, which: int, block: torch.Tensor):
        """block: engine.prior_sufficient_stats(resp, colors) -- the same
        [s0 | s1 | s2] layout, computed on the device (cg_prior_stats)."""
        if block.numel() != self.block:
            raise ValueError("prior block has %d entries, expected %d" % (block.numel(), self.block))
        o = 8 + which * self.block
        self.vec[o:o + self.block] += block.to(self.vec.device)

    def prior_stats(self, which: int):
        o = 8 + which * self.block
        v = self.vec[o:o + self.block].cpu().numpy()
        K = self.K
        s0 = v[:K]
        s1 = v[K:K + K * COLOR_DIM].reshape(K, COLOR_DIM)
        s2 = v[K + K * COLOR_DIM:].reshape(K, -1)
        return s0, s1, s2

    def all_reduce(self):
        """SUM over ranks, in place (one NCCL all-reduce of ~2.7 KB per sweep)."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.vec, op=dist.ReduceOp.SUM)
        return self.vec


def max_over_ranks(x: float, device=None) -> float:
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
