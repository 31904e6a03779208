"""Stability driver: several UASE runs on the same layers, aggregated with Borda across runs.

The reference has no driver for this; its README prescribes the procedure ("we incorporate
stability into n2v2r ... by running it multiple times and aggregating the results",
README.md:54; "node2vec2rank runs multiple times for every parameter combination and then all
the rankings are aggregated into one using the Borda scheme", README.md:173) and BASELINE.json
configs[4] names the workload (16 seeds x 4-layer SBM, rank aggregation across runs, 8 GPUs).
One run = N2V2R.fit_transform_rank() with `seed` = one entry of `seeds` (model.py:36-38 seeds
the start vector of the SVD); the aggregate is N2V2R.aggregate_transform() (model.py:149-201)
over the union of all runs' columns, i.e. S x len(dims) x len(metrics) rankings per comparison.

The layers are built once and stay resident in HBM; only the solver's start block changes per
run.  Under torch.distributed (config "seed_sharded": true, one process per GPU) every rank
holds all layers, the seeds are split over the ranks in contiguous blocks
(dist.layer_partition), score tables are all-gathered and every rank computes the same Borda.
No data-path collective runs inside a solve.
"""
from __future__ import annotations

import os
import time

import numpy as np
import pandas as pd
import torch

from . import _native as N
from . import device as D
from . import dist
from .model import N2V2R, check_solver_outcome


class StabilityRun:
    """Multi-seed UASE runs + Borda across runs.

    graphs, nodes, config: exactly what N2V2R takes (model.py:18); config["seed"] is ignored in
    favour of `seeds`.  Attributes after run(): pairwise_ranks (dict key -> DataFrame, columns
    "seed-{s}_dim-{d}_distance-{m}", seeds outer), pairwise_aggregate_ranks (dict key ->
    DataFrame 'borda_ranks' int64, node order), per_seed_aggregate_ranks (dict key -> DataFrame,
    one Borda column per seed), solver_infos (list of dicts, this rank's runs)."""

    def __init__(self, graphs: list, nodes: list, config: dict, seeds):
        self.seeds = [int(s) for s in seeds]
        if not self.seeds:
            raise ValueError("at least one seed is needed")
        self.config = dict(config)
        if int(self.config.get("seed_batch", 6)) > 1 and self.config.get("layer_storage", "auto") == "auto":
            self.config["layer_storage"] = "planes"       # seed batching multiplies bf16 planes (three operands per layer pass)
        self.node_names = nodes
        self._model = N2V2R(graphs=graphs, nodes=nodes, config=self.config)
        self.save_dir = self._model.save_dir
        self.pairwise_ranks = None
        self.pairwise_aggregate_ranks = None
        self.per_seed_aggregate_ranks = None
        self.solver_infos = []
        self.batch_stats = []
        self._scores = None                 # (S, C, cols, n) float64, device

    def _my_seeds(self):
        if bool(self.config.get("seed_sharded")) and dist.is_distributed():
            part = dist.layer_partition(len(self.seeds), torch.distributed.get_world_size(), torch.distributed.get_rank())
            return [self.seeds[i] for i in part], True
        return list(self.seeds), False

    def scores_device(self):
        """Run every seed of this rank and return the score tensor of ALL seeds, (S, C, cols, n) float64."""
        cfg = self.config
        m = self._model
        layers = m._device_layers()
        n = layers[0].n
        k = m.max_embed_dim
        if k >= n:
            raise ValueError(f"`k` must be an integer satisfying `0 < k < min(A.shape)` (k={k}, n={n}).")
        if k > D.MAX_EMBED_DIM:
            raise NotImplementedError(f"max(embed_dimensions) = {k}: embedding dimensions up to {D.MAX_EMBED_DIM} are supported")
        if m.comp_strategy not in N.STRATEGIES:
            raise UnboundLocalError("cannot access local variable 'embed_one' where it is not associated with a value")
        l = int(cfg.get("block_width") or D.default_block_width(k, n))
        _, dims, mets = D.column_specs(m.embed_dimensions, m.distance_metrics)
        strategy = N.STRATEGIES[m.comp_strategy]
        mine, sharded = self._my_seeds()
        local = []
        self.solver_infos = []
        self.batch_stats = []
        # "seed_batch" (additive config key, default 6): that many solves run concurrently and share every pass over a layer,
        # three operands per pass (device.uase_batched); each solve is bit-identical to its unbatched run.  1 = one after the other.
        group = max(1, min(16, int(cfg.get("seed_batch", 6))))
        batchable = group > 1 and all(L.storage == "planes" for L in layers)
        for g0 in range(0, len(mine), group if batchable else 1):
            chunk = mine[g0:g0 + (group if batchable else 1)]
            if batchable and len(chunk) > 1:
                solved, bstats = D.uase_batched(layers, k, chunk, l=l, tol=float(cfg.get("solver_tol", 1e-5)),
                                                max_iters=int(cfg.get("solver_max_iters", 400)), check_dims=m.embed_dimensions)
                self.batch_stats.append(bstats)
            else:
                solved = []
                for seed in chunk:
                    emb, sigma, _, info = D.uase(layers, k, l=l, tol=float(cfg.get("solver_tol", 1e-5)),
                                                 max_iters=int(cfg.get("solver_max_iters", 400)), seed=seed,
                                                 check_dims=m.embed_dimensions)
                    solved.append((emb, sigma, info))
            for seed, (emb, _, info) in zip(chunk, solved):
                info["seed"] = seed
                check_solver_outcome(info, cfg)
                self.solver_infos.append(info)
                local.append(D.distances(emb, dims, mets, strategy))          # (C, cols, n)
        n_comp = len(layers) if strategy == N.ONE_VS_REST else len(layers) - 1
        shape = (max(n_comp, 0), len(dims), n)
        if local:
            stacked = torch.stack(local)
        else:
            stacked = torch.empty((0,) + shape, dtype=torch.float64, device=N.require_gpu())
        if sharded:
            stacked = dist.gather_blocks(stacked, len(self.seeds))
        self._scores = stacked
        return stacked

    def run(self, method="Borda"):
        """All runs + aggregation.  Returns pairwise_aggregate_ranks."""
        if method.casefold() != "borda":
            raise NotImplementedError('Aggregation method not found. Available methods: Borda')
        verbose = self.config.get("verbose", 0)
        tic = time.time()
        if verbose >= 0:
            print(f"\nRunning n2v2r stability analysis with {len(self.seeds)} seeds ...")
        scores = self.scores_device()
        S, C, cols, n = scores.shape
        names, _, _ = D.column_specs(self._model.embed_dimensions, self._model.distance_metrics)
        # Borda across runs: every comparison ranks the union of all seeds' columns (seeds outer)
        union = scores.permute(1, 0, 2, 3).reshape(C, S * cols, n)
        borda = D.rank_borda(union)[0].cpu().numpy()                      # (C, n) int64
        per_seed = D.rank_borda(scores.reshape(S * C, cols, n))[0].reshape(S, C, n).cpu().numpy()
        host = D.to_host(union)
        all_names = [f"seed-{s}_{c}" for s in self.seeds for c in names]
        self.pairwise_ranks, self.pairwise_aggregate_ranks, self.per_seed_aggregate_ranks = {}, {}, {}
        for c in range(C):
            key = str(c + 1)
            self.pairwise_ranks[key] = pd.DataFrame(host[c].T, index=self.node_names, columns=all_names)
            self.pairwise_aggregate_ranks[key] = pd.DataFrame(borda[c], index=self.node_names, columns=["borda_ranks"])
            self.per_seed_aggregate_ranks[key] = pd.DataFrame(per_seed[:, c, :].T, index=self.node_names,
                                                              columns=[f"seed-{s}_borda_ranks" for s in self.seeds])
        if verbose >= 0:
            print(f"n2v2r aggregated {S * cols} rankings for {C} comparison(s) in {round(time.time() - tic, 2)} seconds")
        if self.save_dir and (not dist.is_distributed() or torch.distributed.get_rank() == 0):
            for key in self.pairwise_ranks:
                self.pairwise_ranks[key].to_csv(os.path.join(self.save_dir, key + ".tsv"), sep='\t', index=True)
                self.pairwise_aggregate_ranks[key].to_csv(os.path.join(self.save_dir, key + "_agg.tsv"), sep='\t', index=True)
        return self.pairwise_aggregate_ranks

    def agreement(self):
        """Spearman correlation of every seed's own Borda ranking with the across-run aggregate:
        dict key -> ndarray (S,).  A stability read-out, not part of the reference."""
        if self.pairwise_aggregate_ranks is None:
            raise ValueError("No n2v2r embeddings found")
        out = {}
        for key, agg in self.pairwise_aggregate_ranks.items():
            a = agg.iloc[:, 0].rank().to_numpy()
            per = self.per_seed_aggregate_ranks[key]
            out[key] = np.asarray([np.corrcoef(a, per.iloc[:, j].rank().to_numpy())[0, 1] for j in range(per.shape[1])])
        return out
