"""Mirror of node2vec2rank/model.py (class N2V2R, :17-311) with fit / rank / Borda on the B200.

Public surface kept verbatim: constructor argument order (graphs, nodes, config), attributes
(node_embeddings, pairwise_ranks, pairwise_aggregate_ranks, pairwise_signed_ranks,
pairwise_signed_aggregate_ranks, prior_singed_ranks, save_dir), methods, prints, output files
and exceptions.  `graphs` may be float32 ndarrays (what the reference's DataLoader returns) or
device.DeviceLayer objects (DataLoader with "device_resident": true).

Additive, optional config keys (defaults keep reference behaviour):
  block_width (int, default: max dim + >=24 guard vectors rounded up to 16; + >=8 when every layer is held as row lists): solver block l
  solver_tol (float, default 1e-5), solver_max_iters (int, default 400)
  storage_planes (int | None): bf16 planes per weighted layer (None = exact)
  layer_storage ("auto" | "rows" | "tiles" | "planes" | "e4m3", default "auto"): row lists of the fp32 non-zeros (large layers
      under 4.5 % density), compressed tiles (up to 8 %), dense bf16 planes, one e4m3 plane (binarised) (device.make_layer)
  layer_sharded (bool, default False): under torch.distributed (one process per GPU) `graphs` holds only
      this rank's contiguous block of layers (dist.layer_partition); Y is all-reduced per operator
      application, embeddings are all-gathered, every rank returns the full tables
"""
from __future__ import annotations

import json
import os
import random
import time
import warnings
from datetime import datetime

import numpy as np
import pandas as pd
import torch

from . import _native as N
from . import device as D
from . import dist


SOLVER_BOUND_WARN = 1e-2      # a stagnated solve whose subspace-angle bound is above this is reported to the user


def check_solver_outcome(info, config):
    """The reference's svds raises ArpackNoConvergence when ARPACK gives up; here a solve that hit max_iters, or stopped on
    the stagnation rules with a loose angle bound, still returns its embedding but never silently: a RuntimeWarning, or a
    RuntimeError under the additive config key "solver_strict"."""
    msg = None
    if info["status"] == N.ENOTCONV or info["converged"] == 0:
        msg = (f"UASE solver stopped at max_iters = {info['iterations']} operator applications without converging "
               f"(subspace-angle bound {info['angle_bound']:.2e}); embeddings and rankings may be inaccurate")
    elif info["converged"] == 2 and not (info["angle_bound"] <= SOLVER_BOUND_WARN):
        msg = (f"UASE solver stopped because the residuals stagnated (fp32 floor) with subspace-angle bound "
               f"{info['angle_bound']:.2e} > {SOLVER_BOUND_WARN:g}: some swept dimension cuts a cluster of singular values")
    if msg:
        if config.get("solver_strict"):
            raise RuntimeError(msg)
        warnings.warn(msg, RuntimeWarning, stacklevel=3)


class N2V2R:
    def __init__(self, graphs: list, nodes: list, config: dict):
        self.config = config
        self.node_names = nodes
        self.graphs = graphs
        self.num_graphs = len(graphs)
        self.embed_dimensions = self.config['embed_dimensions']
        self.max_embed_dim = max(self.embed_dimensions)
        self.distance_metrics = self.config['distance_metrics']
        self.save_dir = None
        self.comp_strategy = self.config['comp_strategy']

        self.node_embeddings = None
        self.pairwise_ranks = None
        self.pairwise_signed_ranks = None
        self.pairwise_aggregate_ranks = None
        self.pairwise_signed_aggregate_ranks = None
        self.prior_singed_ranks = None
        self.singular_values = None
        self.solver_info = None
        self._layers = None            # device-resident layers (built lazily, reused by DeDi)
        self.transfer_stats = None     # how the host graphs crossed PCIe (device.layers_from_host.last_stats)
        self._emb_dev = None

        if self.config["seed"]:        # falsy seed (None or 0) => unseeded, model.py:36
            random.seed(self.config["seed"])
            np.random.seed(self.config["seed"])

        now = datetime.now().strftime(r"%m_%d_%Y_%H_%M_%S")
        if self.config["save_dir"]:
            self.save_dir = os.path.join(self.config["save_dir"], now)
            print(self.save_dir)
            if self._writes_files():
                os.makedirs(self.save_dir)          # raises when two runs start in the same second, like the reference
                with open(os.path.join(self.save_dir, "config.json"), 'w', encoding="utf-8") as f:
                    json.dump(self.config, f)

    def _writes_files(self):
        """Under torch.distributed (layer_sharded / seed_sharded runs) every rank holds the full tables; rank 0 writes them."""
        return not dist.is_distributed() or torch.distributed.get_rank() == 0

    # ---- device residency -----------------------------------------------------------------
    def _device_layers(self):
        if self._layers is None:
            host = [None if isinstance(g, D.DeviceLayer) else np.asarray(g) for g in self.graphs]
            arrays = [a for a in host if a is not None]
            planes = self.config.get("storage_planes")
            # the block the solve will use if the layers end up as row lists: sizes their panels (device.rows_panel_cols)
            n_hint = arrays[0].shape[0] if arrays else 0
            l_hint = int(self.config.get("block_width") or D.default_block_width(self.max_embed_dim, max(n_hint, 1), "rows"))
            if arrays and all(a.dtype == np.float32 and a.ndim == 2 and a.shape == arrays[0].shape for a in arrays):
                # H2D of layer t+1 overlaps the load pass of layer t; large sparse layers cross PCIe packed (K0c)
                built = iter(D.layers_from_host(arrays, n_planes=planes, transfer=self.config.get("host_transfer", "auto"),
                                                storage=self.config.get("layer_storage", "auto"), block_width=l_hint))
                self.transfer_stats = D.layers_from_host.last_stats
                self._layers = [g if a is None else next(built) for g, a in zip(self.graphs, host)]
                return self._layers
            layers = []
            for g, a in zip(self.graphs, host):
                if a is None:
                    layers.append(g)
                else:
                    x = D.to_device(a)
                    if x.dtype != torch.float32:        # the reference hands fp32 to the SVD (preprocessing_utils.py:141)
                        x = x.to(torch.float32)
                    layers.append(D.make_layer(x, n_planes=planes, storage=self.config.get("layer_storage", "auto"), block_width=l_hint))
            self._layers = layers
        return self._layers

    def _global_layers(self, layers):
        """(total number of layers over all ranks, index of this rank's first layer)."""
        counts = torch.tensor([len(layers)], device=N.require_gpu())
        allc = [torch.zeros_like(counts) for _ in range(torch.distributed.get_world_size())]
        torch.distributed.all_gather(allc, counts)
        vals = [int(c.item()) for c in allc]
        if min(vals) == 0:
            raise ValueError(f"layer_sharded: every rank needs at least one layer (layers per rank: {vals})")
        return sum(vals), sum(vals[:torch.distributed.get_rank()])

    def _solver_seed(self):
        seed = self.config.get("seed")
        if seed:
            return int(seed)
        return int(np.random.randint(0, 2**31 - 1))     # unseeded run: draw from the global RNG like svds' v0

    # ---- fit: replaces model.py:51-55 (csc_matrix + spectral_embedding.UASE) ----------------
    def __fit(self, graphs):
        layers = self._device_layers()
        n = layers[0].n
        k = self.max_embed_dim
        if k >= n:                                      # scipy svds: "k must satisfy 0 < k < min(A.shape)"
            raise ValueError(f"`k` must be an integer satisfying `0 < k < min(A.shape)` (k={k}, n={n}).")
        if k > D.MAX_EMBED_DIM:
            raise NotImplementedError(f"max(embed_dimensions) = {k}: the B200 solver works on blocks of at most 80 vectors "
                                      f"(k + >= 8 guard vectors), i.e. embedding dimensions up to {D.MAX_EMBED_DIM}")
        storages = {L.storage for L in layers}
        l = int(self.config.get("block_width") or D.default_block_width(k, n, "rows" if storages == {"rows"} else None))
        T_global, offset = len(layers), 0
        sharded = bool(self.config.get("layer_sharded")) and dist.is_distributed()
        if sharded:
            T_global, offset = self._global_layers(layers)
            lw = torch.tensor([l], device=N.require_gpu())
            torch.distributed.all_reduce(lw, op=torch.distributed.ReduceOp.MAX)      # every rank must use the same block
            l = int(lw.item())
        emb, sigma, _, info = D.uase(layers, k, l=l, tol=float(self.config.get("solver_tol", 1e-5)),
                                     max_iters=int(self.config.get("solver_max_iters", 400)),
                                     seed=self._solver_seed(), check_dims=self.embed_dimensions,
                                     T_global=T_global, layer_offset=offset)
        check_solver_outcome(info, self.config)
        if sharded:
            emb = dist.gather_layers(emb, T_global)
        self.num_graphs = T_global
        self._emb_dev = emb
        self.singular_values = sigma
        self.solver_info = info
        # float64 container holding fp32-precision values, like se.UASE's YA
        self.node_embeddings = D.to_host(emb.double())           # widened on the device (exact)

    # ---- rank: replaces model.py:57-96 + model_utils.py:39-67 --------------------------------
    def __rank(self, pairwise_ranks_dict, comp_strategy):
        if comp_strategy not in N.STRATEGIES:
            raise UnboundLocalError("cannot access local variable 'embed_one' where it is not associated with a value")
        names, dims, mets = D.column_specs(self.embed_dimensions, self.distance_metrics)
        strategy = N.STRATEGIES[comp_strategy]
        scores = D.distances(self._emb_dev, dims, mets, strategy)        # (C, cols, n) float64
        host = D.to_host(scores)
        for c in range(scores.shape[0]):
            key = str(c + 1)                                             # '1'..'T-1' or '1'..'T'
            table = pd.DataFrame(host[c].T, index=self.node_names, columns=names, copy=False)   # (cols, n) block, no copy
            pairwise_ranks_dict.setdefault(key, []).append(table)

    def fit_transform_rank(self):
        """Differential scores of every node for every comparison: dict key -> DataFrame
        (index = node names, one column per (dim, metric)); writes <key>.tsv."""
        if self.config["verbose"] >= 0:
            print(f"\nRunning n2v2r with dimensions {self.embed_dimensions} and distance metrics {self.distance_metrics} ...")
        tic_n2v2r = time.time()
        pairwise_ranks_dict = {}
        tic_uase = time.time()
        self.__fit(graphs=self.graphs)
        toc_uase = time.time()
        if self.config["verbose"] == 1:
            print(f"""\tMulti-layer embedding in {round(toc_uase - tic_uase, 2)} seconds""")
        self.__rank(pairwise_ranks_dict=pairwise_ranks_dict, comp_strategy=self.comp_strategy)
        self.pairwise_ranks = {key: pd.concat(pairwise_ranks_dict[key], axis=1) for key in pairwise_ranks_dict}
        num_rankings = sum([len(self.pairwise_ranks[key].columns) for key in self.pairwise_ranks])
        toc_n2v2r = time.time()
        if self.config["verbose"] >= 0:
            print(f"""n2v2r computed {num_rankings} rankings for {len(self.pairwise_ranks)} comparison(s) in {round(toc_n2v2r - tic_n2v2r, 2)} seconds""")
        if self.config["save_dir"] and self._writes_files():
            for key, rank in self.pairwise_ranks.items():
                rank.to_csv(os.path.join(self.save_dir, key + ".tsv"), sep='\t', index=True)
        return self.pairwise_ranks

    # ---- Borda: replaces model.py:149-201 + model_utils.py:22-36 -----------------------------
    def aggregate_transform(self, method='Borda'):
        """Borda aggregation of every comparison's columns: dict key -> DataFrame('borda_ranks', int64)
        in node order; writes <key>_agg.tsv.  Ties are broken by node position (stable), NaN last."""
        if self.pairwise_ranks:
            out = {}
            start_time = time.time()
            if self.config["verbose"] >= 0:
                print("\nRank aggregation with Borda ...")
            dev = N.require_gpu()
            if method.casefold() != 'borda':
                raise NotImplementedError('Aggregation method not found. Available methods: Borda')
            keys = list(self.pairwise_ranks)
            tables = []
            for comparison_key in keys:
                # always rank what the (user-visible, possibly edited) table holds, like model.py:171-175
                table = self.pairwise_ranks[comparison_key]
                block = np.ascontiguousarray(table.to_numpy(dtype=np.float64).T)      # no copy for the tables __rank built
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")                                    # read-only view: only read for the H2D copy
                    tables.append(torch.from_numpy(block).to(dev))
            if len({tuple(t.shape) for t in tables}) == 1:
                bordas = D.rank_borda(torch.stack(tables))[0].cpu().numpy()          # every comparison in one launch
            else:
                bordas = [D.rank_borda(t)[0].cpu().numpy() for t in tables]
            for comparison_key, borda in zip(keys, bordas):
                out[comparison_key] = pd.DataFrame(borda, index=self.node_names, columns=['borda_ranks'])
            self.pairwise_aggregate_ranks = out
            exec_time_agg = round(time.time() - start_time, 2)
            if self.config["verbose"] == 1:
                print(f"\tFinished aggregation in {exec_time_agg} seconds")
            if self.config["save_dir"] and self._writes_files():
                for k, rank in self.pairwise_aggregate_ranks.items():
                    rank.to_csv(os.path.join(self.save_dir, k + "_agg.tsv"), sep='\t', index=True)
        else:
            raise ValueError("No n2v2r embeddings found")
        return self.pairwise_aggregate_ranks

    # ---- signed ranks: model.py:203-280; the sign transfer runs on the device (K11b) -----------
    def _signed_table(self, table, prior, dev):
        """signed_transform_single (model_utils.py:7-19) for every column of `table` at once: rows whose node is missing
        from the prior are dropped, the rest keep their score where the prior is > 0 and are negated otherwise."""
        keep = table.index.isin(prior.index)
        kept = table.index[keep]
        p = torch.from_numpy(np.ascontiguousarray(prior.loc[kept].to_numpy(dtype=np.float64))).to(dev)
        vals = table.to_numpy()
        is_int = np.issubdtype(vals.dtype, np.integer)
        block = np.ascontiguousarray(vals[keep].T.astype(np.int64 if is_int else np.float64))
        if block.shape[1] == 0:
            return kept, block.T
        out = D.signed_scores(torch.from_numpy(block).to(dev), p)
        return kept, D.to_host(out).T

    def signed_ranks_transform(self, prior_signed_ranks: pd.Series = None):
        if prior_signed_ranks is None:
            if self.prior_singed_ranks:
                prior_signed_ranks = self.prior_singed_ranks
            else:
                raise ValueError("""Prior signed ranks needed, run degree_difference_ranking beforehand
                                 or provide them in arguments.""")
        if self.pairwise_ranks:
            print("\nSigned ranks transformation ...")
            start_time = time.time()
            dev = N.require_gpu()
            signed, signed_agg = {}, {}
            for index, comparison_key in enumerate(self.pairwise_ranks):
                table = self.pairwise_ranks[comparison_key]
                prior = prior_signed_ranks[index]
                kept, vals = self._signed_table(table, prior, dev)
                signed[comparison_key] = pd.DataFrame(vals, index=list(kept), columns=table.columns, copy=False)
                if self.pairwise_aggregate_ranks:
                    agg = self.pairwise_aggregate_ranks[comparison_key].iloc[:, [0]]
                    kept_a, vals_a = self._signed_table(agg, prior, dev)
                    signed_agg[comparison_key] = pd.DataFrame(vals_a[:, 0], index=list(kept_a), columns=["signed_agg_ranks"])
        self.pairwise_signed_ranks = signed
        if self.pairwise_aggregate_ranks:
            self.pairwise_signed_aggregate_ranks = signed_agg
        exec_time_signed = round(time.time() - start_time, 2)
        if self.config["verbose"] == 1:
            print(f"\tFinished signed transformation in {exec_time_signed} seconds")
        if self.config["save_dir"] and self._writes_files():
            for k in self.pairwise_signed_ranks:
                self.pairwise_signed_ranks[k].to_csv(os.path.join(self.save_dir, k + "_signed.tsv"), sep='\t', index=True)
            if self.pairwise_aggregate_ranks:
                for k in self.pairwise_signed_aggregate_ranks:
                    self.pairwise_signed_aggregate_ranks[k].to_csv(os.path.join(self.save_dir, str(k) + "_agg_signed.tsv"),
                                                                   sep='\t', index=True)
        return self.pairwise_signed_ranks

    # ---- DeDi: model.py:282-311; the column sums are a by-product of the load pass (K1b) -----
    def degree_difference_ranking(self):
        layers = self._device_layers()
        sums = []
        for L in layers:
            if L.colsum is None:
                L.colsum = D.column_sums(L)
            sums.append(L.colsum)
        if bool(self.config.get("layer_sharded")) and dist.is_distributed():
            sums = list(dist.gather_layers(torch.stack(sums)[:, :, None], self._global_layers(layers)[0])[:, :, 0])
        sums = [c.cpu().numpy().astype(np.float32) for c in sums]           # np.sum of an fp32 matrix is fp32
        pairwise = {}
        for i in range(1, len(sums)):
            DeDi = sums[i - 1] - sums[i]
            ranking = pd.DataFrame.from_dict({"DeDi": DeDi, "absDeDi": np.abs(DeDi)})
            ranking.index = self.node_names
            pairwise[str(i)] = ranking
        self.prior_singed_ranks = [v.iloc[:, 0] for k, v in pairwise.items()]
        if self.config["save_dir"] and self._writes_files():
            for k, rank in pairwise.items():
                rank.to_csv(os.path.join(self.save_dir, k + "_degDif.tsv"), sep='\t', index=True)
        return pairwise
