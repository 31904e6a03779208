"""Mirror of node2vec2rank/dataloader.py (DataLoader :14-109) with the per-graph transform
(:70-76) running on the B200.  Same config keys, attributes, prints and errors.

Additive config keys (all optional, defaults keep the reference behaviour):
  device_resident (bool, default False): keep the transformed graphs in HBM as DeviceLayer
      objects (bf16 planes) instead of float32 ndarrays; N2V2R accepts both.
  storage_planes (int 1..3 | None): bf16 planes per weighted layer (None = as many as needed
      to be exact; binarised layers always use one).
  edge_list_parser ("device" | "networkx", default "device"): edge lists are parsed by pandas and scattered into the
      dense matrix on the device (K0b); "networkx" keeps the reference's nx.read_weighted_edgelist path.
"""
from __future__ import annotations

import os
import time

import numpy as np
import pandas as pd

from .preprocessing_utils import match_networks, network_transform, network_transform_device


class DataLoader():
    def __init__(self, config):
        self.config = config
        self.graphs = []
        self.interest_nodes = []
        self.__graph_filenames = config["graph_filenames"]
        self.__load_graphs()

    def get_graphs(self):
        return self.graphs

    def get_nodes(self):
        return self.interest_nodes

    def __load_graphs(self):
        tic_loading = time.time()
        print('Loading graphs in memory ...')
        if self.config["is_edge_list"] and self.config.get("edge_list_parser", "device") == "device":
            row_nodes = col_nodes = self.__load_edge_lists_device()
        else:
            for i, graph_filename in enumerate(self.__graph_filenames):
                self.graphs.append(self.__load_graph(graph_filename=graph_filename, graph_index=i))
            self.graphs = match_networks(self.graphs)
            row_nodes = self.graphs[0].index.to_numpy()
            col_nodes = self.graphs[0].columns.to_numpy()

        if np.size(row_nodes) != np.size(col_nodes):
            if self.config["project_unipartite_on"].casefold() == 'rows':
                self.interest_nodes = row_nodes
                print('\tGraphs are non-square and will be projected on row nodes')
            elif self.config["project_unipartite_on"].casefold() == 'columns':
                self.interest_nodes = col_nodes
                print('\tGraphs are non-square and will be projected on column nodes')
            else:
                raise ValueError("Impossible transformation")
        else:
            self.interest_nodes = col_nodes

        n = np.size(self.interest_nodes)
        print(f"""\tThere are {n} common nodes and resulting networks will have size {n} by {n}""")

        cfg = self.config
        if any([cfg['binarize'], cfg['threshold'] is not None, cfg['absolute'],
                int(cfg['top_percent_keep']) != 100, cfg['project_unipartite_on'] is not None]):
            print('Transforming graphs ...')
        variables = [("absoluting", cfg['absolute']), ("thresholding", cfg['threshold'] is not None),
                     ('to_unipartite', cfg['project_unipartite_on'] is not None),
                     ('sparcifying', int(cfg['top_percent_keep']) != 100), ("binarize", cfg['binarize'])]
        for var_name, var_value in variables:
            if var_value:
                print(f"\t {var_name}")

        kwargs = dict(binarize=cfg['binarize'], threshold=cfg['threshold'], absolute=cfg['absolute'],
                      top_percent_keep=cfg['top_percent_keep'], project_unipartite_on=cfg['project_unipartite_on'])
        if cfg.get("device_resident", False):
            self.graphs = [network_transform_device(g, storage_planes=cfg.get("storage_planes"), **kwargs)[0]
                           for g in self.graphs]
        else:
            self.graphs = [network_transform(g, **kwargs) for g in self.graphs]

        toc_loading = time.time()
        print(f"""Finished loading in {round(toc_loading - tic_loading, 2)} seconds \n""")

    def __load_edge_lists_device(self):
        """Edge-list files -> dense float64 matrices on the device, already restricted to the common nodes.

        Replaces nx.read_weighted_edgelist + nx.to_numpy_array + DataFrame + match_networks (dataloader.py:100-104,
        preprocessing_utils.py:291-306) for edge lists: the file is parsed by pandas' C reader, node labels are numbered in
        order of first appearance (networkx's node order: u then v, line by line), duplicate lines of an undirected edge
        are resolved on the host (the last one wins, as nx.Graph.add_edge does) and K0b scatters the edges straight into
        the matrix of the final node order -- no per-graph dense intermediate, no host reindex.  Returns the node labels."""
        from . import device as D
        parsed = []
        for i, graph_filename in enumerate(self.__graph_filenames):
            path = os.path.join(self.config["data_dir"], graph_filename)
            sep = self.config["separator"]
            try:
                df = pd.read_csv(path, sep=sep if sep is not None else r"\s+", header=None, comment="#", dtype=str,
                                 keep_default_na=False, na_filter=False, skip_blank_lines=True)
            except pd.errors.EmptyDataError:
                df = pd.DataFrame({0: [], 1: [], 2: []}, dtype=str)
            except pd.errors.ParserError as exc:                            # ragged lines: networkx raises IndexError
                raise IndexError(f"Edge data and data_keys (('weight', float),) are not the same length: {exc}") from exc
            if df.shape[1] != 3:
                raise IndexError(f"Edge data {list(range(max(df.shape[1] - 2, 0)))} and data_keys (('weight', float),) are not the same length")
            try:
                w = df[2].astype(np.float64).to_numpy()
            except ValueError as exc:
                raise TypeError(f"Failed to convert weight data to type {float}.") from exc
            ends = np.column_stack([df[0].to_numpy(dtype=object), df[1].to_numpy(dtype=object)])
            codes, nodes = pd.factorize(ends.ravel())                      # order of first appearance: u0, v0, u1, v1, ...
            iu, iv = codes[0::2], codes[1::2]
            lo, hi = np.minimum(iu, iv), np.maximum(iu, iv)
            keep = ~pd.DataFrame({"a": lo, "b": hi}).duplicated(keep="last").to_numpy()
            parsed.append((pd.Index(nodes), lo[keep], hi[keep], w[keep]))
            print(f"""\tThere are {len(nodes)} row nodes and {len(nodes)} column nodes in graph {i+1}""")
        common = set.intersection(*[set(p[0].to_list()) for p in parsed])
        final = pd.Index([x for x in parsed[0][0].to_list() if x in common])
        n = len(final)
        for nodes, lo, hi, w in parsed:
            pos = final.get_indexer(nodes).astype(np.int32)               # -1: node not shared by every graph
            self.graphs.append(D.dense_from_edges(pos[lo], pos[hi], w, n) if n else np.zeros((0, 0)))
        return final.to_numpy()

    def __load_graph(self, graph_filename, graph_index):
        extension_type = graph_filename.split(".")[-1]
        path = os.path.join(self.config["data_dir"], graph_filename)
        if not self.config["is_edge_list"]:
            if extension_type != "h5":
                graph_pd = pd.read_csv(path, index_col=0, header=0, sep=self.config["separator"])
            else:
                graph_pd = pd.read_hdf(path)
        else:
            import networkx as nx
            edge_list_graph = nx.read_weighted_edgelist(path, delimiter=self.config["separator"], nodetype=str)
            adj_matrix = nx.to_numpy_array(edge_list_graph)
            graph_pd = pd.DataFrame(adj_matrix, index=edge_list_graph.nodes, columns=edge_list_graph.nodes)

        graph_pd = graph_pd.T if self.config["transpose"] else graph_pd
        num_rows, num_cols = graph_pd.shape
        print(f"""\tThere are {num_rows} row nodes and {num_cols} column nodes in graph {graph_index+1}""")
        return graph_pd
