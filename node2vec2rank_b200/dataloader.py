"""Mirror of node2vec2rank/dataloader.py (DataLoader :14-109) with the per-graph transform
(:70-76) running on the B200.  Same config keys, attributes, prints and errors.

Additive config keys (all optional, defaults keep the reference behaviour):
  device_resident (bool, default False): keep the transformed graphs in HBM as DeviceLayer
      objects (bf16 planes) instead of float32 ndarrays; N2V2R accepts both.
  storage_planes (int 1..3 | None): bf16 planes per weighted layer (None = as many as needed
      to be exact; binarised layers always use one).
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
