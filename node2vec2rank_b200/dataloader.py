"""Mirror of node2vec2rank/dataloader.py (DataLoader :14-109) with the per-graph transform
(:70-76) running on the B200.  Same config keys, attributes, prints and errors.

Additive config keys (all optional, defaults keep the reference behaviour):
  device_resident (bool, default False): keep the transformed graphs in HBM as DeviceLayer
      objects (bf16 planes) instead of float32 ndarrays; N2V2R accepts both.
  storage_planes (int 1..3 | None): bf16 planes per weighted layer (None = as many as needed
      to be exact; binarised layers always use one).
  csv_parser ("device" | "pandas", default "device"): adjacency files are parsed in row chunks through pinned memory into
      device matrices and reindexed to the common nodes on the device (K0d); "pandas" keeps the reference's whole-file
      pd.read_csv + host reindex.  .h5 files always take the pandas path.
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
        elif (not self.config["is_edge_list"] and self.config.get("csv_parser", "device") == "device"
              and all(fn.split(".")[-1] != "h5" for fn in self.__graph_filenames)):
            row_nodes, col_nodes = self.__load_adjacency_device()
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
        pending, self.graphs = self.graphs, []
        while pending:                      # one graph at a time: device-side inputs are built, transformed and released in turn
            g = pending.pop(0)
            if callable(g):
                g = g()
            if cfg.get("device_resident", False):
                from . import device as D
                dims = cfg.get("embed_dimensions") or []
                hint = cfg.get("block_width") or (D.default_block_width(max(dims), g.shape[1], "rows") if len(dims) else None)
                self.graphs.append(network_transform_device(g, storage_planes=cfg.get("storage_planes"),
                                                            storage=cfg.get("layer_storage", "auto"), block_width=hint, **kwargs)[0])
            else:
                self.graphs.append(network_transform(g, **kwargs))
            del g

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
        if n == 0:
            raise ValueError("the graphs have no nodes in common")
        for nodes, lo, hi, w in parsed:
            pos = final.get_indexer(nodes).astype(np.int32)               # -1: node not shared by every graph
            # a thunk: the dense float64 matrix of a graph exists only while the transform loop works on it
            self.graphs.append(lambda pos=pos, lo=lo, hi=hi, w=w: D.dense_from_edges(pos[lo], pos[hi], w, n))
        return final.to_numpy()

    def __load_adjacency_device(self):
        """Adjacency files -> float64 device matrices restricted to the common nodes, in the order of the first graph.

        Replaces pd.read_csv (whole file, float64 DataFrame) + optional .T + match_networks' .loc reindex
        (dataloader.py:86-91, :103, preprocessing_utils.py:291-306): device.csv_to_device parses row chunks through pinned
        staging buffers into a device matrix in file order, K0d (device.gather_submatrix) picks / reorders / transposes the
        common rows and columns on the device.  The entries of self.graphs are thunks: each graph is reindexed (and, when the
        files are too large to sit in HBM together, parsed) only when the transform loop asks for it, so at most one
        float64 intermediate is alive besides the outputs.  Duplicate labels fall back to the pandas path (.loc semantics)."""
        import torch
        from . import device as D
        sep, transpose = self.config["separator"], bool(self.config["transpose"])
        paths = [os.path.join(self.config["data_dir"], fn) for fn in self.__graph_filenames]
        sizes = [os.path.getsize(p) for p in paths]
        free = torch.cuda.mem_get_info()[0] if torch.cuda.is_available() else 0
        keep_all = 4.0 * sum(sizes) < 0.4 * free       # a float64 cell is 8 bytes, its text rarely under 2: all files fit together
        parsed = []
        for i, path in enumerate(paths):
            if keep_all:
                x, rows, cols = D.csv_to_device(path, sep)
            else:                                      # labels only now (header + first column); the cells are parsed later
                x = None
                cols = pd.read_csv(path, index_col=0, header=0, sep=sep, nrows=0).columns
                rows = pd.Index(pd.read_csv(path, usecols=[0], header=0, sep=sep).iloc[:, 0])
            if transpose:
                rows, cols = cols, rows
            if not (rows.is_unique and cols.is_unique):
                del parsed, x
                self.graphs = []
                for j, fn in enumerate(self.__graph_filenames):
                    self.graphs.append(self.__load_graph(graph_filename=fn, graph_index=j))
                self.graphs = match_networks(self.graphs)
                return self.graphs[0].index.to_numpy(), self.graphs[0].columns.to_numpy()
            print(f"""\tThere are {len(rows)} row nodes and {len(cols)} column nodes in graph {i+1}""")
            parsed.append([x, rows, cols, path])
        common_cols = set.intersection(*[set(p[2].to_list()) for p in parsed])
        common_rows = set.intersection(*[set(p[1].to_list()) for p in parsed])
        final_rows = parsed[0][1][parsed[0][1].isin(common_rows)]
        final_cols = parsed[0][2][parsed[0][2].isin(common_cols)]
        if len(final_rows) == 0 or len(final_cols) == 0:
            raise ValueError("the graphs have no nodes in common")

        def thunk(entry):
            def build():
                x, rows, cols, path = entry
                if x is None:
                    x = D.csv_to_device(path, sep)[0]
                entry[0] = None                        # the file-order matrix is released once the graph is built
                ri, ci = rows.get_indexer(final_rows), cols.get_indexer(final_cols)
                same = (not transpose and len(ri) == x.shape[0] and len(ci) == x.shape[1]
                        and np.array_equal(ri, np.arange(len(ri))) and np.array_equal(ci, np.arange(len(ci))))
                return x.contiguous() if same else D.gather_submatrix(x, ri, ci, transpose=transpose)
            return build

        self.graphs = [thunk(p) for p in parsed]
        return final_rows.to_numpy(), final_cols.to_numpy()

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
