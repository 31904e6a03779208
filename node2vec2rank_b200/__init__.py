"""node2vec2rank_b200 -- B200-native UASE embed-and-rank path of pmandros/node2vec2rank.

Host-side mirror of the reference's Python surface (DataLoader, N2V2R, network_transform,
compute_pairwise_distances, borda_aggregate_parallel) over hand-written sm_100a kernels in
libn2v2r_b200.so (include/n2v2r_b200.h).  No CPU fallback: compute entry points raise when
the library or a B200 is missing.
"""
from . import _native

__version__ = "0.1.0"


def __getattr__(name):          # lazy: importing the package must not need torch/pandas
    if name == "DataLoader":
        from .dataloader import DataLoader
        return DataLoader
    if name == "N2V2R":
        from .model import N2V2R
        return N2V2R
    if name in ("network_transform", "match_networks", "bipartite_to_unipartite_projection"):
        from . import preprocessing_utils
        return getattr(preprocessing_utils, name)
    if name in ("compute_pairwise_distances", "borda_aggregate_parallel", "signed_transform_single"):
        from . import model_utils
        return getattr(model_utils, name)
    raise AttributeError(name)
