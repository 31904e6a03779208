"""Command line of the reference (node2vec2rank/node2vec2rank.py:9-81): `--config <json>`, the JSON's
sections flattened into one dict, then load -> fit/rank -> Borda -> DeDi with the same prints and files."""
from __future__ import annotations

import argparse
import json
import sys


def load_config(path):
    with open(path, 'r', encoding='utf-8') as f:
        sections = json.load(f)
    return {param: value for _, params in sections.items() for param, value in params.items()}


def main(argv=None):
    parser = argparse.ArgumentParser(description="Node2Vec2Rank (B200): embed multi-layer graphs with UASE and rank nodes")
    parser.add_argument('--config', help='path to the config file; every other option comes from the file')
    args = parser.parse_args(argv)
    if args.config is None:
        print("The following argument is required: --config")
        parser.print_help()
        return 1
    config = load_config(args.config)
    from .dataloader import DataLoader
    from .model import N2V2R
    dataloader = DataLoader(config=config)
    model = N2V2R(graphs=dataloader.get_graphs(), config=config, nodes=dataloader.get_nodes())
    model.fit_transform_rank()
    model.aggregate_transform()
    model.degree_difference_ranking()
    return 0


if __name__ == "__main__":
    sys.exit(main())
