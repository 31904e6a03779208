"""Drop-in for `python node2vec2rank/node2vec2rank.py --config <json>` (reference node2vec2rank.py:9-81)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from node2vec2rank_b200.cli import main  # noqa: E402

if __name__ == "__main__":
    sys.exit(main())
