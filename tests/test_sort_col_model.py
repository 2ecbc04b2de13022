"""Host-side check of the index math of the column-counter radix pass (csrc/sort_col.cuh) on its
numpy model (tools/sort_col_model.py): counter layout, packed-byte scan with the 256-keys-in-one-group
corner, 16-bit run-start table, XOR swizzle, partial tiles.  No GPU needed."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tools"))


def test_tile_model_is_a_stable_partition():
    import sort_col_model
    sort_col_model.self_test()
