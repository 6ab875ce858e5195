"""Same import surface as the reference's tffm/fm_ops.py; implemented by
fast_tffm_b200.fm_ops (torch custom ops over libfm_b200.so)."""
from fast_tffm_b200.fm_ops import (FmParseError, PackedLines, fm_grad, fm_loss_grad, fm_lookup,  # noqa: F401
                                   fm_parser, fm_poses, fm_scorer, fm_sparse_adagrad, fm_unique, pack_lines)
