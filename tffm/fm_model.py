"""Same import surface as the reference's tffm/fm_model.py."""
from fast_tffm_b200.fm_model import Model  # noqa: F401
