"""Drop-in module names of the reference (`from tffm.fm_ops import fm_parser, fm_scorer`,
`from tffm.fm_model import Model`), served by the B200 implementation in fast_tffm_b200."""
