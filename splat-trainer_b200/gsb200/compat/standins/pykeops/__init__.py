"""pykeops stand-in: import stub only (the calls behind it are replaced by gsb200: io.py / ops.kmeans / ops.estimate_scale)."""
