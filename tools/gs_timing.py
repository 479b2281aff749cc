"""Forward 3DGS rasteriser: timing of the bench's gs_render block (ours and the reference build).  python tools/gs_timing.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench_extra import gs_render_block  # noqa: E402

kf = int(sys.argv[1]) if len(sys.argv) > 1 else 4
print(json.dumps(gs_render_block(keyframes=kf)))
