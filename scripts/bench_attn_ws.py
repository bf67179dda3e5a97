"""A/B of the experimental weight-stationary (tcgen05.mma.ws, M=32) decode-attention path vs the default."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from scripts.bench_attn import bench  # noqa: E402

for shape in ((64, 16, 16, 256, 1040), (64, 32, 32, 128, 1040), (64, 32, 8, 128, 4096), (64, 2, 2, 256, 1040)):
    for ws in ("0", "1", "0", "1"):
        os.environ["LLMSS_ATTN_WS"] = ws
        print("ws=" + ws, end=" ")
        bench(*shape, bs=64)
