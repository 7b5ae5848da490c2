"""Which smem word does the tensor core read for logical A(m, k) in MN-major no-swizzle mode?"""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.argv = [sys.argv[0], "none"] + sys.argv[1:]
from tools.tc_probe import run, w_img  # noqa: E402

lbo, sbo = int(sys.argv[2]), int(sys.argv[3])
N, K = 16, 8
W = np.zeros((N, K), np.float32)
for k in range(K):
    W[k, k] = 1.0
bimg = w_img(W)
words = list(range(0, 12)) + [16, 28, 32, 36, 60, 64, 128, 256, 512, 516, 544, 1024, 2048, 4096]
for w in words:
    a = np.zeros(16384, np.float32)
    a[w] = 1.0
    d = run(a, bimg, [lbo, sbo, 128, N * 16, 128, 2 * N * 16], 1, 0, N, 1)
    nz = np.argwhere(np.abs(d) > 0.5)
    print(f"word {w:5d} (byte {4*w:6d}) -> (m,k) = {[tuple(int(v) for v in x) for x in nz]}")
