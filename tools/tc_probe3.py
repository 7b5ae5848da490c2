import sys

import numpy as np

sys.path.insert(0, ".")
sys.argv = [sys.argv[0], "none"] + sys.argv[1:]
from tools.tc_probe import run, w_img  # noqa: E402

lbo, sbo, a_mn = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
N, K = 16, 8
W = np.zeros((N, K), np.float32)
for k in range(K):
    W[k, k] = 1.0
bimg = w_img(W)
n_words = 16384
lo = (np.arange(n_words) % 1024 + 1).astype(np.float32)
hi = (np.arange(n_words) // 1024 + 1).astype(np.float32)
dlo = run(lo, bimg, [lbo, sbo, 128, N * 16, 128, 2 * N * 16], a_mn, 0, N, 1)
dhi = run(hi, bimg, [lbo, sbo, 128, N * 16, 128, 2 * N * 16], a_mn, 0, N, 1)
word = ((dhi - 1) * 1024 + (dlo - 1)).astype(np.int64)
np.set_printoptions(linewidth=200)
print("a_mn", a_mn, "lbo", lbo, "sbo", sbo, " nonzero D:", int((dlo != 0).sum()), "cols>=8 nonzero:", int((dlo[:, 8:] != 0).sum()))
print("word index read for (m, k), m = 0..11:")
print(word[:12, :8])
print("m = 32..35, 64..65, 127:")
print(word[[32, 33, 34, 35, 64, 65, 127], :8])
