import sys, numpy as np, torch
sys.path.insert(0, ".")
from tests.slab_parity import run
out, status, ref = run(0, 1, torch.device("cuda:0"))
for k in ("u", "v", "w", "th", "p"):
    d = np.abs(out[k] - ref[k])
    idx = np.argwhere(d > 0)
    print(k, d.max(), len(idx), idx[:5].tolist(), idx[-3:].tolist() if len(idx) else "")
print(status, ref["status"])
