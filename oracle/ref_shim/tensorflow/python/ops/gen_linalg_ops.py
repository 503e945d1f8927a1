import numpy as np


def qr(x, full_matrices=False, name=None):
    q, r = np.linalg.qr(np.asarray(x, dtype=np.float64), mode="complete" if full_matrices else "reduced")
    return q.astype(np.float32), r.astype(np.float32)
