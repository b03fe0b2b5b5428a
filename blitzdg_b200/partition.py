"""Element partitioning for multi-GPU runs."""
import numpy as np


def rcb_partition(cx, cy, nparts):
    """Recursive coordinate bisection of element centroids into `nparts` balanced, contiguous parts."""
    K = len(cx)
    part = np.zeros(K, dtype=np.int32)

    def rec(idx, p0, n):
        if n == 1:
            part[idx] = p0
            return
        nl = n // 2
        xs, ys = cx[idx], cy[idx]
        coord = xs if (xs.max() - xs.min()) >= (ys.max() - ys.min()) else ys
        order = np.argsort(coord, kind="stable")
        cut = (len(idx) * nl) // n
        rec(idx[order[:cut]], p0, nl)
        rec(idx[order[cut:]], p0 + nl, n - nl)

    rec(np.arange(K), 0, int(nparts))
    return part
