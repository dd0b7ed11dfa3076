"""Velocity sets on the host (numpy), for building synthetic states and
decomposition-independent inputs. Ordering/weights are interface facts of the
reference: src/lb_d3q19.h:27-40, src/lb_d3q27.h:27-43, src/lb_d2q9.h:26-35."""

import numpy as np

_CV19 = [(0, 0, 0),
         (1, 1, 0), (1, 0, 1), (1, 0, 0), (1, 0, -1), (1, -1, 0), (0, 1, 1),
         (0, 1, 0), (0, 1, -1), (0, 0, 1), (0, 0, -1), (0, -1, 1), (0, -1, 0),
         (0, -1, -1), (-1, 1, 0), (-1, 0, 1), (-1, 0, 0), (-1, 0, -1), (-1, -1, 0)]
_CV9 = [(0, 0, 0), (1, 1, 0), (1, 0, 0), (1, -1, 0), (0, 1, 0), (0, -1, 0),
        (-1, 1, 0), (-1, 0, 0), (-1, -1, 0)]


def cv(nvel):
    if nvel == 19:
        return np.array(_CV19, dtype=np.int64)
    if nvel == 9:
        return np.array(_CV9, dtype=np.int64)
    if nvel == 27:
        out = [(0, 0, 0)]
        for x in (-1, 0, 1):
            for y in (-1, 0, 1):
                for z in (-1, 0, 1):
                    if (x, y, z) != (0, 0, 0):
                        out.append((x, y, z))
        return np.array(out, dtype=np.int64)
    raise ValueError("nvel must be 9, 19 or 27")


def wv(nvel):
    c2 = (cv(nvel) ** 2).sum(axis=1)
    table = {19: [12 / 36, 2 / 36, 1 / 36], 27: [64 / 216, 16 / 216, 4 / 216, 1 / 216],
             9: [16 / 36, 4 / 36, 1 / 36]}[nvel]
    return np.array([table[c] for c in c2])
