#!/usr/bin/env python3
"""Regenerate configuration/elasticconnections.txt for a scene whose file is missing (the reference ships
configuration/position.txt + velocity.txt — "config B'", 49 153 particles — but its elasticconnections.txt is a lost
large blob, /root/reference/.MISSING_LARGE_BLOBS).

Rule = the reference generator's geometric spring rule (/root/reference/src/owHelper.cpp:972-1001, 1385-1387), without
its worm-specific muscle colouring: for every elastic particle i, scan the NON-liquid particles j != i in index order;
each j with r_ij <= r0*sqrt(2.7) (float arithmetic as there) becomes a spring (j + 0.1, r_ij*simulationScale*0.95, 0, 0),
at most MAX_NEIGHBOR_COUNT = 32 per particle; unused rows are (-1, 0, 0, 0). Output in the loader's text format
(owHelper.cpp:1511-1540): 32 rows per elastic particle, elastic particles in index order.

  python tools/make_connections.py position.txt elasticconnections.txt      (or import and call connections(pos))
"""
import sys

import numpy as np

H = np.float32(3.34)
R0 = np.float32(0.5) * H
MAXN = 32


def simulation_scale():
    mass = np.float32(3.25e-14)
    return np.float32(0.004) * np.power(mass, np.float32(1.0 / 3.0), dtype=np.float32) / np.power(np.float32(0.00025), np.float32(1.0 / 3.0), dtype=np.float32)


def connections(pos):
    """pos: (N, 4) float32 (x, y, z, type). Returns (32 * Ne, 4) float32, elastic particles in index order."""
    from scipy.spatial import cKDTree
    pos = np.asarray(pos, np.float32)
    t = pos[:, 3].astype(np.int32)
    elastic = np.nonzero(t == 2)[0]
    cand = np.nonzero(t != 1)[0]                      # liquid particles are skipped as spring candidates (:986)
    tree = cKDTree(pos[cand, :3].astype(np.float64))
    limit = np.float32(R0 * np.float32(np.sqrt(2.7)))  # `r0*sqrt(2.7)`: double sqrt, float product
    limit = np.float32(np.float64(R0) * np.sqrt(2.7))
    scale = simulation_scale()
    out = np.zeros((MAXN * len(elastic), 4), np.float32)
    out[:, 0] = -1.0
    near = tree.query_ball_point(pos[elastic, :3].astype(np.float64), float(limit) * 1.001)
    for k, i in enumerate(elastic):
        js = np.sort(cand[np.asarray(near[k], np.int64)])
        js = js[js != i]
        d = pos[i, :3] - pos[js, :3]                   # float32, as the generator: dx2 = (xi - xj); dx2 *= dx2
        r = np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]).astype(np.float64)).astype(np.float32)
        keep = r <= limit
        js, r = js[keep][:MAXN], r[keep][:MAXN]
        rows = out[MAXN * k: MAXN * k + len(js)]
        rows[:, 0] = js.astype(np.float32) + np.float32(0.1)
        rows[:, 1] = (r.astype(np.float64) * float(scale) * 0.95).astype(np.float32)   # r_ij*simulationScale*0.95 in double
    return out


def write_txt(path, con):
    with open(path, "w") as f:
        for r in con:
            f.write("%e\t%e\t%e\t%e\n" % tuple(float(v) for v in r))


if __name__ == "__main__":
    p = np.loadtxt(sys.argv[1], dtype=np.float32).reshape(-1, 4)
    c = connections(p)
    write_txt(sys.argv[2], c)
    print("%d elastic particles, %d springs" % (len(c) // MAXN, int((c[:, 0] >= 0).sum())))
