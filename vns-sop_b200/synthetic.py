"""Synthetic scaled SOP instances (BASELINE.json config 5, SURVEY.md 8: "20k nodes / 2k clusters").

The reference ships nothing of this size; the generator follows the construction of its GTSP-derived
files (datasets/sop/*.sop): points in the plane, proximity clusters (or a random assignment, the `RND`
files), cluster 0 = a single depot node that is both START_SET and END_SET, profit = cluster size
(rule p1), EDGE_WEIGHT_TYPE CEIL_2D (or raw Euclid for a real-valued table), TMAX = a fraction of a
nearest-neighbour tour through all cluster centroids.  The result is an ordinary instances.SopFile:
write_sop() emits a file any reference-compatible loader reads, build_problem() gives the flat model.
"""
from __future__ import annotations

import numpy as np

from . import instances as ins

DEFAULT_SEED = 20190147


def nn_tour(centroids: np.ndarray, start: int = 0) -> np.ndarray:
    """Nearest-neighbour order over the centroids, starting at `start` (O(C^2), numpy)."""
    C = len(centroids)
    left = np.ones(C, bool)
    order = np.empty(C, np.int64)
    cur = start
    for k in range(C):
        order[k] = cur
        left[cur] = False
        if k + 1 == C:
            break
        d = ((centroids - centroids[cur]) ** 2).sum(1)
        d[~left] = np.inf
        cur = int(np.argmin(d))
    return order


def generate(n_points: int = 20000, cluster_size: int = 10, seed: int = DEFAULT_SEED, box: float = 1e4,
             random_clusters: bool = False, edge_weight_type: str = "CEIL_2D", budget_fraction: float = 0.4,
             name: str | None = None) -> ins.SopFile:
    rs = np.random.Generator(np.random.MT19937(seed))
    pts = np.floor(rs.uniform(0.0, box, size=(n_points, 2)))          # integer coordinates like the TSPLIB sources
    n_clusters = n_points // cluster_size
    if random_clusters:
        perm = rs.permutation(n_points)
    else:
        # proximity clusters: sort into a sqrt(C) x sqrt(C) grid (boustrophedon), cut into contiguous groups
        g = max(1, int(round(np.sqrt(n_clusters))))
        cell = np.minimum((pts / (box / g)).astype(np.int64), g - 1)
        col = np.where(cell[:, 1] % 2 == 0, cell[:, 0], g - 1 - cell[:, 0])
        perm = np.lexsort((pts[:, 0], col, cell[:, 1]))
    groups = [perm[k * cluster_size:(k + 1) * cluster_size] for k in range(n_clusters)]
    if n_clusters * cluster_size < n_points:
        groups[-1] = np.concatenate([groups[-1], perm[n_clusters * cluster_size:]])
    depot = np.floor(pts.mean(0))
    coords = np.vstack([depot[None, :], pts])                           # node 0 = depot
    set_nodes = [np.array([0], np.int32)] + [(gidx + 1).astype(np.int32) for gidx in groups]
    set_profit = np.array([0.0] + [float(len(gidx)) for gidx in groups])
    cent = np.vstack([coords[s].mean(0) for s in set_nodes])
    order = nn_tour(cent, 0)
    tour_len = float(np.sqrt(((cent[order] - cent[np.roll(order, -1)]) ** 2).sum(1)).sum())
    sf = ins.SopFile(name=name or "synth_%d_%d%s" % (n_points, n_clusters, "_RND" if random_clusters else ""),
                     comment="synthetic scaled SOP (seed %d)" % seed, tmax=int(budget_fraction * tour_len),
                     start_set=0, end_set=0, edge_weight_type=edge_weight_type, coords=coords,
                     set_profit=set_profit, set_nodes=set_nodes)
    return sf


def generate_dop(n_centres: int = 2000, headings: int = 32, seed: int = DEFAULT_SEED, box: float = 1e4,
                 radius_fraction: float = 0.005, budget_fraction: float = 0.4, name: str | None = None) -> ins.Problem:
    """DOP flavour of the scaled instance ("2000 centres x 32 headings", rho = 0.5 % of the bounding
    box): every cluster is one location with `headings` heading samples, cluster 0 = the depot (start and goal, so
    the goal set is duplicated at the end like SOPLoader does for g == s, sop_loader.cpp:118-156), profit = 1 per
    location.  The distance table is NOT materialised here (8 N^2 bytes = 33 GB at the default size): the Problem
    carries the centres and `dubins_radius`, and the engine builds the table on the device
    (vnsgpu_build_table_dubins).  Returns the flat Problem directly."""
    rs = np.random.Generator(np.random.MT19937(seed))
    ctr = np.floor(rs.uniform(0.0, box, size=(n_centres, 2)))
    depot = np.floor(ctr.mean(0))
    centers = np.vstack([depot[None, :], ctr, depot[None, :]])            # start, locations, goal (= the depot again)
    C = n_centres + 2
    m = np.full(C, headings, dtype=np.int32)
    off = np.zeros(C + 1, dtype=np.int32)
    off[1:] = np.cumsum(m)
    order = nn_tour(centers[:-1], 0)
    tour_len = float(np.sqrt(((centers[order] - centers[np.roll(order, -1)]) ** 2).sum(1)).sum())
    reward = np.ones(C)
    reward[0] = reward[-1] = 0.0
    node_id = np.arange(C * headings, dtype=np.int32)
    node_id[-headings:] = node_id[:headings]                               # the goal's samples are the depot's nodes
    return ins.Problem(name=name or "synth_dop_%d_x%d" % (n_centres, headings), C=C, m=m, off=off, reward=reward, start=0,
                       goal=C - 1, budget=float(int(budget_fraction * tour_len)), node_id=node_id,
                       orig_cluster_id=np.concatenate([np.arange(C - 1), [0]]).astype(np.int32), rule=-1,
                       centers=centers, dubins_radius=radius_fraction * box)


def centroid_tour_state(pb: ins.Problem, fraction: float = 0.85):
    """A long feasible-looking tour without running greedy_insertion (which is O(C^3) evaluations at this
    size): the nearest-neighbour order over the cluster centroids, cut where the centroid-to-centroid length
    reaches `fraction` of the budget.  Returns (vns vector, T) for VnsSopPath.set()."""
    cent = pb.centers if pb.xy is None else np.vstack([pb.xy[pb.off[c]:pb.off[c + 1]].mean(0) for c in range(pb.C)])
    inner = [c for c in range(pb.C) if c not in (pb.start, pb.goal)]
    order = nn_tour(cent[[pb.start] + inner], 0)[1:]                    # indices into [start] + inner
    seq = [inner[k - 1] for k in order]
    length, T, prev = 0.0, 0, cent[pb.start]
    for c in seq:
        step = float(np.sqrt(((cent[c] - prev) ** 2).sum()))
        back = float(np.sqrt(((cent[pb.goal] - cent[c]) ** 2).sum()))
        if length + step + back > fraction * pb.budget:
            break
        length += step
        prev = cent[c]
        T += 1
    vns = np.array(seq, dtype=np.int32)
    return vns, T
