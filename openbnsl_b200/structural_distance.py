"""Structural distances between learned and true graphs, without pgmpy / networkx.

Mirror of the reference's ``helpers/structural_distance.py`` (``PDAG2CPDAG`` :43-76,
``structural_errors`` :79-137) for graphs over variable indices: an argument is a
``base.PDAG`` of this package (or of the reference module: anything with ``num_vars`` and
``has_edge(from, to)``) or an ``n x n`` 0/1 matrix with ``adj[u, v] != 0  <=>  u -> v`` (an
undirected edge is both arcs).  The reference compares every learned graph with the truth
through these two functions (``pytest/structure_learning/test_pc.py:31-33``,
``test_exhaustive_search.py:22-30``); SURVEY.md section 8f rank 2 asks for them so that the
CPDAG-level parity harness needs no pgmpy.

Matrix formulation instead of the reference's edge-set loops: with D the strictly directed
arcs and A the adjacency (either direction), the "CPDAG" the reference builds keeps exactly
the arcs that take part in an unshielded collider Y -> X <- Z (Y, Z strictly directed
parents of X, not adjacent) and makes every other edge undirected -- it does NOT propagate
Meek's rules (use ``structure_learning.apply_meeks_rules`` for that).
"""
import numpy as np

__all__ = ["adjacency", "PDAG2CPDAG", "pdag2cpdag", "structural_errors"]


def adjacency(g):
    """0/1 matrix of a PDAG-like object (``num_vars`` + ``has_edge``) or of a square array."""
    if hasattr(g, "has_edge") and hasattr(g, "num_vars"):
        n = int(g.num_vars)
        return np.array([[bool(g.has_edge(u, v)) for v in range(n)] for u in range(n)], dtype=bool).reshape(n, n)
    a = np.asarray(g)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError("expected a PDAG or a square adjacency matrix")
    a = a.astype(bool)
    np.fill_diagonal(a, False)          # the reference ignores self loops (u != v, :27-31)
    return a


def pdag2cpdag(g):
    """Skeleton + unshielded colliders of ``g`` as a 0/1 matrix (``PDAG2CPDAG``, :43-76)."""
    a = adjacency(g)
    directed = a & ~a.T                  # strictly directed arcs (pdag_removed, :58-64)
    adjacent = a | a.T
    out = adjacent.copy()                # every edge in both directions (:52-56)
    # collider[y, x]: y -> x and some other strict parent z of x is not adjacent to y (:66-75)
    d = directed.astype(np.int64)
    spouses = (~adjacent).astype(np.int64)
    np.fill_diagonal(spouses, 0)
    collider = directed & ((spouses @ d) > 0)
    out &= ~collider.T                   # drop x -> y where y -> x is a collider arc
    return out.astype(np.uint8)


def PDAG2CPDAG(g):
    """Same name as the reference helper; returns the 0/1 matrix of :func:`pdag2cpdag`."""
    return pdag2cpdag(g)


def structural_errors(ground_truth, predicted):
    """SHD and its parts between two graphs, each first mapped through :func:`pdag2cpdag`.

    Keys as the reference (:79-137): SHD = EE + ME + DE, DE = ED + MD + RD;
    ME/EE missing / extra skeleton edges, ED undirected in truth but directed in the prediction,
    MD directed in truth but undirected in the prediction, RD reversed.
    """
    gt = pdag2cpdag(ground_truth).astype(bool)
    pr = pdag2cpdag(predicted).astype(bool)
    if gt.shape != pr.shape:
        raise ValueError("Node sets differ. n_true=%d, n_pred=%d" % (gt.shape[0], pr.shape[0]))
    upper = np.triu(np.ones_like(gt, dtype=bool), 1)
    skel_t, skel_p = gt | gt.T, pr | pr.T
    und_t, und_p = gt & gt.T, pr & pr.T
    dir_t, dir_p = gt & ~gt.T, pr & ~pr.T
    me = int((skel_t & ~skel_p & upper).sum())
    ee = int((skel_p & ~skel_t & upper).sum())
    ed = int((und_t & (dir_p | dir_p.T) & upper).sum())
    md = int((dir_t & und_p).sum())
    rd = int((dir_t & dir_p.T).sum())
    de = ed + md + rd
    return {"SHD": me + ee + de, "ME": me, "EE": ee, "DE": de, "ED": ed, "MD": md, "RD": rd}
