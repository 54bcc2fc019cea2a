"""Synthetic discrete Bayesian networks and forward sampling.

The reference's tests draw data with ``pgmpy.utils.get_example_model(name).simulate(n, seed)``
and sort the columns by name (pytest/structure_learning/test_pc.py:17-19,
test_exhaustive_search.py:20-22).  pgmpy and the .bif files are not available
offline, so the named structures are restated here:

* ``asia``   -- textbook structure AND textbook CPTs (8 binary nodes, 8 arcs);
* ``alarm``  -- the 37-node / 46-arc structure with its cardinalities (509
  free parameters); CPT values are seeded Dirichlet(1) draws ("Alarm-structure
  synthetic data", BASELINE.json configs[1]);
* ``hepar2_like`` / ``munin1_like`` / ``random_binary`` -- seeded random DAGs
  matching the summary statistics of the named networks.

Everything is returned with variables in lexicographic name order, codes as
``uint8`` and state names chosen so that string order equals code order, which
is what the reference's DataframeWrapper would compute
(src/base/dataframe_wrapper.cpp:36-52).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Sequence, Tuple

import numpy as np


@dataclasses.dataclass
class Network:
    names: List[str]                 # sorted lexicographically
    cards: np.ndarray                # [n] int
    parents: List[List[int]]         # parents[v] = indices into names
    cpts: List[np.ndarray]           # cpts[v]: [prod(card[parents]), card[v]]

    @property
    def n(self) -> int:
        return len(self.names)

    def edges(self) -> List[Tuple[int, int]]:
        return [(p, v) for v in range(self.n) for p in self.parents[v]]

    def topological_order(self) -> List[int]:
        indeg = [len(p) for p in self.parents]
        children = [[] for _ in range(self.n)]
        for v, ps in enumerate(self.parents):
            for p in ps:
                children[p].append(v)
        order, stack = [], [v for v in range(self.n) if indeg[v] == 0]
        while stack:
            v = stack.pop()
            order.append(v)
            for c in children[v]:
                indeg[c] -= 1
                if indeg[c] == 0:
                    stack.append(c)
        if len(order) != self.n:
            raise ValueError("network has a cycle")
        return order

    def sample(self, n_rows: int, seed: int = 0) -> np.ndarray:
        """Forward-sample ``n_rows`` rows. Returns uint8 codes, shape [n_rows, n], C order."""
        rng = np.random.default_rng(seed)
        out = np.zeros((n_rows, self.n), dtype=np.uint8)
        for v in self.topological_order():
            cfg = np.zeros(n_rows, dtype=np.int64)
            for p in self.parents[v]:          # first parent most significant
                cfg = cfg * int(self.cards[p]) + out[:, p]
            cum = np.cumsum(self.cpts[v], axis=1)
            cum[:, -1] = 1.0
            u = rng.random(n_rows)
            out[:, v] = (u[:, None] >= cum[cfg][:, :-1]).sum(axis=1).astype(np.uint8)
        return out


def state_names(card: int) -> List[str]:
    """State labels whose lexicographic order equals code order."""
    width = len(str(card - 1))
    return ["s%0*d" % (width, i) for i in range(card)]


def _from_named(cards: Dict[str, int], arcs: Sequence[Tuple[str, str]],
                cpt_fn) -> Network:
    names = sorted(cards)
    idx = {s: i for i, s in enumerate(names)}
    parents: List[List[int]] = [[] for _ in names]
    for a, b in arcs:
        parents[idx[b]].append(idx[a])
    for ps in parents:
        ps.sort()
    card_arr = np.array([cards[s] for s in names], dtype=np.int64)
    cpts = [cpt_fn(v, names[v], [names[p] for p in parents[v]], card_arr) for v in range(len(names))]
    return Network(names, card_arr, parents, cpts)


# ----------------------------------------------------------------------------
# Asia (Lauritzen & Spiegelhalter 1988): states "no" < "yes" => code 0 = no.
# ----------------------------------------------------------------------------
_ASIA_ARCS = [("asia", "tub"), ("smoke", "lung"), ("smoke", "bronc"), ("tub", "either"),
              ("lung", "either"), ("either", "xray"), ("either", "dysp"), ("bronc", "dysp")]


def asia() -> Network:
    cards = {s: 2 for s in ("asia", "bronc", "dysp", "either", "lung", "smoke", "tub", "xray")}

    def cpt(v, name, pnames, card_arr):
        def row(p_yes):
            return [1.0 - p_yes, p_yes]
        if name == "asia":
            return np.array([row(0.01)])
        if name == "smoke":
            return np.array([row(0.5)])
        if name == "tub":      # parent asia: no, yes
            return np.array([row(0.01), row(0.05)])
        if name == "lung":     # parent smoke
            return np.array([row(0.01), row(0.1)])
        if name == "bronc":    # parent smoke
            return np.array([row(0.3), row(0.6)])
        if name == "either":   # parents sorted: lung, tub ; OR gate
            assert pnames == ["lung", "tub"]
            return np.array([row(0.0), row(1.0), row(1.0), row(1.0)])
        if name == "xray":     # parent either
            return np.array([row(0.05), row(0.98)])
        if name == "dysp":     # parents sorted: bronc, either
            assert pnames == ["bronc", "either"]
            return np.array([row(0.1), row(0.7), row(0.8), row(0.9)])
        raise KeyError(name)

    return _from_named(cards, _ASIA_ARCS, cpt)


# ----------------------------------------------------------------------------
# Alarm structure (Beinlich et al. 1989): 37 nodes, 46 arcs, 509 parameters.
# ----------------------------------------------------------------------------
_ALARM_CARDS = {}
for _s in ("HISTORY HYPOVOLEMIA LVFAILURE ERRLOWOUTPUT ERRCAUTER INSUFFANESTH ANAPHYLAXIS "
           "KINKEDTUBE FIO2 PULMEMBOLUS SHUNT DISCONNECT CATECHOL").split():
    _ALARM_CARDS[_s] = 2
for _s in ("CVP PCWP LVEDVOLUME STROKEVOLUME HRBP HREKG HRSAT TPR PVSAT SAO2 PAP INTUBATION "
           "MINVOLSET ARTCO2 HR CO BP").split():
    _ALARM_CARDS[_s] = 3
for _s in "EXPCO2 MINVOL PRESS VENTMACH VENTTUBE VENTLUNG VENTALV".split():
    _ALARM_CARDS[_s] = 4

_ALARM_ARCS = [tuple(a.split(">")) for a in (
    "LVFAILURE>HISTORY LVEDVOLUME>CVP LVEDVOLUME>PCWP HYPOVOLEMIA>LVEDVOLUME LVFAILURE>LVEDVOLUME "
    "HYPOVOLEMIA>STROKEVOLUME LVFAILURE>STROKEVOLUME ERRLOWOUTPUT>HRBP HR>HRBP ERRCAUTER>HREKG "
    "HR>HREKG ERRCAUTER>HRSAT HR>HRSAT ANAPHYLAXIS>TPR ARTCO2>EXPCO2 VENTLUNG>EXPCO2 "
    "INTUBATION>MINVOL VENTLUNG>MINVOL FIO2>PVSAT VENTALV>PVSAT PVSAT>SAO2 SHUNT>SAO2 "
    "PULMEMBOLUS>PAP INTUBATION>SHUNT PULMEMBOLUS>SHUNT INTUBATION>PRESS KINKEDTUBE>PRESS "
    "VENTTUBE>PRESS MINVOLSET>VENTMACH DISCONNECT>VENTTUBE VENTMACH>VENTTUBE INTUBATION>VENTLUNG "
    "KINKEDTUBE>VENTLUNG VENTTUBE>VENTLUNG INTUBATION>VENTALV VENTLUNG>VENTALV VENTALV>ARTCO2 "
    "ARTCO2>CATECHOL INSUFFANESTH>CATECHOL SAO2>CATECHOL TPR>CATECHOL CATECHOL>HR HR>CO "
    "STROKEVOLUME>CO CO>BP TPR>BP").split()]


def _dirichlet_cpts(seed: int, alpha: float = 1.0):
    rng = np.random.default_rng(seed)

    def cpt(v, name, pnames, card_arr):
        # one independent stream per node so structure edits do not shift draws
        q = 1
        for p in pnames:
            q *= int(_cards_lookup[p])
        return rng.dirichlet(np.full(int(card_arr[v]), alpha), size=q)

    _cards_lookup: Dict[str, int] = {}
    return cpt, _cards_lookup


def alarm(cpt_seed: int = 1234) -> Network:
    cpt, lookup = _dirichlet_cpts(cpt_seed)
    lookup.update(_ALARM_CARDS)
    net = _from_named(_ALARM_CARDS, _ALARM_ARCS, cpt)
    n_params = sum((int(net.cards[v]) - 1) * net.cpts[v].shape[0] for v in range(net.n))
    assert net.n == 37 and len(net.edges()) == 46 and n_params == 509
    return net


# ----------------------------------------------------------------------------
# Seeded random networks ("-like" generators)
# ----------------------------------------------------------------------------
def random_network(n: int, n_arcs: int, max_indegree: int, cards: Sequence[int],
                   seed: int, alpha: float = 1.0, prefix: str = "v") -> Network:
    """Random DAG over a seeded topological order with exactly ``n_arcs`` arcs
    (if feasible) and in-degree <= max_indegree; Dirichlet(alpha) CPT rows."""
    rng = np.random.default_rng(seed)
    width = len(str(n - 1))
    names = ["%s%0*d" % (prefix, width, i) for i in range(n)]   # already sorted
    order = rng.permutation(n)
    pos = np.empty(n, dtype=np.int64)
    pos[order] = np.arange(n)
    parents: List[List[int]] = [[] for _ in range(n)]
    arcs = set()
    attempts = 0
    while len(arcs) < n_arcs and attempts < 200 * n_arcs:
        attempts += 1
        a, b = rng.integers(0, n, size=2)
        if a == b:
            continue
        if pos[a] > pos[b]:
            a, b = b, a
        if (a, b) in arcs or len(parents[b]) >= max_indegree:
            continue
        arcs.add((int(a), int(b)))
        parents[int(b)].append(int(a))
    for ps in parents:
        ps.sort()
    card_arr = np.asarray(cards, dtype=np.int64)
    assert card_arr.shape == (n,)
    cpts = []
    for v in range(n):
        q = int(np.prod(card_arr[parents[v]])) if parents[v] else 1
        cpts.append(rng.dirichlet(np.full(int(card_arr[v]), alpha), size=q))
    return Network(names, card_arr, parents, cpts)


def hepar2_like(seed: int = 70) -> Network:
    """70 nodes, 123 arcs, max in-degree 6, cardinalities in {2,3,4}."""
    rng = np.random.default_rng(seed + 1)
    cards = rng.choice([2, 2, 2, 3, 3, 4], size=70)
    return random_network(70, 123, 6, cards, seed, prefix="h")


def munin1_like(seed: int = 186) -> Network:
    """186 nodes, 273 arcs, max in-degree 3, cardinalities 2..21 (a few at 21)."""
    rng = np.random.default_rng(seed + 1)
    cards = rng.choice([2, 3, 4, 5, 5, 6, 7, 8, 10, 12], size=186)
    cards[rng.choice(186, size=6, replace=False)] = 21
    return random_network(186, 273, 3, cards, seed, prefix="m")


def random_binary(n: int, seed: int = 24, arcs_per_node: float = 1.5, max_indegree: int = 3) -> Network:
    """n binary variables, ~1.5 arcs/node, max in-degree 3 (BASELINE.json configs[2])."""
    return random_network(n, int(round(arcs_per_node * n)), max_indegree, [2] * n, seed, prefix="b")


def to_reference_frame(codes: np.ndarray, names: Sequence[str], cards: Sequence[int]):
    """Duck-typed 'DataFrame' the reference's DataframeWrapper accepts
    (src/base/dataframe_wrapper.cpp:7-30): ``.columns`` and ``.values`` (object dtype,
    Fortran order, Python str cells)."""
    import types
    cols = []
    for v in range(codes.shape[1]):
        lut = np.array(state_names(int(cards[v])), dtype=object)
        cols.append(lut[codes[:, v]])
    obj = np.empty(codes.shape, dtype=object, order="F")
    for v, c in enumerate(cols):
        obj[:, v] = c
    return types.SimpleNamespace(columns=list(names), values=obj)
