"""CPU tests of the host side: C-ABI exports, error behaviour without a GPU,
DataframeWrapper semantics, the PC / exhaustive-search drivers over the test
seam (checker-backed evaluator), OracleGraph PC (the reference's own test_pc)."""
import ctypes
import re
import os
import types

import numpy as np
import pytest

import oracle
import openbnsl_b200
from helpers import OracleEvaluator, load_golden, pdag_to_adj, ranked
from openbnsl_b200 import capi, networks as nw

L = openbnsl_b200.openbnsllib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_c_abi_exports_every_declared_symbol():
    with open(os.path.join(ROOT, "include", "openbnsl_b200.h")) as f:
        declared = set(re.findall(r"OBN_API\s+[\w\s\*]+?\b(obn_\w+)\s*\(", f.read()))
    assert declared and declared == set(capi.EXPORTS)
    lib = ctypes.CDLL(capi.LIB_PATH)
    for s in declared:
        assert hasattr(lib, s), s
    assert b"sm_100a" in capi.load().obn_version()


def test_header_is_plain_c(tmp_path):
    """The drop-in boundary is a C ABI: the header must compile as C99 (no C++ types, no torch types)."""
    import shutil
    import subprocess
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    src = tmp_path / "hdr.c"
    src.write_text('#include "openbnsl_b200.h"\nint main(void) { obn_profile p; (void)p; return OBN_OK; }\n')
    r = subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-I",
                        os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_two_half_ci_call_rejects_a_missing_ctx():
    # obn_ci_tests_submit / _collect validate before touching a device, like every other entry point
    lib = capi.load()
    assert lib.obn_ci_tests_collect(None, None, None, None) == capi.OBN_ERR_INVALID
    assert b"ctx" in lib.obn_last_error()
    assert lib.obn_ci_tests_submit(None, None, None, None, 0, None, None, 1, 1) == capi.OBN_ERR_INVALID


def test_module_surface_matches_reference():
    assert L.cuda_enabled is True
    for sub, names in {"base": ["DataframeWrapper", "ContingencyTable", "PDAG"],
                       "score": ["BDeu", "K2", "BIC", "AIC", "calculate_local_score"],
                       "citest": ["ChiSquare", "GSquare", "OracleGraph", "citest"],
                       "structure_learning": ["exhaustive_search", "pc", "simulated_annealing", "rai"]}.items():
        for n in names:
            assert hasattr(getattr(L, sub), n), (sub, n)
    assert repr(L.score.BDeu()) == "BDeu(1.000000)" and repr(L.citest.GSquare()) == "GSquare(level=0.050000)"
    g = L.base.PDAG(3)
    g.add_edge(0, 1)
    assert g.has_edge(0, 1) and not g.has_edge(1, 0) and g.num_vars == 3
    g.remove_edge(0, 1)
    assert not g.has_edge(0, 1)


def test_dataframe_wrapper_semantics():
    # codes are ranks of lexicographically sorted strings: "10" < "2" (dataframe_wrapper.cpp:38-51)
    vals = np.array([["b", "10"], ["a", "2"], ["b", "2"], ["c", "10"]], dtype=object)
    for order in ("F", "C"):
        frame = types.SimpleNamespace(columns=["first", "second"], values=np.array(vals, order=order))
        df = L.base.DataframeWrapper(frame)
        assert df.num_of_vars == 2 and df.num_of_datapoints == 4
        assert df.col_idx2str == ["first", "second"] and df.col_str2idx == {"first": 0, "second": 1}
        assert df.val_idx2str == [["a", "b", "c"], ["10", "2"]]
        assert df.num_of_values == [3, 2]
        assert df.data_column_major == [[1, 0, 1, 2], [0, 1, 1, 0]]
        assert df.data_row_major == [[1, 0], [0, 1], [1, 1], [2, 0]]
    ints = types.SimpleNamespace(columns=["v"], values=np.array([[10], [2], [2], [10]], dtype=np.int64))
    assert L.base.DataframeWrapper(ints).data_column_major == [[0, 1, 1, 0]]
    with pytest.raises(ValueError):
        L.base.DataframeWrapper(types.SimpleNamespace(columns=["a"], values=np.array([[1.5]])))
    with pytest.raises(ValueError, match="pandas"):
        L.base.DataframeWrapper(object())
    import pandas as pd
    pdf = pd.DataFrame({"b": ["x", "y", "x"], "a": ["1", "1", "0"]})
    df = L.base.DataframeWrapper(pdf[sorted(pdf.columns)].astype(object))
    assert df.col_idx2str == ["a", "b"] and df.data_column_major == [[1, 1, 0], [0, 1, 0]]


def test_dataframe_wrapper_ingest_paths_agree_with_a_plain_ranking():
    # the ingest is written around a pointer / value cache (dataframe_wrapper.cpp); whatever the cells are --
    # shared str objects, one object per cell, non-ASCII text, bytes, integers -- the codes must be the ranks of
    # the sorted distinct strings (src/base/dataframe_wrapper.cpp:36-70)
    rng = np.random.default_rng(0)
    states = ["b", "a", "10", "9", "é", "aa", "", "Z", "z0"]
    idx = rng.integers(0, len(states), size=(500, 3))
    shared = np.array(states, dtype=object)[idx]                                   # few objects, many references
    fresh = np.empty(idx.shape, dtype=object)
    for r in range(idx.shape[0]):
        for c in range(idx.shape[1]):
            fresh[r, c] = "".join(list(states[idx[r, c]]))                        # a new object per cell
    want = []
    for c in range(3):
        order = sorted(set(states[i] for i in idx[:, c]), key=lambda t: t.encode("utf-8"))
        want.append([order.index(states[i]) for i in idx[:, c]])
    for vals in (shared, fresh, np.asfortranarray(fresh)):
        df = L.base.DataframeWrapper(types.SimpleNamespace(columns=["x", "y", "z"], values=vals))
        assert df.data_column_major == want
        assert df.val_idx2str[0] == sorted(set(states[i] for i in idx[:, 0]), key=lambda t: t.encode("utf-8"))
    ref = oracle.reference_modules()[0]
    if ref is not None:                                                            # the unmodified reference agrees
        r = ref.base.DataframeWrapper(types.SimpleNamespace(columns=["x", "y", "z"], values=np.asfortranarray(fresh)))
        assert r.data_column_major == want and r.val_idx2str == df.val_idx2str
    # bytes cells go through the same string cast as in the reference
    b = np.array([[b"q"], [b"p"], [b"q"]], dtype=object)
    assert L.base.DataframeWrapper(types.SimpleNamespace(columns=["v"], values=b)).data_column_major == [[1, 0, 1]]
    # integers rank by their decimal strings, negative ones included; more than 1024 rows cycle the value cache
    ints = rng.integers(-3, 12, size=(3000, 2))
    di = L.base.DataframeWrapper(types.SimpleNamespace(columns=["i", "j"], values=ints))
    for c in range(2):
        order = sorted(set(str(v) for v in ints[:, c]))
        assert di.data_column_major[c] == [order.index(str(v)) for v in ints[:, c]]
    # uint8 codes: the 257th distinct value is an error, not a wrap-around
    many = np.array([["v%03d" % i] for i in range(257)], dtype=object)
    with pytest.raises(ValueError, match="more than 256"):
        L.base.DataframeWrapper(types.SimpleNamespace(columns=["v"], values=many))
    with pytest.raises(ValueError, match="cast value to string"):
        L.base.DataframeWrapper(types.SimpleNamespace(columns=["v"], values=np.array([[1.5], [object()]], dtype=object)))


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        capi.Context(0)
    df = L.base.DataframeWrapper.from_codes(np.zeros((4, 2), dtype=np.uint8), ["a", "b"], [1, 1])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        L.structure_learning.pc(df, L.citest.GSquare(0.05), 1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        L.base.ContingencyTable([0], df)


def test_argument_errors_match_reference():
    df = L.base.DataframeWrapper.from_codes(np.zeros((4, 3), dtype=np.uint8), ["a", "b", "c"], [1, 1, 1])
    with pytest.raises(ValueError, match="max_parents"):          # exhaustive_search.cpp:18-19
        L.structure_learning.exhaustive_search(df, L.score.BDeu(1.0), 3)
    with pytest.raises(ValueError, match="Unsupported score type"):   # local_score.h:43
        L.structure_learning.exhaustive_search(df, L.score.BIC(), 1)
    with pytest.raises(ValueError, match="sorted"):               # contingency_table.h:37-38
        L.base.ContingencyTable([1, 0], df)


@pytest.mark.parametrize("name", ["asia_10k_s0", "alarm_3k_s1"])
def test_pc_driver_matches_golden(name):
    g, net, codes, cards = load_golden(name)
    ods = oracle.Dataset(codes, cards)
    for e in g["pc"]:
        T = L.citest.ChiSquare(e["level"]) if e["kind"] == 0 else L.citest.GSquare(e["level"])
        for tb in (0, 7):        # default batching and tiny rounds must agree
            got = L._testing.pc(net.n, OracleEvaluator(ods), T, e["max_cond_vars"], e["stable"], None, tb)
            assert pdag_to_adj(got, net.n).tolist() == e["adj"], (e["kind"], e["stable"], tb)


def test_es_driver_matches_golden():
    g, net, codes, cards = load_golden("asia_10k_s0")
    ods = oracle.Dataset(codes, cards)
    for e in g["exhaustive_search"]:
        got = L._testing.exhaustive_search(net.n, OracleEvaluator(ods), L.score.BDeu(e["ess"]), e["max_parents"])
        assert pdag_to_adj(got, net.n).tolist() == e["adj"]


def _skeleton_cpdag(net):
    """CPDAG of the true DAG at the level pc() returns (skeleton + v-structures, no Meek)."""
    n = net.n
    und = np.zeros((n, n), dtype=np.uint8)
    for p, v in net.edges():
        und[p, v] = und[v, p] = 1
    adj = und.copy()
    for v in range(n):
        ps = net.parents[v]
        for i in range(len(ps)):
            for j in range(i + 1, len(ps)):
                a, b = ps[i], ps[j]
                if not und[a, b]:
                    adj[v, a] = 0
                    adj[v, b] = 0
    return adj


@pytest.mark.parametrize("netf", [nw.asia, nw.alarm])
@pytest.mark.parametrize("stable", [True, False])
def test_pc_with_oracle_graph_recovers_structure(netf, stable):
    # the reference's own PC test: d-separation oracle, max_cond_vars = n (pytest/structure_learning/test_pc.py:11-33)
    net = netf()
    truth = L.base.PDAG(net.n)
    for p, v in net.edges():
        truth.add_edge(p, v)
    df = L.base.DataframeWrapper.from_codes(np.zeros((2, net.n), dtype=np.uint8), net.names, [1] * net.n)
    mc = net.n if net.n <= 8 else 4
    got = pdag_to_adj(L.structure_learning.pc(df, L.citest.OracleGraph(truth), mc, stable), net.n)
    assert (got == _skeleton_cpdag(net)).all()


def test_depth0_oracle_graph_facts():
    net = nw.asia()                     # names sorted: asia bronc dysp either lung smoke tub xray
    truth = L.base.PDAG(net.n)
    for p, v in net.edges():
        truth.add_edge(p, v)
    i = {s: k for k, s in enumerate(net.names)}
    df = L.base.DataframeWrapper.from_codes(np.zeros((2, net.n), dtype=np.uint8), net.names, [1] * net.n)
    g0 = pdag_to_adj(L.structure_learning.pc(df, L.citest.OracleGraph(truth), 0, True), net.n)
    assert g0[i["asia"], i["smoke"]] == 0          # marginally independent roots
    assert g0[i["tub"], i["either"]] == 1 and g0[i["tub"], i["lung"]] == 0
