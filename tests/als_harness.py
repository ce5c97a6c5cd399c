"""Test-side restatement of the reference's ALS REST resources over a pluggable compute backend.

The resources themselves stay Java in production; this harness exists so the
reference's endpoint-level golden values (tests/golden/als_fixed_model.json) can be replayed against
(a) the CPU oracle and (b) the CUDA library through its C ABI, with identical orchestration.

Each function cites the reference resource it follows; paths are relative to
app/oryx-app-serving/src/main/java/com/cloudera/oryx/app/serving/als/ in the reference tree.
"""
from __future__ import annotations

import json
import math
import os

import numpy as np

import oracle

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name: str):
    with open(os.path.join(GOLDEN_DIR, name + ".json")) as f:
        return json.load(f)


class OracleBackend:
    """Compute backend made of the CPU oracle's functions."""

    name = "oracle"

    def __init__(self, features: int, y: np.ndarray, hash_vectors=None, max_bits=0):
        self.features = features
        self.y = np.ascontiguousarray(y, dtype=np.float32)
        self.hv = None if hash_vectors is None or len(hash_vectors) == 0 else np.ascontiguousarray(hash_vectors, np.float32)
        self.max_bits = max_bits
        self.part = None if self.hv is None else oracle.lsh_index_rows(self.hv, self.y)

    def topn(self, kind, q, how_many, excluded=None):
        cand = None
        if self.hv is not None:
            main = oracle.lsh_index(self.hv, q)
            cand = oracle.lsh_candidates(self.hv.shape[0], self.max_bits, main)
        rows, scores = oracle.topn(kind, q, self.y, how_many, part=self.part, cand=cand,
                                   excluded=excluded)
        return rows, scores

    def scores_all(self, kind, q):
        return oracle.scores_all(kind, q, self.y)

    def gram(self):
        return oracle.transpose_times_self(self.y)

    def solver(self, g):
        return oracle.Solver(g)

    def updated_xu(self, solver, value, xu, yi, implicit):
        return oracle.compute_updated_xu(solver, value, xu, yi, implicit)

    def dot(self, x, y):
        return oracle.dot(x, y)

    def norm(self, x):
        return oracle.norm(x)

    def close(self):
        pass


def test_rescorer():
    """TestALSRescorerProvider.TEST_RESCORER (reference test tree, TestALSRescorerProvider.java:25-35)."""

    def is_filtered(item_id: str) -> bool:
        return ord(item_id[-1]) % 2 == 0

    def rescore(item_id: str, score: float) -> float:
        return float("nan") if is_filtered(item_id) else score * 2.0

    return is_filtered, rescore


class ServingHarness:
    """String-ID bookkeeping the Java ALSServingModel keeps (user vectors, known items, id<->row)."""

    def __init__(self, golden_model: dict, backend_factory):
        m = golden_model
        self.features = m["features"]
        self.implicit = m["implicit"]
        self.x = {f"U{i}": np.asarray(v, dtype=np.float32) for i, v in enumerate(m["X"])}
        self.item_ids = [f"I{j}" for j in range(len(m["Y"]))]
        self.item_row = {iid: j for j, iid in enumerate(self.item_ids)}
        self.y = np.asarray(m["Y"], dtype=np.float32)
        self.known = {}
        for u, row in enumerate(m["known"]):
            self.known[f"U{u}"] = {f"I{j}" for j, c in enumerate(row) if c > 0}
        self.backend = backend_factory(self.features, self.y)

    # ---- ALSServingModel.topN with allowedPredicate / rescoreFn (model/ALSServingModel.java:289-335)
    def _topn(self, kind, q, how_many, excluded_ids=(), rescorer=None):
        excl_rows = sorted(self.item_row[i] for i in excluded_ids if i in self.item_row)
        if rescorer is None:
            rows, scores = self.backend.topn(kind, q, how_many, excluded=excl_rows or None)
            return [(self.item_ids[r], float(s)) for r, s in zip(rows, scores)]
        # Arbitrary host Rescorer: full score vector from the backend, TopNConsumer logic here
        # (oryx_scores_all keeps callback semantics on the host).
        is_filtered, rescore = rescorer
        scores = self.backend.scores_all(kind, q)
        vals = np.full(len(scores), np.nan)
        for r, s in enumerate(scores):
            iid = self.item_ids[r]
            if r in excl_rows or is_filtered(iid):
                continue
            vals[r] = rescore(iid, float(s))
        rows, out = oracle.topn_values(vals, how_many)
        return [(self.item_ids[r], float(s)) for r, s in zip(rows, out)]

    @staticmethod
    def _page(pairs, how_many, offset):
        # AbstractALSResource.toIDValueResponse (AbstractALSResource.java:37-43)
        return pairs[offset:offset + how_many]

    # ---- Recommend.java:74-114
    def recommend(self, user, how_many=10, offset=0, consider_known=False, rescorer=None):
        q = self.x[user]
        excl = () if consider_known else self.known.get(user, ())
        return self._page(self._topn(oracle.DOT, q, how_many + offset, excl, rescorer), how_many, offset)

    # ---- RecommendToMany.java:60-112 (mean vector, DotsFunction.java:34-45)
    def recommend_to_many(self, users, how_many=10, offset=0, consider_known=False, rescorer=None):
        q = oracle.dots_mean(np.stack([self.x[u] for u in users]))
        excl = set()
        if not consider_known:
            for u in users:
                excl |= self.known.get(u, set())
        return self._page(self._topn(oracle.DOT, q, how_many + offset, excl, rescorer), how_many, offset)

    # ---- Similarity.java:62-104 (CosineAverageFunction.java:39-56; multi-vector ctor even for one item)
    def similarity(self, items, how_many=10, offset=0, rescorer=None):
        q = oracle.cosine_average(np.stack([self.y[self.item_row[i]] for i in items]), single_ctor=False)
        return self._page(self._topn(oracle.COSINE, q, how_many + offset, set(items), rescorer), how_many, offset)

    # ---- Estimate.java:54-71
    def estimate(self, user, items):
        xu = self.x[user]
        return [0.0 if i not in self.item_row else self.backend.dot(self.y[self.item_row[i]], xu)
                for i in items]

    # ---- SimilarityToItem.java:48-68
    def similarity_to_item(self, to_item, items):
        t = self.y[self.item_row[to_item]]
        tn = self.backend.norm(t)
        out = []
        for i in items:
            if i not in self.item_row:
                out.append(0.0)
            else:
                v = self.y[self.item_row[i]]
                out.append(self.backend.dot(v, t) / (tn * self.backend.norm(v)))
        return out

    # ---- EstimateForAnonymous.buildTemporaryUserVector (EstimateForAnonymous.java:73-87)
    def build_temporary_user_vector(self, context, xu):
        solver = self.backend.solver(self.backend.gram())  # ALSServingModel.getYTYSolver :357-373
        for item, value in context:
            yi = self.y[self.item_row[item]] if item in self.item_row else None
            new_xu = self.backend.updated_xu(solver, value, xu, yi, self.implicit)
            if new_xu is not None:
                xu = new_xu
        return xu

    # ---- EstimateForAnonymous.java:48-62
    def estimate_for_anonymous(self, to_item, context):
        t = self.y[self.item_row[to_item]]
        xu = self.build_temporary_user_vector(context, None)
        return 0.0 if xu is None else self.backend.dot(xu, t)

    # ---- RecommendToAnonymous.java:64-100
    def recommend_to_anonymous(self, context, how_many=10, offset=0, rescorer=None):
        xu = self.build_temporary_user_vector(context, None)
        assert xu is not None
        excl = {i for i, _ in context}
        return self._page(self._topn(oracle.DOT, xu, how_many + offset, excl, rescorer), how_many, offset)

    # ---- RecommendWithContext.java:62-106
    def recommend_with_context(self, user, context, how_many=10, offset=0, consider_known=False, rescorer=None):
        xu = self.build_temporary_user_vector(context, self.x[user])
        excl = {i for i, _ in context}
        if not consider_known:
            excl |= self.known.get(user, set())
        return self._page(self._topn(oracle.DOT, xu, how_many + offset, excl, rescorer), how_many, offset)

    # ---- Because.java:56-86
    def because(self, user, item, how_many=10, offset=0):
        iv = self.y[self.item_row[item]]
        ivn = self.backend.norm(iv)
        pairs = []
        for k in sorted(self.known.get(user, ())):
            ov = self.y[self.item_row[k]]
            pairs.append((k, self.backend.dot(iv, ov) / (ivn * self.backend.norm(ov))))
        pairs.sort(key=lambda p: -p[1])
        return self._page(pairs, how_many, offset)

    # ---- MostSurprising.java:58-81
    def most_surprising(self, user, how_many=10, offset=0):
        xu = self.x[user]
        pairs = [(k, self.backend.dot(xu, self.y[self.item_row[k]])) for k in sorted(self.known.get(user, ()))]
        pairs.sort(key=lambda p: p[1])
        return self._page(pairs, how_many, offset)

    # ---- PopularRepresentativeItems.java:46-60
    def popular_representative_items(self):
        out = []
        unit = np.zeros(self.features, dtype=np.float32)
        for i in range(self.features):
            unit[i] = 1.0
            top = self._topn(oracle.DOT, unit, 1)
            out.append(top[0][0] if top else None)
            unit[i] = 0.0
        return out


def check_fixed_model_goldens(h: ServingHarness, score_tol_floor: float = 0.0):
    """Replays every endpoint golden in als_fixed_model.json against harness ``h``."""
    g = load_golden("als_fixed_model")
    rescorer = test_rescorer()

    def tol(c):
        return max(c.get("tol", 1e-6), score_tol_floor)

    def check_top(res, c):
        if "count" in c:
            assert len(res) == c["count"], (c["_cite"], res)
        vals = [v for _, v in res]
        assert all(vals[i] >= vals[i + 1] for i in range(len(vals) - 1)), (c["_cite"], res)
        if "top_id" in c:
            assert res[0][0] == c["top_id"], (c["_cite"], res)
            assert abs(res[0][1] - c["top_value"]) <= tol(c), (c["_cite"], res)

    for c in g["recommend"]:
        if "how_many_cases" in c:
            for hm, exp in c["how_many_cases"]:
                assert len(h.recommend(c["user"], how_many=hm)) == exp, c["_cite"]
            continue
        if "offset_cases" in c:
            for hm, off, exp in c["offset_cases"]:
                assert len(h.recommend(c["user"], how_many=hm, offset=off)) == exp, c["_cite"]
            continue
        res = h.recommend(c["user"], consider_known=c.get("consider_known", False),
                          rescorer=rescorer if c.get("rescorer") else None)
        check_top(res, c)
    for c in g["recommend_to_many"]:
        res = h.recommend_to_many(c["users"], consider_known=c.get("consider_known", False),
                                  rescorer=rescorer if c.get("rescorer") else None)
        check_top(res, c)
    for c in g["similarity"]:
        res = h.similarity(c["items"])
        check_top(res, c)
        if "id_at" in c:
            assert res[c["id_at"][0]][0] == c["id_at"][1], (c["_cite"], res)
        if "value_at" in c:
            assert abs(res[c["value_at"][0]][1] - c["value_at"][1]) <= tol(c), (c["_cite"], res)
    for c in g["estimate"]:
        res = h.estimate(c["user"], c["items"])
        assert len(res) == len(c["values"])
        for a, b in zip(res, c["values"]):
            assert abs(a - b) <= tol(c), (c["_cite"], res)
    for c in g["similarity_to_item"]:
        res = h.similarity_to_item(c["to_item"], c["items"])
        for a, b in zip(res, c["values"]):
            assert abs(a - b) <= tol(c), (c["_cite"], res)
    for c in g["estimate_for_anonymous"]:
        v = h.estimate_for_anonymous(c["to_item"], [tuple(p) for p in c["context"]])
        assert abs(v - c["value"]) <= tol(c), (c["_cite"], v)
    for c in g["recommend_to_anonymous"]:
        res = h.recommend_to_anonymous([tuple(p) for p in c["context"]],
                                       rescorer=rescorer if c.get("rescorer") else None)
        check_top(res, c)
    for c in g["recommend_with_context"]:
        res = h.recommend_with_context(c["user"], [tuple(p) for p in c["context"]])
        check_top(res, c)
    for c in g["because"]:
        res = h.because(c["user"], c["item"])
        assert res[0][0] == c["top_id"] and abs(res[0][1] - c["top_value"]) <= tol(c), (c["_cite"], res)
    for c in g["most_surprising"]:
        res = h.most_surprising(c["user"])
        assert res[0][0] == c["top_id"] and abs(res[0][1] - c["top_value"]) <= tol(c), (c["_cite"], res)
    for c in g["popular_representative_items"]:
        res = h.popular_representative_items()
        assert len(res) == c["count"] and res[0] in c["first_in"] and res[1] == c["second"], (c["_cite"], res)


def random_model(n_items: int, features: int, seed: int) -> np.ndarray:
    """LoadTestALSModelFactory's recipe (reference test tree, model/LoadTestALSModelFactory.java:38-51):
    every component i.i.d. N(0,1), fp32."""
    rng = np.random.default_rng(seed)
    return rng.standard_normal((n_items, features), dtype=np.float32)


def random_hash_vectors(num_hashes: int, features: int, seed: int) -> np.ndarray:
    """Stand-in for LocalitySensitiveHash's hash-vector choice (LocalitySensitiveHash.java:80-104):
    greedy most-orthogonal of >=1000 Gaussian candidates.  The reference draws them from Well19937c,
    which is not reproducible here, so hash vectors are an INPUT to both oracle and library."""
    rng = np.random.default_rng(seed)
    chosen = []
    for i in range(num_hashes):
        best, best_score, since = None, math.inf, 0
        while since < 1000:
            cand = rng.standard_normal(features).astype(np.float32)
            cn = oracle.norm(cand)
            score = sum(abs(oracle.dot(c, cand)) / oracle.norm(c) / cn for c in chosen)
            if score < best_score:
                best, best_score, since = cand, score, 0
                if score == 0.0:
                    break
            else:
                since += 1
        chosen.append(best)
    return np.stack(chosen) if chosen else np.zeros((0, features), dtype=np.float32)
