"""Generates tests/golden/reward_cosine.npz by running the REFERENCE's own code.

Run in the build container only (it imports /root/reference, which does not exist on the GPU box):

    python tests/golden/make_reward_golden.py

The reference module cannot be imported as shipped: it logs in to the HF hub at
import and needs sentence_transformers / evaluate / wandb.  None of that touches the arithmetic we
pin, so those modules are stubbed; ``SentenceTransformer.encode`` is replaced by a lookup of seeded
synthetic embeddings.  ``RewardModel.calculate_factual_accuracy`` (reference :63-71),
``calculate_relevance`` (:73-79) and the ground-truth cosine inside ``calculate_reward`` (:99-104) then
run VERBATIM, with the real sklearn ``cosine_similarity`` (sklearn 1.9.0 here).  Their outputs are the
golden values: max_j cos(response, doc_j), cos(response, query), cos(response, ground_truth).
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference/reinforcement_learning_optimization_after_rag.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reward_cosine.npz")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _FakeST:
    table = {}

    def __init__(self, *a, **k):
        pass

    def encode(self, text):
        return _FakeST.table[text]


def load_reference():
    _stub("wandb", init=lambda *a, **k: None, log=lambda *a, **k: None, finish=lambda *a, **k: None)
    _stub("sentence_transformers", SentenceTransformer=_FakeST)
    _stub("evaluate", load=lambda *a, **k: None)
    import huggingface_hub                      # real package (transformers needs it); only login() is neutralised
    huggingface_hub.login = lambda *a, **k: None
    if "datasets" not in sys.modules:
        try:
            import datasets  # noqa: F401
        except Exception:
            _stub("datasets", load_dataset=None, Dataset=None)
    spec = importlib.util.spec_from_file_location("ref_rl", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = load_reference()
    rng = np.random.default_rng(20261106)
    cases = []
    # (n samples, m docs, d) — d values are the embedding sizes of BASELINE.json's configs
    for n, m, d in [(6, 2, 384), (5, 5, 768), (4, 4, 1024), (3, 3, 4096), (4, 0, 768), (3, 3, 100)]:
        resp = rng.standard_normal((n, d)).astype(np.float32)
        docs = rng.standard_normal((n, m, d)).astype(np.float32) * rng.uniform(0.1, 5.0, (n, m, 1)).astype(np.float32)
        if m:
            docs[0, 0] = resp[0] * 3.0          # cosine exactly ~1
            docs[1, m - 1] = -resp[1]           # cosine ~-1
            if n > 2:
                docs[2, 0] = 0.0                # zero-norm document: sklearn leaves it, cosine 0
        query = rng.standard_normal((n, d)).astype(np.float32)
        gt = rng.standard_normal((n, d)).astype(np.float32)
        cases.append((resp, docs, query, gt))

    rm = ref.RewardModel("stub")
    out = {}
    for ci, (resp, docs, query, gt) in enumerate(cases):
        n, m = docs.shape[0], docs.shape[1]
        fact = np.zeros(n, dtype=np.float32)
        rel = np.zeros(n, dtype=np.float32)
        gts = np.zeros(n, dtype=np.float32)
        for i in range(n):
            _FakeST.table = {"resp": resp[i], "query": query[i], "gt": gt[i]}
            names = []
            for j in range(m):
                _FakeST.table[f"doc{j}"] = docs[i, j]
                names.append(f"doc{j}")
            fact[i] = rm.calculate_factual_accuracy("resp", names)          # reference :63-71
            rel[i] = rm.calculate_relevance("resp", "query")                # reference :73-79
            _, parts = rm.calculate_reward("resp", "query", names, ground_truth="gt")        # :93-123
            gts[i] = parts["ground_truth_similarity"]
        out[f"c{ci}_resp"], out[f"c{ci}_docs"], out[f"c{ci}_query"], out[f"c{ci}_gt"] = resp, docs, query, gt
        out[f"c{ci}_fact"], out[f"c{ci}_rel"], out[f"c{ci}_gtsim"] = fact, rel, gts
    out["ncases"] = np.array(len(cases))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items() if k.endswith(("fact", "rel"))})


if __name__ == "__main__":
    main()
