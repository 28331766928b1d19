"""CPU: the oracle (oracle/bert_oracle.c) is pinned bit-exactly to the reference's own outputs."""
import numpy as np
import pytest

from conftest import rand_tokens
from golden_util import golden_names, load_golden


# The C restatement is scalar (~10x slower than the reference's AVX2 build): on the BERT-base / BERT-large cases the CPU suite
# re-evaluates the shortest sentences up to this many GFLOP per case (at least two sentences) (the whole case is compared on the GPU box against the
# CUDA path); the small shapes are checked in full.
ORACLE_GFLOP_BUDGET = 25.0


def _gflop(hp, L):
    d, f, Lr = hp.n_embd, hp.n_intermediate, hp.n_layer
    return (Lr * (2 * L * (4 * d * d + 2 * d * f) + 4 * L * L * d) + 2 * L * d) / 1e9


@pytest.mark.parametrize("name", golden_names())
def test_oracle_matches_reference_golden_bit_exact(name, orc, make_model):
    c = load_golden(name)
    path = make_model(c["shape"], c["ftype"], c["seed"], c["n_vocab"], c["dist"])
    o = orc.Oracle(path)
    spent, done = 0.0, 0
    for i in np.argsort([len(t) for t in c["tokens"]], kind="stable"):
        t, e = c["tokens"][i], c["emb"][i]
        spent += _gflop(o.hp, len(t))
        if done >= 2 and spent > ORACLE_GFLOP_BUDGET:
            break
        got = o.eval(t)
        assert np.array_equal(got, e), (name, len(t), float(np.abs(got - e).max()))
        done += 1
    assert done >= 2


@pytest.mark.parametrize("ftype", ["f32", "f16", "q4_0", "q4_1"])
def test_oracle_matches_compiled_reference_live(ftype, orc, make_model):
    if not orc.available_reference():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    path = make_model("mini", ftype, seed=21)
    o, r = orc.Oracle(path), orc.Reference(path, n_threads=2)
    rng = np.random.default_rng(5)
    for L in [1, 4, 40, 97]:
        t = rand_tokens(rng, 500, L)
        assert np.array_equal(o.eval(t), r.eval(t))
    r.close()


def test_reference_output_is_thread_invariant(orc, make_model):
    if not orc.available_reference():
        pytest.skip("oracle/_ref not built")
    path = make_model("mini", "q4_0", seed=21)
    t = rand_tokens(np.random.default_rng(1), 500, 64)
    a = orc.Reference(path, n_threads=1)
    b = orc.Reference(path, n_threads=4)
    assert np.array_equal(a.eval(t), b.eval(t))


def test_oracle_discriminates_sentences(orc, make_model):
    """different sentences must not look identical, otherwise the cosine gate is meaningless"""
    path = make_model("mini", "f16", seed=21)
    o = orc.Oracle(path)
    rng = np.random.default_rng(2)
    a, b = o.eval(rand_tokens(rng, 500, 32)), o.eval(rand_tokens(rng, 500, 32))
    assert float(a @ b) < 0.999
