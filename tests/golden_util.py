"""Loading the committed golden cases (tests/golden/*.npz, made by tests/golden/gen_golden.py)."""
import glob
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def golden_names():
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(HERE, "golden", "*.npz")))


def load_golden(name):
    z = np.load(os.path.join(HERE, "golden", name + ".npz"))
    lens = z["lens"]
    toks, off = [], 0
    for L in lens:
        toks.append(z["tokens"][off:off + L].astype(np.int32))
        off += L
    emb = z["emb"]
    # emb_portable: the same sentences through the reference's portable (non-AVX2) build; spread = 1 - cos between the two
    # builds of the reference itself = the reproducibility floor of that model and sentence (tests/golden/gen_golden.py)
    port = z["emb_portable"] if "emb_portable" in z.files else None
    spread = None
    if port is not None:
        a, b = emb.astype(np.float64), port.astype(np.float64)
        spread = 1.0 - (a * b).sum(1) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))
    return dict(shape=str(z["shape"]), ftype=str(z["ftype"]), seed=int(z["seed"]), n_vocab=int(z["n_vocab"]),
                dist=str(z["dist"]), tokens=toks, emb=emb, emb_portable=port, spread=spread)
