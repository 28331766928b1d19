"""Driver of tools/numerics_study.c: 1-cos of the final embedding against the reference's committed outputs (tests/golden)
when single GPU-style rounding points are switched on (see the C file's header).  CPU only.

    python tools/numerics_study.py base_q4_0 [max_len] """
import ctypes
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle as O  # noqa: E402
from golden_util import load_golden  # noqa: E402

MASKS = [(0, "exact operands, other summation order"), (1, "a~ f16"), (2, "w~ f16"), (3, "a~ and w~ f16"), (4, "QKV f16"),
         (8, "residual f16"), (16, "pre-LN f16"), (4 + 8, "QKV + residual f16"), (1 + 2 + 4 + 8, "a~ w~ QKV residual f16 (fused-LN path)"),
         (1 + 2 + 4 + 8 + 16, "all (unfused path)")]


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "base_q4_0"
    max_len = int(sys.argv[2]) if len(sys.argv) > 2 else 33
    so = os.path.join(tempfile.gettempdir(), "libstudy.so")
    subprocess.check_call(["gcc", "-O2", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tools", "numerics_study.c"), "-lm", "-lpthread"])
    c = load_golden(name)
    g = O.load_gguf_io()
    path = os.path.join(tempfile.gettempdir(), f"study_{name}.gguf")
    if not os.path.exists(path):
        g.make_model(path, c["shape"], c["ftype"], seed=c["seed"], n_vocab=c["n_vocab"], dist=c["dist"])
    o = O.Oracle(path)
    lib = ctypes.CDLL(so)
    lib.study_forward.argtypes = [ctypes.c_void_p, O._i32p, ctypes.c_int, O._f32p, O._f32p, ctypes.c_int]
    lib.oracle_init()
    sel = [(t, e) for t, e in zip(c["tokens"], c["emb"]) if len(t) <= max_len]
    print(name, "sentences L =", [len(t) for t, _ in sel])
    for mask, what in MASKS:
        row = []
        for t, e in sel:
            tok = np.ascontiguousarray(t, np.int32)
            out = np.zeros(o.hp.n_embd, np.float32)
            lib.study_forward(ctypes.byref(o.model), tok.ctypes.data_as(O._i32p), len(tok), O._fp(out), None, mask)
            row.append(1.0 - float(out.astype(np.float64) @ e.astype(np.float64) / (np.linalg.norm(out.astype(np.float64)) * np.linalg.norm(e.astype(np.float64)))))
        print(f"mask {mask:2d} {what:45s}", " ".join(f"{v:.1e}" for v in row), flush=True)


if __name__ == "__main__":
    main()
