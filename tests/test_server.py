"""The embedding server: examples/server.cpp's wire protocol on the B200 engine with
dynamic batching across connections.  CPU: the batching queue with a stub backend.  GPU: the real server binary."""
import os
import socket
import struct
import subprocess
import threading

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dynamic_batcher_host_logic(tmp_path):
    exe = str(tmp_path / "batcher_test")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-pthread", "-o", exe, os.path.join(ROOT, "tests", "cpp", "batcher_test.cpp")])
    out = subprocess.run([exe, "48", "25", "16", "200"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stdout + out.stderr
    f = dict(zip(out.stdout.split()[0::2], out.stdout.split()[1::2]))
    assert int(f["requests"]) == 48 * 25 and int(f["wrong"]) == 0 and int(f["over"]) == 0
    assert int(f["batches"]) < int(f["requests"]) // 2        # concurrent requests were merged into shared launches
    assert 1 < int(f["largest"]) <= 16
    # max_batch = 1 degenerates to the reference's one-text-per-call behaviour
    out = subprocess.run([exe, "8", "10", "1", "0"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0 and "batches 80 largest 1" in out.stdout, out.stdout


@pytest.mark.gpu
def test_server_protocol_and_batching(binding, make_model):
    """n_embd handshake, one text per message, n_embd floats back (examples/server.cpp:100-118, sample_client.py), from
    16 concurrent connections; answers equal bert_encode_batch on the same texts"""
    path = make_model("mini64", "f16", seed=51)
    words = ["[UNK]"] + [f"w{i}" for i in range(40)]
    rng = np.random.default_rng(5)
    texts = [" ".join(rng.choice(words, int(n))) for n in rng.integers(1, 60, 16 * 6)]
    m = binding.BertB200(path)
    want = m.encode_batch(texts)
    m.close()
    env = dict(os.environ, BERT_B200_DEVICES="0")
    srv = subprocess.Popen([os.path.join(ROOT, "build", "bert-server"), "-m", path, "--port", "0"], stdout=subprocess.PIPE, text=True, env=env)
    try:
        line = srv.stdout.readline()
        assert "Server running on port" in line, line
        port = int(line.split("port")[1].split()[0])
        got = [None] * len(texts)
        errs = []

        def client(k):
            try:
                s = socket.create_connection(("127.0.0.1", port), timeout=60)
                n_embd = struct.unpack("i", s.recv(4, socket.MSG_WAITALL))[0]
                assert n_embd == want.shape[1]
                for i in range(k, len(texts), 16):
                    s.sendall(texts[i].encode())
                    got[i] = np.frombuffer(s.recv(4 * n_embd, socket.MSG_WAITALL), np.float32).copy()
                s.close()
            except Exception as e:  # noqa: BLE001
                errs.append(repr(e))

        th = [threading.Thread(target=client, args=(k,)) for k in range(16)]
        [t.start() for t in th]
        [t.join(120) for t in th]
        assert not errs, errs
        for i, g in enumerate(got):
            assert g is not None and g.shape == (want.shape[1],)
            assert np.allclose(g, want[i], atol=2e-4), (i, float(np.abs(g - want[i]).max()))
    finally:
        srv.kill()
        srv.wait()
