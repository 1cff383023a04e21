"""GGUF tensor loader: tb200_gguf_* must see exactly the tensors, shapes, types and bytes that
gguf-py wrote (the package llama.cpp ships, llama.cpp/gguf-py) — CPU; and a matrix loaded from the file must multiply
like the same bytes handed over directly — GPU."""
import numpy as np
import pytest

import helpers as H
from oracle import binding as B

gguf = pytest.importorskip("gguf")

GGML_TO_GGUF = {H.F16: gguf.GGMLQuantizationType.F16, H.Q4_0: gguf.GGMLQuantizationType.Q4_0, H.Q8_0: gguf.GGMLQuantizationType.Q8_0,
                H.Q4_K: gguf.GGMLQuantizationType.Q4_K, H.F32: gguf.GGMLQuantizationType.F32}


def write_model(path, tensors, alignment=None):
    """tensors: name -> (ggml type, ne (innermost first), raw uint8 bytes)"""
    w = gguf.GGUFWriter(str(path), "llama")
    w.add_string("general.name", "tb200-test")
    w.add_uint32("llama.block_count", 1)
    w.add_array("tokenizer.ggml.tokens", ["a", "bc", "def"])
    w.add_array("llama.rope.scales", [1.0, 2.0, 3.5])
    if alignment:
        w.add_custom_alignment(alignment)
    for name, (ty, ne, raw) in tensors.items():
        if ty == H.F32:
            w.add_tensor(name, raw.view(np.float32).reshape(list(reversed(ne))))
        elif ty == H.F16:
            w.add_tensor(name, raw.view(np.float16).reshape(list(reversed(ne))))
        else:
            row = B.row_bytes(ty, ne[0])
            w.add_tensor(name, raw.reshape(list(reversed(ne[1:])) + [row]), raw_dtype=GGML_TO_GGUF[ty])
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()


def model_tensors():
    K = 512
    t = {}
    for i, (wt, rows) in enumerate(((H.Q4_0, 96), (H.Q8_0, 40), (H.Q4_K, 64), (H.F16, 33))):
        t[f"blk.0.w{i}.weight"] = (wt, [K, rows], H.make_weights(wt, rows, K, seed=20 + i, realistic_q4k=False))
    t["blk.0.ffn_gate_exps.weight"] = (H.Q4_0, [K, 24, 4], H.make_weights(H.Q4_0, 24 * 4, K, seed=30))
    t["output_norm.weight"] = (H.F32, [K], np.linspace(0.5, 1.5, K, dtype=np.float32).view(np.uint8))
    return t


@pytest.fixture(scope="module")
def tbmod():
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    return tb


@pytest.mark.parametrize("alignment", [None, 64])
def test_loader_matches_gguf_py(tbmod, tmp_path, alignment):
    tensors = model_tensors()
    path = tmp_path / "m.gguf"
    write_model(path, tensors, alignment)
    f = tbmod.GgufFile(path)
    assert sorted(f.names()) == sorted(tensors)
    reader = gguf.GGUFReader(str(path))
    by_name = {t.name: t for t in reader.tensors}
    for name, (ty, ne, raw) in tensors.items():
        got_ty, got_ne, _, nbytes = f.info(name)
        assert got_ty == ty and got_ne[:len(ne)] == ne and all(v == 1 for v in got_ne[len(ne):])
        assert nbytes == raw.size
        assert np.array_equal(f.raw(name), raw), name                                    # what we wrote
        assert np.array_equal(f.raw(name), np.asarray(by_name[name].data).view(np.uint8).reshape(-1)), name   # what gguf-py reads back
    f.close()


def test_loader_rejects_garbage(tbmod, tmp_path):
    p = tmp_path / "bad.gguf"
    p.write_bytes(b"GGUF" + bytes(40))
    with pytest.raises(tbmod.TB200Error):
        tbmod.GgufFile(p)
    p.write_bytes(b"NOPE" + bytes(64))
    with pytest.raises(tbmod.TB200Error):
        tbmod.GgufFile(p)
    with pytest.raises(tbmod.TB200Error):
        tbmod.GgufFile(tmp_path / "missing.gguf")
    good = tmp_path / "m.gguf"
    write_model(good, model_tensors())
    data = good.read_bytes()
    p.write_bytes(data[:len(data) - 4096])          # data section cut short
    f = tbmod.GgufFile(p)                          # the header still parses; tensors past the cut must be refused, not read
    refused = 0
    for name in f.names():
        try:
            f.info(name)
        except tbmod.TB200Error:
            refused += 1
    assert 0 < refused < len(f.names())
    f.close()


@pytest.mark.gpu
def test_gguf_weights_multiply_like_direct_upload(tmp_path):
    import torch
    import llamafile_b200 as tb
    if not tb.link():
        pytest.fail("no sm_100 device")
    o = B.oracle()
    tensors = model_tensors()
    path = tmp_path / "m.gguf"
    write_model(path, tensors)
    f = tb.GgufFile(path)
    K = 512
    for name, (ty, ne, raw) in tensors.items():
        if ty == H.F32 or len(ne) != 2:
            continue
        rows = ne[1]
        for T in (1, 40):
            X = H.make_acts(T, K, seed=T)
            w = f.weights(name)
            out = torch.full((T, rows), float("nan"), device="cuda")
            tb.mul_mat(w, torch.from_numpy(X).cuda(), T, out)
            tb.synchronize()
            H.assert_close(out.cpu().numpy(), o.mul_mat(raw, ty, X), f"{name} T={T}")
            w.free()
        # a rank's row shard
        lo, hi = rows // 4, rows // 4 + rows // 2
        w = f.weights(name, lo, hi)
        X = H.make_acts(2, K, seed=5)
        out = torch.full((2, hi - lo), float("nan"), device="cuda")
        tb.mul_mat(w, torch.from_numpy(X).cuda(), 2, out)
        tb.synchronize()
        rb = B.row_bytes(ty, K)
        H.assert_close(out.cpu().numpy(), o.mul_mat(raw[lo * rb:hi * rb], ty, X), f"{name} rows {lo}:{hi}")
        w.free()
    # stacked experts -> mixmul
    ty, ne, raw = tensors["blk.0.ffn_gate_exps.weight"]
    w = f.weights("blk.0.ffn_gate_exps.weight")
    assert w.rows == ne[1] * ne[2]
    tokens, experts, rows = 5, ne[2], ne[1]
    X = H.make_acts(tokens, K, seed=9).reshape(tokens, 1, K)
    plan = np.array([[t % experts, (t + 1) % experts] for t in range(tokens)], np.int32)
    out = torch.full((tokens, 2, rows), float("nan"), device="cuda")
    assert tb.mixmul(w, experts, torch.from_numpy(X).cuda(), 1, torch.from_numpy(plan).cuda(), 2, tokens, out)
    tb.synchronize()
    rb = B.row_bytes(ty, K)
    got = out.cpu().numpy()
    for t in range(tokens):
        for j in range(2):
            e = plan[t, j]
            H.assert_close(got[t, j][None], o.mul_mat(raw[e * rows * rb:(e + 1) * rows * rb], ty, X[t]), f"expert {e} token {t}")
    w.free()
    f.close()
