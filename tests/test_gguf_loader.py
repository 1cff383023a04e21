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


def raw_gguf(tensors, kvs=b"", n_kv=0, version=3, n_tensors=None, data=b"", pad_to=32):
    """A GGUF file written by hand (ggml.c:20896-21160): tensors = [(name, [ne...], ggml type, offset)]."""
    import struct
    def s(b): return struct.pack("<Q", len(b)) + b
    out = b"GGUF" + struct.pack("<IQQ", version, len(tensors) if n_tensors is None else n_tensors, n_kv) + kvs
    for name, ne, ty, off in tensors:
        out += s(name.encode()) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", e & (2**64 - 1)) for e in ne) + struct.pack("<IQ", ty, off & (2**64 - 1))
    out += bytes(-len(out) % pad_to)
    return out + data


def test_loader_refuses_hostile_headers(tbmod, tmp_path):
    """Extents, offsets and counts come from the file: nothing may wrap, be allocated for, or be read past the mapping."""
    import struct
    p = tmp_path / "h.gguf"
    data = bytes(range(256)) * 16                      # 4096 bytes of tensor data

    def opens(blob):
        p.write_bytes(blob)
        return tbmod.GgufFile(p)

    def refused_at_open(blob):
        p.write_bytes(blob)
        with pytest.raises(tbmod.TB200Error):
            tbmod.GgufFile(p)

    # the well-formed twin first: one F32 vector of 1024 elements = the whole data section
    f = opens(raw_gguf([("v", [1024], H.F32, 0)], data=data))
    assert f.info("v")[3] == 4096 and np.array_equal(f.raw("v"), np.frombuffer(data, np.uint8))
    f.close()
    # offsets: just past the end, far past, and 2^64 - 32 (data_start + offset wraps to a small number)
    for off in (32, 1 << 40, -32, -4096 + 32):
        f = opens(raw_gguf([("v", [1024], H.F32, off)], data=data))
        with pytest.raises(tbmod.TB200Error, match="outside the file"):
            f.info("v")
        f.close()
    # extents whose byte size wraps: 2^62 floats x 4 = 2^64 = 0 bytes; rows x row bytes wrapping; 2^63 as a "negative" extent
    for ne in ([1 << 62], [1024, 1 << 52], [32, 1 << 29, 1 << 28], [1 << 61, 3]):
        f = opens(raw_gguf([("v", ne, H.F32, 0)], data=data))
        with pytest.raises(tbmod.TB200Error, match="outside the file"):
            f.info("v")
        f.close()
    refused_at_open(raw_gguf([("v", [1 << 63], H.F32, 0)], data=data))
    refused_at_open(raw_gguf([("v", [1024, 1 << 54], H.F32, 0)], data=data))
    refused_at_open(raw_gguf([("v", [32, 1 << 31, 1 << 31], H.F32, 0)], data=data))
    refused_at_open(raw_gguf([("v", [1 << 32, 1 << 32], H.F32, 0)], data=data))      # element count does not fit int64
    refused_at_open(raw_gguf([("v", [4, 4, 4, 4, 4], H.F32, 0)], data=data))          # rank 5
    refused_at_open(raw_gguf([("v", [], H.F32, 0)], data=data))                       # rank 0
    # counts: a header announcing 16M tensors / kvs in a 4 KB file, an array of 2^60 u64s, a string of 2^63 bytes
    refused_at_open(raw_gguf([], n_tensors=1 << 24, data=data))
    refused_at_open(raw_gguf([], n_tensors=1 << 60, data=data))
    refused_at_open(raw_gguf([], n_kv=1 << 24, data=data))
    key = struct.pack("<Q", 1) + b"k"
    refused_at_open(raw_gguf([], kvs=key + struct.pack("<IIQ", 9, 10, 1 << 60), n_kv=1, data=data))
    refused_at_open(raw_gguf([], kvs=key + struct.pack("<IQ", 8, 1 << 63), n_kv=1, data=data))
    refused_at_open(raw_gguf([], kvs=key + struct.pack("<I", 13), n_kv=1, data=data))             # unknown value type
    refused_at_open(raw_gguf([], kvs=key + struct.pack("<IIQ", 9, 9, 1), n_kv=1, data=data))      # array of arrays: not in the format
    refused_at_open(raw_gguf([("v", [1024], H.F32, 0)], version=1, data=data))
    refused_at_open(raw_gguf([("v", [1024], H.F32, 0)], version=4, data=data))
    # every metadata value type the format has, skipped correctly (the tensor after them is found where it is)
    kvs, n = b"", 0
    for ty, payload in ((0, b"\x01"), (1, b"\xff"), (2, b"\x01\x02"), (3, b"\x01\x02"), (4, bytes(4)), (5, bytes(4)), (6, bytes(4)), (7, b"\x01"),
                        (8, struct.pack("<Q", 3) + b"abc"), (10, bytes(8)), (11, bytes(8)), (12, bytes(8)),
                        (9, struct.pack("<IQ", 6, 3) + bytes(12)), (9, struct.pack("<IQ", 8, 2) + struct.pack("<Q", 1) + b"x" + struct.pack("<Q", 0)),
                        (9, struct.pack("<IQ", 12, 0))):
        k = f"key.{n}".encode()
        kvs += struct.pack("<Q", len(k)) + k + struct.pack("<I", ty) + payload
        n += 1
    k = b"general.alignment"
    kvs += struct.pack("<Q", len(k)) + k + struct.pack("<II", 4, 128)
    f = opens(raw_gguf([("v", [1024], H.F32, 0)], kvs=kvs, n_kv=n + 1, data=data, pad_to=128))
    assert np.array_equal(f.raw("v"), np.frombuffer(data, np.uint8))
    f.close()


def test_loader_survives_every_truncation(tbmod, tmp_path):
    """The first n bytes of a real file, for every n in the header and a sample beyond: open() either refuses or returns a file
    whose tensors are each readable in full or refused - never a crash, never a read past the mapping."""
    good = tmp_path / "m.gguf"
    tensors = model_tensors()
    write_model(good, tensors)
    blob = good.read_bytes()
    f = tbmod.GgufFile(good)
    f.close()
    p = tmp_path / "t.gguf"
    header_end = blob.index(b"blk.0.w0.weight") + 600
    opened = 0
    for n in list(range(0, min(header_end, len(blob)))) + list(range(header_end, len(blob), 997)):
        p.write_bytes(blob[:n])
        try:
            f = tbmod.GgufFile(p)
        except tbmod.TB200Error:
            continue
        opened += 1
        for name in f.names():
            try:
                raw = f.raw(name)
            except tbmod.TB200Error:
                continue
            assert np.array_equal(raw, tensors[name][2]), (n, name)
        f.close()
    assert opened > 0


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
