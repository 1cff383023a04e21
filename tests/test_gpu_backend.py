"""GPU (B200): the module as a ggml backend. The reference's own ggml (oracle/_ref, pristine ggml.c /
ggml-alloc.c / ggml-backend.c) is the HOST: it allocates the graph's tensors in the TB200 buffer type, uploads them with
ggml_backend_tensor_set and calls ggml_backend_graph_compute, i.e. the backend object is exercised through its interface
table by the code that would exercise it in llamafile. Every case runs the SAME graph on ggml's CPU backend and compares
the way the reference tests its backends (llama.cpp/tests/test-backend-ops.cpp:966-968: NMSE <= 5e-4; here also the
scaled-error gate of the rest of the suite)."""
import numpy as np
import pytest

import helpers as H
from ggml_host import Graph, ggml
from oracle import binding as B

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def host(tb):
    G = ggml()
    if G is None:
        pytest.skip("oracle/_ref (the reference's ggml) is not built")
    assert tb.backend_link(G.ref_backend_api()), "tb200_backend_link found no device"
    backend = tb.backend_init(0)
    cpu = G.ggml_backend_cpu_init()
    G.ggml_backend_cpu_set_n_threads(cpu, 4)
    yield G, backend, cpu
    G.ggml_backend_free(backend)
    G.ggml_backend_free(cpu)


def run_both(host, build, inputs):
    """build(graph) -> (dict name -> tensor, list of outputs [(tensor, shape)]); inputs: name -> array. Returns [(tb200, cpu)] per output."""
    G, backend, cpu = host
    res = []
    for bk in (backend, cpu):
        g = Graph(bk)
        tensors, outs = build(g)
        g.build([t for t, _ in outs])
        for name, arr in inputs.items():
            g.set(tensors[name], arr)
        assert g.compute() == 0
        res.append([g.get(t, shape) for t, shape in outs])
        g.free()
    return list(zip(*res))


@pytest.mark.parametrize("wt", H.WEIGHT_TYPES)
@pytest.mark.parametrize("N,K,T", [(16, 256, 1), (16, 256, 16), (130, 1024, 1), (130, 1024, 9), (384, 512, 33), (64, 2048, 300)])
def test_mul_mat_vs_cpu_backend(host, wt, N, K, T):
    """llama.cpp/tests/test-backend-ops.cpp:2236-2318 (test_mul_mat) shapes and beyond: decode kernels, skinny GEMM, tcgen05 prefill."""
    W = H.make_weights(wt, N, K, seed=N + T, realistic_q4k=False)
    X = H.make_acts(T, K, seed=T)

    def build(g):
        a, b = g.tensor(wt, K, N), g.tensor(H.F32, K, T)
        return {"a": a, "b": b}, [(g.mul_mat(a, b), (T, N))]
    (got, ref), = run_both(host, build, {"a": W, "b": X})
    assert H.nmse(got, ref) <= H.NMSE_TOL
    H.assert_close(got, ref, f"{H.TYPE_NAME[wt]} N={N} K={K} T={T}")


def test_supports_op_matrix(host):
    G, backend, cpu = host
    g = Graph(backend)
    a4, af32, b, bh = g.tensor(H.Q4_0, 256, 32), g.tensor(H.F32, 256, 32), g.tensor(H.F32, 256, 3), g.tensor(H.F16, 256, 3)
    a3 = g.tensor(H.Q4_0, 256, 32, 2)
    b3 = g.tensor(H.F32, 256, 3, 2)
    assert G.ggml_backend_supports_op(backend, g.mul_mat(a4, b))
    assert not G.ggml_backend_supports_op(backend, g.mul_mat(af32, b))       # F32 weights: handed back, as llamafile_sgemm does here
    assert not G.ggml_backend_supports_op(backend, g.mul_mat(a4, bh))        # activations must be F32
    assert not G.ggml_backend_supports_op(backend, g.mul_mat(a3, b3))        # batched matmuls (attention) stay on another backend
    # the matmul's neighbours (fp32 matrices): taken; anything else is not
    nw = g.tensor(H.F32, 256)
    assert G.ggml_backend_supports_op(backend, g.add(b, b)) and G.ggml_backend_supports_op(backend, g.mul(b, nw))
    assert G.ggml_backend_supports_op(backend, g.rms_norm(b, 1e-5)) and G.ggml_backend_supports_op(backend, g.silu(b))
    assert not G.ggml_backend_supports_op(backend, g.add(b3, b3))            # batched tensors
    assert not G.ggml_backend_supports_op(backend, g.add(bh, bh))            # not fp32
    assert not G.ggml_backend_supports_op(backend, G.ggml_rms_norm(g.ctx, b3, 1e-5))
    assert G.ggml_backend_name(backend).startswith(b"TB200")
    g.free()


def decode_chain(K, F, V, types=None):
    """One transformer layer's matmuls + the output projection as a dependent chain of one-token MUL_MAT nodes."""
    shapes = {"wq": (K, K), "wk": (K // 4, K), "wv": (K // 4, K), "wo": (K, K), "gate": (F, K), "up": (F, K), "down": (K, F), "out": (V, K)}
    types = types or {}

    def build(g):
        t = {n: g.tensor(types.get(n, H.Q4_0), k, rows) for n, (rows, k) in shapes.items()}
        t["x"] = g.tensor(H.F32, K, 1)
        q, kk, v = g.mul_mat(t["wq"], t["x"]), g.mul_mat(t["wk"], t["x"]), g.mul_mat(t["wv"], t["x"])
        o = g.mul_mat(t["wo"], q)
        ga, up = g.mul_mat(t["gate"], o), g.mul_mat(t["up"], o)
        d = g.mul_mat(t["down"], ga)
        out = g.mul_mat(t["out"], d)
        outs = [(q, (1, K)), (kk, (1, K // 4)), (v, (1, K // 4)), (o, (1, K)), (ga, (1, F)), (up, (1, F)), (d, (1, K)), (out, (1, V))]
        return t, outs
    return shapes, build


def test_decode_chain_is_one_persistent_launch(tb, host):
    """Consecutive one-token Q4_0 matmuls run as ONE kernel launch (the decode plan), cached per graph; the values follow the inputs."""
    G, backend, cpu = host
    K, F, V = 512, 1280, 3000
    shapes, build = decode_chain(K, F, V)
    rng = np.random.default_rng(5)
    inputs = {n: H.make_weights(H.Q4_0, rows, k, seed=i) * 1 for i, (n, (rows, k)) in enumerate(shapes.items())}
    # scale the block scales up so that activations keep their magnitude through the chain
    for n, (rows, k) in shapes.items():
        blk = inputs[n].reshape(-1, 18)
        d = blk[:, :2].copy().view(np.float16).astype(np.float32) * (2.5 / (0.02 * np.sqrt(k)))
        blk[:, :2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    inputs["x"] = rng.standard_normal((1, K)).astype(np.float32)
    pairs = run_both(host, build, inputs)
    for i, (got, ref) in enumerate(pairs):
        H.assert_close(got, ref, f"chain output {i}")
    # launch accounting + replay on the TB200 backend
    g = Graph(backend)
    tensors, outs = build(g)
    g.build([t for t, _ in outs], weights_usage=True)
    for name, arr in inputs.items():
        g.set(tensors[name], arr)
    assert g.compute() == 0
    launches, plan_nodes = tb.backend_last_graph_stats(backend)
    assert plan_nodes == 8 and launches <= 9, (launches, plan_nodes)        # first compute: weight repacks + ONE plan launch
    first = g.get(outs[-1][0], outs[-1][1])
    assert g.compute() == 0
    launches, plan_nodes = tb.backend_last_graph_stats(backend)
    assert (launches, plan_nodes) == (1, 8)                                  # replay: one persistent kernel for the 8 matmuls
    assert np.array_equal(first, g.get(outs[-1][0], outs[-1][1]))
    x2 = rng.standard_normal((1, K)).astype(np.float32)
    g.set(tensors["x"], x2)
    assert g.compute() == 0
    assert not np.array_equal(first, g.get(outs[-1][0], outs[-1][1]))
    H.assert_close(g.get(outs[0][0], outs[0][1]), B.oracle().mul_mat(inputs["wq"], H.Q4_0, x2), "wq after new input")
    # new weights at the same tensor: the resident (repacked) copy must not survive
    Wn = H.make_weights(H.Q4_0, K, K, seed=99)
    g.set(tensors["wq"], Wn)
    assert g.compute() == 0
    H.assert_close(g.get(outs[0][0], outs[0][1]), B.oracle().mul_mat(Wn, H.Q4_0, x2), "wq after new weights")
    g.free()


def test_decode_chain_q4_K_and_mixed_types(tb, host):
    """A Q4_K chain runs as one persistent launch too; in a Q4_K_M-like mix (ffn_down and the output matrix Q6_K) the runs of
    plannable nodes end where the type changes: two plan launches + two one-shot matmuls (each = quantizer + kernel), values
    equal to ggml's CPU backend either way."""
    G, backend, cpu = host
    K, F, V = 512, 1280, 3000
    rng = np.random.default_rng(8)
    for types, want_plan_nodes in (({n: H.Q4_K for n in ("wq", "wk", "wv", "wo", "gate", "up", "down", "out")}, 8),
                                   ({**{n: H.Q4_K for n in ("wq", "wk", "wv", "wo", "gate", "up")}, "down": H.Q6_K, "out": H.Q6_K}, 6)):
        shapes, build = decode_chain(K, F, V, types)
        inputs = {}
        for i, (n, (rows, k)) in enumerate(shapes.items()):
            w = rng.standard_normal((rows, k)).astype(np.float32) * (1.6 / np.sqrt(k))    # keeps the hidden state's magnitude through the chain
            r = B.reference()
            inputs[n] = (r.quantize_q4_K(w) if types[n] == H.Q4_K else r.quantize_q6_K(w)) if r is not None else H.make_weights(types[n], rows, k, seed=i)
        inputs["x"] = rng.standard_normal((1, K)).astype(np.float32)
        pairs = run_both(host, build, inputs)
        for i, (got, ref) in enumerate(pairs):
            H.assert_close(got, ref, f"chain output {i} ({len(set(types.values()))} types)")
        g = Graph(backend)
        tensors, outs = build(g)
        g.build([t for t, _ in outs], weights_usage=True)
        for name, arr in inputs.items():
            g.set(tensors[name], arr)
        assert g.compute() == 0 and g.compute() == 0
        launches, plan_nodes = tb.backend_last_graph_stats(backend)
        assert plan_nodes == want_plan_nodes, (launches, plan_nodes)
        assert launches == (1 if want_plan_nodes == 8 else 1 + 2 * 2), (launches, plan_nodes)
        g.free()


def ffn_block(K, F, wt, tokens, extra_reader=False):
    """llm_build_norm + llm_build_ffn (LLM_FFN_SILU, LLM_FFN_PAR) + the residual add of a llama layer's second half, preceded by the
    attention output projection with ITS residual (llama.cpp/src/llama.cpp:8099-8225): x -> wo x + h -> rms_norm * w -> gate, up ->
    silu(gate) * up -> down + (wo x + h)."""
    shapes = {"wo": (K, K), "gate": (F, K), "up": (F, K), "down": (K, F)}

    def build(g):
        t = {n: g.tensor(wt, k, rows) for n, (rows, k) in shapes.items()}
        t["x"], t["h"], t["nw"] = g.tensor(H.F32, K, tokens), g.tensor(H.F32, K, tokens), g.tensor(H.F32, K)
        o = g.mul_mat(t["wo"], t["x"])
        ffn_inp = g.add(o, t["h"])
        cur = g.mul(g.rms_norm(ffn_inp, 1e-5), t["nw"])
        ga, up = g.mul_mat(t["gate"], cur), g.mul_mat(t["up"], cur)
        act = g.mul(g.silu(ga), up)
        d = g.mul_mat(t["down"], act)
        out = g.add(d, ffn_inp)
        outs = [(ffn_inp, (tokens, K)), (out, (tokens, K))]
        if extra_reader:   # something else reads the normalised activations: that pattern must then NOT be folded away
            outs.append((g.add(cur, t["h"]), (tokens, K)))
        return t, outs
    return shapes, build


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q4_K])
@pytest.mark.parametrize("tokens,extra_reader", [(1, False), (1, True), (5, False)])
def test_ffn_block_neighbours_folded_into_the_plan(tb, host, wt, tokens, extra_reader):
    """The matmul's graph neighbours on this backend (SURVEY.md §8 f4 through f1): at one token the whole block - 4 matmuls, RMSNorm,
    the norm weight, SiLU, the gate*up product and both residual adds, 10 nodes - is ONE persistent kernel launch with no fp32
    intermediate in memory; with a second reader of an intermediate, or more tokens, the same nodes run as kernels of their own.
    Values against ggml's CPU backend either way (SiLU's expf differs by a few ulp: the usual fp32 gate)."""
    G, backend, cpu = host
    K, F = 512, 1280
    shapes, build = ffn_block(K, F, wt, tokens, extra_reader)
    rng = np.random.default_rng(21)
    inputs = {}
    for i, (n, (rows, k)) in enumerate(shapes.items()):
        w = rng.standard_normal((rows, k)).astype(np.float32) * (1.2 / np.sqrt(k))
        inputs[n] = B.oracle().quantize_q4_0(w) if wt == H.Q4_0 else (B.reference().quantize_q4_K(w) if B.reference() is not None else H.make_weights(wt, rows, k, seed=i))
    inputs["x"] = rng.standard_normal((tokens, K)).astype(np.float32)
    inputs["h"] = rng.standard_normal((tokens, K)).astype(np.float32)
    inputs["nw"] = rng.uniform(0.5, 1.5, K).astype(np.float32)
    pairs = run_both(host, build, inputs)
    for i, (got, ref) in enumerate(pairs):
        H.assert_close(got, ref, f"ffn block output {i} tokens={tokens} extra_reader={extra_reader}")
    g = Graph(backend)
    tensors, outs = build(g)
    g.build([t for t, _ in outs], weights_usage=True)
    for name, arr in inputs.items():
        g.set(tensors[name], arr)
    assert g.compute() == 0 and g.compute() == 0
    launches, plan_nodes = tb.backend_last_graph_stats(backend)
    if tokens == 1 and not extra_reader:
        assert (launches, plan_nodes) == (1, 4), (launches, plan_nodes)
    elif tokens == 1:
        assert plan_nodes >= 2 and launches > 1, (launches, plan_nodes)      # the norm pattern runs unfused, the rest still plans
    else:
        assert plan_nodes == 0, (launches, plan_nodes)
    g.free()


def test_cached_plan_follows_a_changed_node_parameter(tb, host):
    """The plan cache is keyed by the run's nodes: addresses, op, first parameter word, row length. The same tensors at the same
    addresses with another rms_norm eps (here: the node's op_params rewritten in place between two computes, what a second graph
    from a reusing allocator looks like to the backend) must not replay the plan that folded the old eps."""
    import ctypes as C
    from llamafile_b200.sgemm import GgmlTensor
    G, backend, cpu = host
    K, F, wt = 512, 1280, H.Q4_0
    shapes = {"gate": (F, K), "up": (F, K)}
    rng = np.random.default_rng(44)
    inputs = {n: B.oracle().quantize_q4_0(rng.standard_normal((rows, k)).astype(np.float32) / np.sqrt(k)) for n, (rows, k) in shapes.items()}
    inputs["x"] = (rng.standard_normal((1, K)) * 1e-2).astype(np.float32)     # small values: eps matters (mean square 1e-4)
    inputs["nw"] = rng.uniform(0.5, 1.5, K).astype(np.float32)
    res = {}
    for bk, name in ((backend, "tb200"), (cpu, "cpu")):
        g = Graph(bk)
        t = {n: g.tensor(wt, k, rows) for n, (rows, k) in shapes.items()}
        t["x"], t["nw"] = g.tensor(H.F32, K, 1), g.tensor(H.F32, K)
        norm = g.rms_norm(t["x"], 1e-5)
        cur = g.mul(norm, t["nw"])
        ga, up = g.mul_mat(t["gate"], cur), g.mul_mat(t["up"], cur)
        g.build([ga, up], weights_usage=True)
        for n, arr in inputs.items():
            g.set(t[n], arr)
        outs = []
        for eps in (1e-5, 1e-3, 1e-5):
            C.cast(norm, C.POINTER(GgmlTensor)).contents.op_params_[0] = C.c_int32.from_buffer_copy(C.c_float(eps)).value
            assert g.compute() == 0 and g.compute() == 0
            if name == "tb200":
                launches, plan_nodes = tb.backend_last_graph_stats(backend)
                assert plan_nodes == 2, (launches, plan_nodes)      # both matmuls ran inside a plan (the norm + its weight as the group's prologue)
            outs.append(g.get(ga, (1, F)))
        res[name] = outs
        g.free()
    for i, (got, ref) in enumerate(zip(res["tb200"], res["cpu"])):
        H.assert_close(got, ref, f"eps case {i}")
    assert np.array_equal(res["tb200"][0], res["tb200"][2])
    assert np.abs(res["cpu"][0] - res["cpu"][1]).max() > 0.05 * np.abs(res["cpu"][0]).max(), "the test's eps values must change the result"


def test_sum_of_two_matmuls_of_one_group_is_not_folded(tb, host):
    """a.x + b.x: both matmuls share their input, i.e. one plan group without a barrier between them, so the add cannot become the
    residual epilogue of either (it would read the other's rows while they are being written); and a chain a.x -> + r -> b -> + (a.x + r)."""
    G, backend, cpu = host
    K = 512
    rng = np.random.default_rng(33)
    Wa, Wb = (B.oracle().quantize_q4_0(rng.standard_normal((K, K)).astype(np.float32) / np.sqrt(K)) for _ in range(2))

    def build(g):
        t = {"a": g.tensor(H.Q4_0, K, K), "b": g.tensor(H.Q4_0, K, K), "x": g.tensor(H.F32, K, 1), "r": g.tensor(H.F32, K, 1)}
        ya, yb = g.mul_mat(t["a"], t["x"]), g.mul_mat(t["b"], t["x"])
        s1 = g.add(ya, yb)
        s2 = g.add(s1, t["r"])
        yc = g.mul_mat(t["b"], s2)
        return t, [(s1, (1, K)), (g.add(yc, s2), (1, K))]
    inputs = {"a": Wa, "b": Wb, "x": rng.standard_normal((1, K)).astype(np.float32), "r": rng.standard_normal((1, K)).astype(np.float32)}
    for _ in range(3):
        for i, (got, ref) in enumerate(run_both(host, build, inputs)):
            H.assert_close(got, ref, f"output {i}")


@pytest.mark.parametrize("wt", [H.Q4_0, H.Q8_0, H.F16])
@pytest.mark.parametrize("tokens,strided_ids", [(1, False), (5, False), (19, True), (64, True), (300, False)])
def test_mul_mat_id_vs_cpu_backend(host, wt, tokens, strided_ids):
    """llama.cpp/tests/test-backend-ops.cpp:995-1062 (test_mul_mat_id): as [K, rows, E], b [K, 1, T], ids [used, T] (dense, or a
    view of the [E, T] top-k result as llama.cpp builds it)."""
    G, backend, cpu = host
    rows, K, E, used = 96, 512, 8, 2
    W = H.make_weights(wt, E * rows, K, seed=tokens)
    X = np.random.default_rng(tokens).standard_normal((tokens, 1, K)).astype(np.float32)
    plan = H.make_plan(tokens, E, used, seed=tokens)
    ids_full = np.zeros((tokens, E), np.int32)
    ids_full[:, :used] = plan

    def build(g):
        as_, b = g.tensor(wt, K, rows, E), g.tensor(H.F32, K, 1, tokens)
        if strided_ids:
            full = g.tensor(H.I32, E, tokens)
            ids = G.ggml_view_2d(g.ctx, full, used, tokens, E * 4, 0)
            t = {"as": as_, "b": b, "ids": full}
        else:
            ids = g.tensor(H.I32, used, tokens)
            t = {"as": as_, "b": b, "ids": ids}
        return t, [(g.mul_mat_id(as_, b, ids), (tokens, used, rows))]
    (got, ref), = run_both(host, build, {"as": W, "b": X, "ids": ids_full if strided_ids else plan})
    H.assert_close(got, ref, f"mul_mat_id {H.TYPE_NAME[wt]} tokens={tokens}")


def test_standalone_host_link(tb):
    """tb200_backend_link(NULL): the module's own minimal host services (what bench.py's e2e leg uses, no ggml in the process)."""
    import subprocess
    import sys
    code = ("import sys; sys.path.insert(0, %r); import llamafile_b200 as tb; assert tb.backend_link(None); b = tb.backend_init(0); "
            "assert tb.lib().tb200_backend_is_tb200(b); print('STANDALONE_OK')") % str(H.__file__.rsplit('/tests/', 1)[0])
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert "STANDALONE_OK" in r.stdout, r.stdout + r.stderr
