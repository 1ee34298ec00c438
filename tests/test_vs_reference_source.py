"""CPU: host logic pinned to the REFERENCE'S OWN CODE where that code cannot be linked (it lives inside member functions of
src/llama-sparkinfer.cpp that need a loaded llama_model and a CUDA context): the statements are cut out of the reference source
AT TEST TIME, wrapped in a function with the variables they use, compiled with g++ and run beside the library on seeded inputs.
Nothing of the reference is stored in this repository; skipped where /root/reference is absent (the GPU box).

  * the cache budget split over the layers, ctor src/llama-sparkinfer.cpp:177-202  == spif_cache_group_budget
  * reorder_tensor_2d / _1d, src/llama-sparkinfer.cpp:293-322                    == spif_reorder_rows
  * the decay adaptation of SingleThreadExecutor::make_anchor, ggml/include/ggml-sparkinfer.hpp:143-175 (header INCLUDED, the
    real executor thread driven)                                                 == spif_balancer_plan's lambda trajectory
  * the swap-list walk of sparkinfer_reload_plan, src/llama-sparkinfer.cpp:45-91: which (incoming group, slot) pairs a
    wanted-set / resident-set difference produces == the pairing rule the shard balancer keeps (ascending order on both
    sides, the incoming group takes the slot of the outgoing one)."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

REF = os.environ.get("SPIF_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "src", "llama-sparkinfer.cpp")
pytestmark = pytest.mark.skipif(not os.path.exists(SRC), reason="reference tree absent")


def cut(text, first_marker, end_marker):
    """the statements from the line containing first_marker up to (not including) the line containing end_marker"""
    lines = text.splitlines()
    a = next(i for i, ln in enumerate(lines) if first_marker in ln)
    b = next(i for i, ln in enumerate(lines) if i > a and end_marker in ln)
    return "\n".join(lines[a:b])


def build(tmp_path, name, code):
    src = tmp_path / f"{name}.cpp"
    src.write_text(code)
    so = tmp_path / f"lib{name}.so"
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-o", str(so), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return ctypes.CDLL(str(so))


def test_cache_budget_equals_the_reference_ctor_loops(tmp_path):
    from sparkinfer_b200 import _lib
    body = cut(open(SRC).read(), "const int  n_group_cache_budget = vram_budget", "auto create_tensor = ")
    assert "ffn_norm_pattern[il]" in body and "before_rr" in body, "the reference's loops moved: update the markers"
    lib = build(tmp_path, "refbudget", f"""
#include <algorithm>
#include <cstdint>
#include <vector>
extern "C" void ref_budget(int64_t vram_budget, int64_t n_bytes_group, int has_ffn_gate, const float * ffn_norm_pattern,
                           int n_layer, int n_group, int * out) {{
{body}
    for (int il = 0; il < n_layer; ++il) {{
        out[il] = n_group_cache[il];
    }}
}}
""")
    L = _lib.load()
    rng = np.random.default_rng(11)
    for case in range(300):
        n_layer = int(rng.integers(1, 81))
        n_group = int(rng.choice([64, 432, 864, 1024]))
        gate = int(rng.integers(0, 2))
        group_bytes = int(rng.choice([16 * 10240, 16 * 5440, 8 * 16384, 16 * 2880]))
        pattern = rng.random(n_layer).astype(np.float32) + 0.05
        pattern /= pattern.sum()
        full = group_bytes * (3 if gate else 2) * n_group * n_layer
        vram = int(full * rng.choice([0.0, 0.03, 0.25, 0.6, 0.97, 1.0, 1.7]))
        want = np.zeros(n_layer, dtype=np.int32)
        got = np.full(n_layer, -7, dtype=np.int32)
        lib.ref_budget(ctypes.c_int64(vram), ctypes.c_int64(group_bytes), gate, pattern.ctypes.data_as(ctypes.c_void_p), n_layer,
                       n_group, want.ctypes.data_as(ctypes.c_void_p))
        assert L.spif_cache_group_budget(vram, group_bytes, gate, pattern.ctypes.data, n_layer, n_group, got.ctypes.data) == 0
        assert np.array_equal(got, want), (case, n_layer, n_group, gate, vram, got, want)


def test_reload_plan_walk_equals_the_reference(tmp_path):
    """sparkinfer_reload_plan's walk, cut from the reference and run on plain arrays: for wanted-but-absent groups W and
    resident-but-unwanted groups C (both as 0/1 vectors over the groups) it pairs the k-th W with the k-th C in ascending order
    and hands the incoming group the outgoing group's cache slot; slot count, neuron_idx and neuron_mask stay consistent.
    This pins what "a swap" means in the reference; the shard balancer keeps the same rule per move (group_in takes the slot
    group_out leaves, buffers never resize), which tests/test_sharding.py::test_balancer_state_machine_c_abi checks on
    spif_balancer_plan / _commit."""
    text = open(SRC).read()
    body = cut(text, "int weight_idx = 0;", "reload_records.push_back")
    assert "group_maps[cache_idx]" in body and "reload_plan[reload_count]" in body, "the reference's walk moved: update the markers"
    lib = build(tmp_path, "refplan", f"""
#include <cstdint>
struct copy_pair {{ int weight_idx, cache_idx; }};
extern "C" int ref_plan(const float * weight_only, const float * cache_only, int32_t * neuron_idx, int32_t * group_maps,
                        int32_t * neuron_mask, int g, int n_g, copy_pair * reload_plan) {{
    int reload_count = 0;
{body}
    return reload_count;
}}
""")
    rng = np.random.default_rng(5)
    for case in range(200):
        n_g, g = int(rng.choice([8, 32, 64])), int(rng.choice([1, 4, 16]))
        m_g = int(rng.integers(1, n_g))
        resident = np.sort(rng.choice(n_g, m_g, replace=False))                 # groups in the cache, slot s holds resident[s]
        group_maps = np.full(n_g, -1, dtype=np.int32)
        group_maps[resident] = np.arange(m_g)
        neuron_mask = np.repeat((group_maps >= 0).astype(np.int32), g)
        neuron_idx = (resident[:, None] * g + np.arange(g)[None, :]).reshape(-1).astype(np.int32)
        k = int(rng.integers(0, min(m_g, n_g - m_g) + 1))
        outgoing = np.sort(rng.choice(resident, k, replace=False))
        incoming = np.sort(rng.choice(np.setdiff1d(np.arange(n_g), resident), k, replace=False))
        weight_only = np.zeros(n_g, dtype=np.float32)
        cache_only = np.zeros(n_g, dtype=np.float32)
        weight_only[incoming], cache_only[outgoing] = 1, 1
        plan = np.zeros((max(k, 1), 2), dtype=np.int32)
        slots_before = group_maps.copy()
        n = lib.ref_plan(weight_only.ctypes.data_as(ctypes.c_void_p), cache_only.ctypes.data_as(ctypes.c_void_p),
                         neuron_idx.ctypes.data_as(ctypes.c_void_p), group_maps.ctypes.data_as(ctypes.c_void_p),
                         neuron_mask.ctypes.data_as(ctypes.c_void_p), g, n_g, plan.ctypes.data_as(ctypes.c_void_p))
        assert n == k
        # the rule the balancer keeps: k-th incoming group (ascending) takes the slot of the k-th outgoing group (ascending)
        for i in range(k):
            assert plan[i, 0] == incoming[i] and plan[i, 1] == slots_before[outgoing[i]]
            assert group_maps[incoming[i]] == slots_before[outgoing[i]] and group_maps[outgoing[i]] == -1
            assert np.array_equal(neuron_idx[plan[i, 1] * g:(plan[i, 1] + 1) * g], incoming[i] * g + np.arange(g))
        assert np.array_equal(np.repeat((group_maps >= 0).astype(np.int32), g), neuron_mask)
        assert (group_maps >= 0).sum() == m_g                                   # slots never resize


def test_row_reorder_equals_the_reference_lambdas(tmp_path):
    """reorder_tensor_2d / reorder_tensor_1d (src/llama-sparkinfer.cpp:293-322), cut from the reference and compiled over a
    minimal stand-in for ggml_tensor (only the fields and helpers the lambdas touch) == spif_reorder_rows on the same bytes."""
    from sparkinfer_b200 import _lib
    body = cut(open(SRC).read(), "auto reorder_tensor_2d = [&]", "auto reorder_if_exists = [&]")
    assert "perm[new_row]" in body and "perm[new_i]" in body, "the reference's lambdas moved: update the markers"
    lib = build(tmp_path, "refreorder", f"""
#include <cstdint>
#include <cstring>
#include <vector>
struct ggml_tensor {{ int type; int64_t ne[4]; size_t nb[4]; void * data; }};
static size_t ggml_row_size(int type, int64_t n) {{   // the four types of the path (ggml.c type traits)
    return type == 0 ? 4 * n : type == 1 || type == 30 ? 2 * n : type == 8 ? n / 32 * 34 : type == 2 ? n / 32 * 18 : 0;
}}
static int64_t ggml_nrows(const ggml_tensor * t) {{ return t->ne[1] * t->ne[2] * t->ne[3]; }}
static size_t  ggml_nbytes(const ggml_tensor * t) {{ return ggml_row_size(t->type, t->ne[0]) * ggml_nrows(t); }}
static void ggml_backend_tensor_get(const ggml_tensor * t, void * dst, size_t off, size_t n) {{ memcpy(dst, (char *) t->data + off, n); }}
static void ggml_backend_tensor_set(ggml_tensor * t, const void * src, size_t off, size_t n) {{ memcpy((char *) t->data + off, src, n); }}
extern "C" void ref_reorder(int type, int64_t ne0, int64_t ne1, void * data, const int32_t * perm_in, int one_d) {{
    ggml_tensor tensor_obj = {{ type, {{ ne0, ne1, 1, 1 }}, {{ 0, 0, 0, 0 }}, data }};
    tensor_obj.nb[0] = ggml_row_size(type, 1);
    tensor_obj.nb[1] = ggml_row_size(type, ne0);
    std::vector<int32_t> perm_v(perm_in, perm_in + (one_d ? ne0 : ne1));
    std::vector<uint8_t> buf_src(ggml_nbytes(&tensor_obj)), buf_dst(ggml_nbytes(&tensor_obj));
    auto * src_buf = buf_src.data();
    auto * dst_buf = buf_dst.data();
{body}
    if (one_d) {{
        reorder_tensor_1d(&tensor_obj, perm_v);
    }} else {{
        reorder_tensor_2d(&tensor_obj, perm_v);
    }}
}}
""")
    L = _lib.load()
    rng = np.random.default_rng(21)
    for type_, cols in ((1, 320), (30, 64), (8, 256), (2, 512), (0, 96)):
        n = int(rng.choice([64, 432, 1024]))
        rb = L.spif_row_size(type_, cols)
        src = rng.integers(0, 256, (n, rb), dtype=np.uint8)
        perm = rng.permutation(n).astype(np.int32)
        want = src.copy()
        lib.ref_reorder(type_, ctypes.c_int64(cols), ctypes.c_int64(n), want.ctypes.data_as(ctypes.c_void_p),
                        perm.ctypes.data_as(ctypes.c_void_p), 0)
        got = np.empty_like(src)
        assert L.spif_reorder_rows(src.ctypes.data, got.ctypes.data, perm.ctypes.data, n, rb) == 0
        assert np.array_equal(got, want) and np.array_equal(got, src[perm])
    bias = rng.normal(size=864).astype(np.float32)                              # 1-D: a bias / score vector, 4 bytes per row
    perm = rng.permutation(864).astype(np.int32)
    want = bias.copy()
    lib.ref_reorder(0, ctypes.c_int64(864), ctypes.c_int64(1), want.ctypes.data_as(ctypes.c_void_p),
                    perm.ctypes.data_as(ctypes.c_void_p), 1)
    got = np.empty_like(bias)
    assert L.spif_reorder_rows(bias.ctypes.data, got.ctypes.data, perm.ctypes.data, 864, 4) == 0
    assert np.array_equal(got, want) and np.array_equal(got, bias[perm])


REFLIB = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libspif_ref.so")


@pytest.mark.skipif(not os.path.exists(REFLIB), reason="oracle/_ref not built")
def test_decay_adaptation_equals_the_reference_executor(tmp_path):
    """The adaptive decay of the balancer (a14).  The reference adapts it in SingleThreadExecutor::make_anchor
    (ggml/include/ggml-sparkinfer.hpp:143-175): lambda *= 1 +- dx, clamped to [0.05, 0.95], raised when work arrived while the
    reload was still running.  Here the reference's header itself is INCLUDED (not restated) in a small harness that drives the
    real executor thread through a raise / lower sequence; spif_balancer_plan is driven through the same sequence by its own
    trigger (the promised balance held / did not hold).  Same arithmetic, same start value and step from the reference's
    constants: the trajectories must be bit-identical, including both clamps."""
    from sparkinfer_b200 import _lib
    src = tmp_path / "reflambda.cpp"
    src.write_text("""
#include "ggml-sparkinfer.hpp"
#include <future>
extern "C" float ref_init_decay(void) { return k_spif_init_dfr_decay; }
extern "C" float ref_dx_decay(void) { return k_spif_dx_dfr_decay; }
extern "C" void ref_lambda_walk(float * decay2, const int * raise, int n, float * out) {
    SingleThreadExecutor ex;
    const auto W = SingleThreadExecutor::SPIF_WAIT_MUL_MAT_SPARSE;
    for (int i = 0; i < n; ++i) {
        if (raise[i]) {  // a kernel is submitted while the anchor is still active: it lands in `pending`
            std::promise<void> gate;
            auto gf = gate.get_future().share();
            ex.post([gf] { gf.wait(); });          // keeps the worker busy until the submit below has happened
            ex.make_anchor(W, decay2);
            auto fut = ex.submit(W, [] {});
            gate.set_value();
            fut.get();                              // runs after the anchor task moved it: decay2 is updated by then
        } else {         // nothing arrives before the anchor task runs
            ex.make_anchor(W, decay2);
            std::promise<void> done;
            auto df = done.get_future();
            ex.post([&done] { done.set_value(); });
            df.get();
        }
        out[i] = decay2[0];
    }
}
""")
    so = tmp_path / "libreflambda.so"
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-shared", "-fPIC", "-pthread", "-I", os.path.join(REF, "ggml", "include"),
                        "-o", str(so), str(src), "-L", os.path.dirname(REFLIB), "-lspif_ref", f"-Wl,-rpath,{os.path.dirname(REFLIB)}"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    lib = ctypes.CDLL(str(so))
    lib.ref_init_decay.restype = ctypes.c_float
    lib.ref_dx_decay.restype = ctypes.c_float
    lam0, dx = lib.ref_init_decay(), lib.ref_dx_decay()
    assert abs(lam0 - 0.67) < 1e-6 and abs(dx - 0.05) < 1e-6
    rng = np.random.default_rng(2)
    seq = np.concatenate([np.ones(10), np.zeros(70), np.ones(5), rng.integers(0, 2, 60)]).astype(np.int32)  # both clamps, then mixed
    decay = np.array([lam0, 1 - lam0], dtype=np.float32)
    want = np.zeros(seq.size, dtype=np.float32)
    lib.ref_lambda_walk(decay.ctypes.data_as(ctypes.c_void_p), seq.ctypes.data_as(ctypes.c_void_p), int(seq.size),
                        want.ctypes.data_as(ctypes.c_void_p))
    assert want.max() == np.float32(0.95) and want.min() == np.float32(0.05)

    # the library: 1 layer, 4 groups on 2 shards, no swaps allowed (max_swaps = 0), so every plan promises the balance it sees
    L = _lib.load()
    sog = np.array([[0, 0, 1, 1]], dtype=np.int32)
    h = ctypes.c_void_p()
    assert L.spif_balancer_create(1, 4, 2, 0, sog.ctypes.data, lam0, dx, ctypes.c_float(0.02), ctypes.byref(h)) == 0
    n_moves, n_swaps = ctypes.c_int(), ctypes.c_int()
    ib, ia = ctypes.c_float(), ctypes.c_float()

    def plan(a):  # shard 0 carries `a` per slot, shard 1 carries 1: imbalance 2a / (a + 1)
        scores = np.array([[[a, a]], [[1, 1]]], dtype=np.float32)
        assert L.spif_balancer_plan(h, scores.ctypes.data, 0, ctypes.c_float(0.0), None, 0, ctypes.byref(n_moves), ctypes.byref(ib),
                                    ctypes.byref(ia), ctypes.byref(n_swaps)) == 0
        assert n_moves.value == 0
        assert L.spif_balancer_commit(h) == 0
        lam = ctypes.c_float()
        assert L.spif_balancer_state(h, ctypes.byref(lam), None, None) == 0
        return lam.value

    a = 1.0
    assert plan(a) == lam0                      # the first plan has no promise to check: no adaptation
    got = []
    for up in seq:
        prev = 2 * a / (a + 1)
        a = a * 1.2 if up else 1.0              # the promised balance is missed by more than drift_tol / holds
        assert not up or 2 * a / (a + 1) > prev * 1.025, "run of raises too long for the imbalance range of this toy layout"
        got.append(plan(a))
    L.spif_balancer_destroy(h)
    assert np.array_equal(np.array(got, dtype=np.float32), want), (got[:12], want[:12])
