"""Per-token sparse-FFN stack: the host-side mirror of how the reference chains the path across layers.

Reference: llm_graph_context::build_sparse_ffn + the look-ahead predictor (src/llama-graph.cpp:932-946): layer 0 predicts
its own mask from x_0, and while layer l runs, the mask of layer l+1 is predicted from layer l's FFN input x_l.
The attention blocks between the FFNs are out of scope, so the x_l are inputs.

Sharding: with world_size G each rank holds the rows of n_ff / G neurons of every matrix (whole groups of
`g` consecutive neurons, dealt by placement.py, moved online by balancer.py), predicts / masks / computes only those, and the partial
y_l is summed over ranks with ONE allreduce per layer.

torch is plumbing only (device memory, streams, NCCL through torch.distributed); every compute step is a call into
libspif_b200.so.  There is no CPU fallback.
"""
import torch

from . import ops


class LayerWeights:
    """device-resident rows of one layer held by this rank (all byte tensors in ggml row layout)"""
    __slots__ = ("up", "gate", "down_t", "pred_up", "pred_down", "pred_down_b", "m", "pred_up_b", "up_b", "gate_b", "down_b")

    def __init__(self, up, gate, down_t, pred_up, pred_down, pred_down_b, m, pred_up_b=None, up_b=None, gate_b=None,
                 down_b=None):
        self.up, self.gate, self.down_t = up, gate, down_t
        self.pred_up, self.pred_down, self.pred_down_b = pred_up, pred_down, pred_down_b
        self.m = m
        # optional F32 biases (llama-graph.cpp:874, 1049-1059, 1136-1139); down_b on ONE shard only: it is added once per sum
        self.pred_up_b, self.up_b, self.gate_b, self.down_b = pred_up_b, up_b, gate_b, down_b


class FfnStack:
    """n_layer sparse FFNs + predictors for one decode step (n_tok tokens)"""

    def __init__(self, arch, w_type, p_type, n_embd, n_ff_local, r, layers, n_tok=1, group=None, device="cuda",
                 flags=0, exchange="nccl", topk=0):
        self.arch, self.w_type, self.p_type = arch, w_type, p_type
        self.n_embd, self.m, self.r, self.n_tok = n_embd, n_ff_local, r, n_tok
        self.layers = layers
        self.L = len(layers)
        self.group = group  # torch.distributed process group (None: single GPU)
        # per-layer sum over the shards: "peer" = fused into the layer's partial-sum kernel over NVLink peer memory
        # (spif_sparse_ffn_sharded), "nccl" = one ncclAllReduce per layer
        self.comm = None
        if group is None:
            self.comm = ops.ShardComm(n_tok * n_embd + r, None, device)  # world 1: the layer kernel's control words + h1
        if group is not None and exchange == "peer":
            try:
                comm = ops.ShardComm(n_tok * n_embd + r, group, device)
            except RuntimeError as e:  # symmetric memory not available here: every rank must take the same decision
                comm = None
                print(f"[sparkinfer_b200] peer exchange unavailable ({type(e).__name__}: {e}); using ncclAllReduce", flush=True)
            ok = torch.tensor([1 if comm is not None else 0], device=device)
            torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN, group=group)
            self.comm = comm if int(ok.item()) == 1 else None
        self.exchange = "none" if group is None else ("peer" if self.comm is not None else "nccl")
        if self.exchange == "peer" and getattr(self.comm, "multicast", 0):
            self.exchange = "peer+multicast"
        self.flags = flags
        self.ws = ops.Workspace(n_embd, n_ff_local, n_tok, device)
        self.x = torch.zeros((self.L, n_tok, n_embd), dtype=torch.float32, device=device)
        self.y = torch.zeros((self.L, n_tok, n_embd), dtype=torch.float32, device=device)
        self.p = torch.zeros((self.L, n_tok, n_ff_local), dtype=torch.float32, device=device)
        self.stats = torch.zeros((self.L, n_tok, 4), dtype=torch.int32, device=device)
        # top-k masking (Llama-3 / SiLU config): mask = the k largest predictor logits of this shard instead of p >= 0.5
        self.topk = int(topk)
        self.logits = torch.zeros((n_tok, n_ff_local), dtype=torch.float32, device=device) if self.topk > 0 else None
        self.graph = None
        self.launches_per_step = 0
        self.early_mask = True
        self._side = torch.cuda.Stream(device=device)
        # the predictor's hidden vector lives in the workspace: a predictor running concurrently with an FFN needs its own
        self.ws_pred = ops.Workspace(n_embd, n_ff_local, n_tok, device)
        # active-row lists handed from launch to launch by the look-ahead layer kernel (double-buffered)
        self.lists = [ops.ActiveList(n_ff_local, device), ops.ActiveList(n_ff_local, device)]

    # -- single pieces ------------------------------------------------------------------------------------------
    def predict(self, l, x_from):
        w = self.layers[l]
        if self.topk > 0:
            ops.predictor(w.pred_up, w.pred_up_b, w.pred_down, w.pred_down_b, self.x[x_from], self.p_type, self.r, self.m,
                          flags=self.flags, ws=self.ws_pred, out=self.p[l], logits_out=self.logits)
            ops.topk_mask(self.logits, self.topk, out=self.p[l])
            return 3
        ops.predictor(w.pred_up, w.pred_up_b, w.pred_down, w.pred_down_b, self.x[x_from], self.p_type, self.r, self.m,
                      flags=self.flags, ws=self.ws_pred, out=self.p[l])
        return 2  # kernels: predictor up GEMV, predictor down GEMV+sigmoid

    def ffn(self, l, want_stats=False, mask_stable=False, exchange=False):
        w = self.layers[l]
        fl = self.flags | (ops.FLAG_MASK_STABLE if mask_stable else 0)
        ops.sparse_ffn(self.arch, self.w_type, w.up, w.gate, w.down_t, self.x[l], self.p[l], self.m, up_b=w.up_b,
                       gate_b=w.gate_b, down_b=w.down_b, flags=fl, ws=self.ws, out=self.y[l], stats=self.stats[l] if want_stats else None,
                       comm=self.comm if (exchange and self.group is not None) else None)
        return 2  # kernels: fused gate/up/act/down, partial reduction (+ shard exchange)

    def allreduce(self, l):
        if self.group is not None and self.comm is None:
            torch.distributed.all_reduce(self.y[l], group=self.group)

    # -- one decode step ----------------------------------------------------------------------------------------
    def step_eager(self, want_stats=False):
        n = self.predict(0, 0)
        for l in range(self.L):
            if l + 1 < self.L:
                n += self.predict(l + 1, l)  # look-ahead, llama-graph.cpp:939-946
            # the mask of layer l was written >= 2 launches ago (the look-ahead predictor of l+1 sits in between)
            n += self.ffn(l, want_stats, mask_stable=(l + 1 < self.L) and self.early_mask, exchange=True)
            self.allreduce(l)
        self.launches_per_step = n
        return n

    def step_two_streams(self):
        """same work as step_eager, issued as two chains that only meet where the data does: the predictor chain
        (needs x only) runs on a side stream, FFN l waits for the event of predictor l.  Captured into one CUDA graph
        the two chains become parallel branches and share the SMs (the predictor kernel is register-light and uses no
        shared memory so that it fits next to the FFN kernel's TMA ring)."""
        main = torch.cuda.current_stream()
        side = self._side
        side.wait_stream(main)
        evs = []
        with torch.cuda.stream(side):
            n = self.predict(0, 0)
            e = torch.cuda.Event()
            e.record(side)
            evs.append(e)
            for l in range(self.L - 1):
                n += self.predict(l + 1, l)
                e = torch.cuda.Event()
                e.record(side)
                evs.append(e)
        for l in range(self.L):
            main.wait_event(evs[l])
            n += self.ffn(l, False, mask_stable=self.early_mask, exchange=True)  # full edge from predictor l
            self.allreduce(l)
        main.wait_stream(side)
        self.launches_per_step = n
        return n

    def step_overlap(self):
        """the reference's data edges with the one overlap they allow: predictor(l+1) reads x_l, which in a real decode exists
        only after FFN(l-1) (its y feeds the attention block that produces x_l), so predictor(l+1) starts after FFN(l-1) and
        runs BESIDE FFN(l) on a second stream; FFN(l) waits for predictor(l).  Unlike step_two_streams the predictor chain cannot
        run ahead of the FFN chain.  Used where the single-launch layer kernel does not apply (quantised FFN weights)."""
        main = torch.cuda.current_stream()
        side = self._side
        side.wait_stream(main)
        with torch.cuda.stream(side):
            n = self.predict(0, 0)
            ev_p = torch.cuda.Event()
            ev_p.record(side)
        ev_f_prev = None
        for l in range(self.L):
            ev_next = None
            if l + 1 < self.L:
                if ev_f_prev is not None:
                    side.wait_event(ev_f_prev)
                with torch.cuda.stream(side):
                    n += self.predict(l + 1, l)
                    ev_next = torch.cuda.Event()
                    ev_next.record(side)
            main.wait_event(ev_p)
            n += self.ffn(l, False, mask_stable=self.early_mask, exchange=True)  # full edge from predictor l
            self.allreduce(l)
            ev_f_prev = torch.cuda.Event()
            ev_f_prev.record(main)
            ev_p = ev_next
        main.wait_stream(side)
        self.launches_per_step = n
        return n

    # -- look-ahead layer kernel: FFN(l) + predictor(l+1) in ONE launch (spif_sparse_ffn_lookahead) ------------------
    def fused_supported(self):
        return (self.n_tok == 1 and self.comm is not None and self.w_type == self.p_type and self.r % 256 == 0
                and self.r <= self.n_embd and self.n_embd <= 8192 and self.w_type in (ops.GGML_TYPE_F16, ops.GGML_TYPE_BF16))

    def _lookahead(self, l_ffn, l_pred, x_from, want_stats=False):
        """one launch: the sparse FFN of layer l_ffn (None: none) and the predictor of layer l_pred (None: none), both
        reading x[x_from]; the predictor's list goes to lists[l_pred % 2], the FFN reads lists[l_ffn % 2]"""
        ffn = pred = None
        n = 1
        if l_ffn is not None:
            w = self.layers[l_ffn]
            ffn = dict(arch=self.arch, w_type=self.w_type, up=w.up, gate=w.gate, down_t=w.down_t, cur=self.x[l_ffn, 0],
                       n_ff=self.m, up_b=w.up_b, gate_b=w.gate_b, down_b=w.down_b, flags=self.flags, out=self.y[l_ffn, 0], stats=self.stats[l_ffn, 0] if want_stats else None)
        if l_pred is not None:
            w = self.layers[l_pred]
            pred = dict(pred_up=w.pred_up, pred_up_b=w.pred_up_b, pred_down=w.pred_down, pred_down_b=w.pred_down_b,
                        cur=self.x[x_from, 0],
                        p_type=self.p_type, r=self.r, m=self.m, flags=self.flags, out=self.p[l_pred, 0],
                        logits_out=self.logits[0] if self.topk > 0 else None)
        ops.sparse_ffn_lookahead(self.comm, self.ws, ffn=ffn, pred=pred,
                                 list_in=self.lists[l_ffn % 2] if l_ffn is not None else None,
                                 list_out=self.lists[l_pred % 2] if (l_pred is not None and self.topk == 0) else None)
        if l_pred is not None and self.topk > 0:  # top-k configs: the list comes from the k largest logits of this shard
            ops.topk_mask(self.logits, self.topk, out=self.p[l_pred])
            ops.mask_to_list(self.p[l_pred, 0], out=self.lists[l_pred % 2])
            n += 2
        return n

    def step_fused(self, want_stats=False):
        """the reference's decode order (llama-graph.cpp:932-946) with ONE launch per layer: predictor(0) from x_0, then
        FFN(l) + predictor(l+1) from x_l.  The true data edges hold by construction (single stream, launch l+1 after l)."""
        n = self._lookahead(None, 0, 0)
        for l in range(self.L):
            n += self._lookahead(l, l + 1 if l + 1 < self.L else None, l, want_stats)
        self.launches_per_step = n
        return n

    def ffn_only_fused(self, want_stats=False):
        """the sparse FFNs alone through the layer kernel (pred == NULL), lists rebuilt from the masks in self.p"""
        n = 0
        for l in range(self.L):
            ops.mask_to_list(self.p[l, 0], out=self.lists[l % 2])
            w = self.layers[l]
            ops.sparse_ffn_lookahead(self.comm, self.ws, ffn=dict(arch=self.arch, w_type=self.w_type, up=w.up, gate=w.gate,
                                                                  down_t=w.down_t, cur=self.x[l, 0], n_ff=self.m, up_b=w.up_b,
                                                                  gate_b=w.gate_b, down_b=w.down_b, flags=self.flags,
                                                                  out=self.y[l, 0], stats=self.stats[l, 0] if want_stats else None),
                                     list_in=self.lists[l % 2])
            n += 2
        return n

    def predict_only_fused(self):
        n = self._lookahead(None, 0, 0)
        for l in range(self.L - 1):
            n += self._lookahead(None, l + 1, l)
        return n

    def ffn_only_eager(self, want_stats=False):
        """the sparse FFNs alone, masks as left in self.p by the last full step"""
        n = 0
        for l in range(self.L):
            n += self.ffn(l, want_stats, mask_stable=self.early_mask)
        return n

    def predict_only_eager(self):
        n = self.predict(0, 0)
        for l in range(self.L - 1):
            n += self.predict(l + 1, l)
        return n

    def capture(self, fn=None):
        """capture fn (default: the full step) into a CUDA graph on a side stream; returns the graph"""
        fn = fn or self.step_eager
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            fn()  # warm-up on the capture stream (sets kernel attributes, NCCL channels)
            s.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                fn()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        if fn in (self.step_eager, self.step_fused):
            self.graph = g
        return g

    def step(self):
        if self.graph is None:
            self.capture()
        self.graph.replay()

    # -- from the model file ----------------------------------------------------------------------------------------------
    ARCH_OF_GGUF = {"prosparse-llama": ops.ARCH_PROSPARSE, "bamboo": ops.ARCH_RELU_GATED, "qwen2": ops.ARCH_RELU_GATED,
                    "opt": ops.ARCH_RELU_UP, "falcon": ops.ARCH_RELU_UP, "llama": ops.ARCH_SILU_GATED}

    @classmethod
    def from_gguf(cls, path, n_tok=1, device="cuda", flags=0, layers=None, group=None, rows_of_layer=None, exchange="peer",
                  topk=0, model_split=None):
        """the sparse-FFN stack of a model GGUF (sparkinfer_b200/convert.py or the reference's converter), resident on `device`.
        What the reference's loader does for these tensors (src/llama-model.cpp:2730-2768): blk.N.ffn_up / ffn_gate / ffn_down
        (transposed: one row per neuron), blk.N.ffn_pred_up / ffn_pred_down (+ .bias), rank from {arch}.pred_lora.
        layers: which blocks (default all).  Sharded (group != None): rows_of_layer[l] = ascending neuron rows this rank holds
        (placement.shard_rows); only those rows are read from the file; ffn_pred_up is replicated; b_down goes to rank 0.
        model_split: path of a model-split file (`-ms`): the rows are loaded in ITS order (blk.N.ffn_reorder_perms, new_row ->
        old_row: what the reference's cache manager does at load time, src/llama-sparkinfer.cpp:283-352), and with a group each
        rank takes its contiguous block of the reordered rows (model_split.shard_rows_from_split); replaces rows_of_layer.
        Rows are staged through pinned host memory.  -> (stack, info dict with norm weights `ffn_norm` [L, n_embd] and rms_eps)"""
        import numpy as np

        from .gguf_io import GGML_F32, ModelGguf
        with ModelGguf(path) as g:
            if g.arch not in cls.ARCH_OF_GGUF:
                raise ValueError(f"{path}: architecture {g.arch} has no sparse-FFN path")
            arch = cls.ARCH_OF_GGUF[g.arch]
            which = list(range(g.n_layer)) if layers is None else list(layers)
            ranks = {int(g.pred_rank[l]) for l in which}
            if len(ranks) != 1 or 0 in ranks:
                raise ValueError(f"{path}: the stack needs one predictor rank for all its layers, pred_lora gives {sorted(ranks)}")
            r = ranks.pop()
            rank0 = group is None or torch.distributed.get_rank(group) == 0
            if model_split is not None:
                from .model_split import read_split, shard_rows_from_split
                _, _, perms = read_split(model_split)
                if perms.shape[1] != g.n_ff or perms.shape[0] < g.n_layer:
                    raise ValueError(f"{model_split}: permutations of {perms.shape[1]} rows x {perms.shape[0]} layers do not fit {path}")
                rows_of_layer = shard_rows_from_split(perms, which, 0 if group is None else torch.distributed.get_rank(group),
                                                      1 if group is None else torch.distributed.get_world_size(group))

            def to_dev(name, rows=None, f32=False):
                t = g.find(name)
                if t is None:
                    return None
                if f32 and t.type != GGML_F32:
                    raise ValueError(f"{name}: bias / norm vectors are F32 in the file (type {t.type})")
                if rows is None:
                    host = g.read_rows(name)
                else:  # a shard's rows / a model-split order: one gather in the library
                    host = g.gather_rows(name, rows)
                pinned = torch.from_numpy(host).pin_memory()
                d = pinned.to(device, non_blocking=True).reshape(-1)
                return d.view(torch.float32) if f32 else d

            lw, norms, w_type, p_type, m = [], [], None, None, None
            for i, l in enumerate(which):
                rows = None if rows_of_layer is None else rows_of_layer[i]
                up_t, pu_t = g.find(f"blk.{l}.ffn_up.weight"), g.find(f"blk.{l}.ffn_pred_up.weight")
                dn_t, pd_t = g.find(f"blk.{l}.ffn_down.weight"), g.find(f"blk.{l}.ffn_pred_down.weight")
                if None in (up_t, pu_t, dn_t, pd_t):
                    raise ValueError(f"{path}: layer {l} lacks ffn_up / ffn_down / ffn_pred_up / ffn_pred_down")
                E, F = int(up_t.ne[0]), int(up_t.ne[1])
                if (int(dn_t.ne[0]), int(dn_t.ne[1])) != (E, F):
                    raise ValueError(f"blk.{l}.ffn_down.weight is {dn_t.ne[0]} x {dn_t.ne[1]}: not stored transposed (one row per "
                                     f"neuron, convert_hf_to_gguf.py:4640-4648)")
                if (int(pu_t.ne[0]), int(pu_t.ne[1]), int(pd_t.ne[0]), int(pd_t.ne[1])) != (E, r, r, F):
                    raise ValueError(f"layer {l}: predictor shapes do not fit n_embd {E}, rank {r}, n_ff {F}")
                if w_type is None:
                    w_type, p_type, m = int(up_t.type), int(pu_t.type), (F if rows is None else len(rows))
                if (int(up_t.type), int(dn_t.type), int(pu_t.type), int(pd_t.type)) != (w_type, w_type, p_type, p_type):
                    raise ValueError(f"layer {l}: mixed tensor types")
                gate = to_dev(f"blk.{l}.ffn_gate.weight", rows)
                if (gate is None) != (arch == ops.ARCH_RELU_UP):
                    raise ValueError(f"layer {l}: ffn_gate does not match architecture {g.arch}")
                lw.append(LayerWeights(
                    up=to_dev(f"blk.{l}.ffn_up.weight", rows), gate=gate, down_t=to_dev(f"blk.{l}.ffn_down.weight", rows),
                    pred_up=to_dev(f"blk.{l}.ffn_pred_up.weight"), pred_down=to_dev(f"blk.{l}.ffn_pred_down.weight", rows),
                    pred_down_b=to_dev(f"blk.{l}.ffn_pred_down.bias", rows, f32=True), m=m,
                    pred_up_b=to_dev(f"blk.{l}.ffn_pred_up.bias", f32=True), up_b=to_dev(f"blk.{l}.ffn_up.bias", rows, f32=True),
                    gate_b=to_dev(f"blk.{l}.ffn_gate.bias", rows, f32=True),
                    down_b=to_dev(f"blk.{l}.ffn_down.bias", f32=True) if rank0 else None))
                norms.append(to_dev(f"blk.{l}.ffn_norm.weight", f32=True))
            torch.cuda.synchronize(device)
            st = cls(arch, w_type, p_type, g.n_embd, m, r, lw, n_tok=n_tok, group=group, device=device, flags=flags,
                     exchange=exchange, topk=topk)
            info = {"arch": g.arch, "n_layer": len(which), "n_embd": g.n_embd, "n_ff": g.n_ff, "r": r, "rms_eps": g.rms_eps,
                    "ffn_norm": torch.stack(norms) if all(n is not None for n in norms) else None}
        return st, info
