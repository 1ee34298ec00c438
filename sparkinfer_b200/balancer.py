"""Online shard balancer: the execution side of placement.py.

Reference: while decoding, build_dfr (src/llama-graph.cpp:910-930) folds every layer's mask into a decayed firing rate per
neuron group, and every `reload_interval` tokens sparkinfer_reload_plan (src/llama-sparkinfer.cpp:45-91) turns the scores
into copy pairs that ggml_sparkinfer_reload executes (host -> GPU cache slots, one group of g rows of up/gate/down and of
the predictor per pair).  Here every group is resident on exactly one GPU shard, so a copy pair becomes a SWAP between
the most- and least-loaded shards: the two ranks exchange the g rows of every matrix of the layer (NCCL send/recv over
NVLink) and each writes the incoming group into the slot the outgoing one leaves -- buffers never resize, tensor
addresses never change, captured CUDA graphs stay valid.

Division of labour:
  observe()    one spif_dfr_update launch per layer on this rank's masks  (CUDA, C-ABI)
  slot_scores  all_gather of the per-slot scores (torch.distributed)
  rebalance()  spif_balancer_plan (host C++ in the library: slot tables, per-layer spif_rebalance_plan, adaptive decay, the
               moves of this rank) -> the row exchange (NCCL send/recv) -> spif_balancer_commit (the anchor)
The plan only depends on gathered data, so every rank derives the same one without a coordinator.
"""
import ctypes

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, placement


def _row_bytes(t, rows):
    n = t.numel() * t.element_size()
    assert n % rows == 0
    return n // rows


class Move(ctypes.Structure):
    """struct spif_move (include/spif_b200.h)"""
    _fields_ = [("layer", ctypes.c_int32), ("round", ctypes.c_int32), ("slot", ctypes.c_int32), ("peer", ctypes.c_int32),
                ("group_out", ctypes.c_int32), ("group_in", ctypes.c_int32)]


class ShardBalancer:
    """Executor around the library's spif_balancer (host C++: slot tables of every layer and shard, swap plans, adaptive decay,
    pending / committed state).  This class owns only the device side: the decayed firing rates (one spif_dfr_update launch per
    layer), the all-gather of the scores, and the row exchange of the planned moves over NCCL send/recv."""

    LAYER_CHUNK = 16

    def __init__(self, layers, shard_of_group, rank, world, group=None, g=16, lam=0.9, lam_step=0.05, drift_tol=0.02):
        """layers: this rank's LayerWeights (m rows each); shard_of_group: per layer int array [n_g] (same on all ranks).
        Slot s of shard r initially holds the s-th smallest group id placed on r (placement.shard_rows order)."""
        self.layers, self.rank, self.world, self.group = layers, rank, world, group
        assert all(getattr(w, "up_b", None) is None and getattr(w, "gate_b", None) is None for w in layers), \
            "the balancer moves weight rows, predictor-down rows and their bias; per-neuron FFN biases (OPT) are not moved"
        self.g = g
        self.L = len(layers)
        self.n_slot = layers[0].m // g
        self.n_g = self.n_slot * world
        sog = np.ascontiguousarray(np.stack([np.asarray(s) for s in shard_of_group]), dtype=np.int32)
        assert sog.shape == (self.L, self.n_g), "equal group count per shard"
        self._lib = _lib.load()
        self._h = ctypes.c_void_p()
        _lib.check(self._lib.spif_balancer_create(self.L, self.n_g, world, rank, sog.ctypes.data, float(lam), float(lam_step),
                                                  float(drift_tol), ctypes.byref(self._h)), "spif_balancer_create")
        dev = layers[0].up.device
        self.dfr = torch.zeros((self.L, self.n_slot), dtype=torch.float32, device=dev)  # per local slot
        self.tokens_seen = 0
        self.bytes_moved = 0

    def __del__(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.spif_balancer_destroy(self._h)
            self._h = None

    # -- bookkeeping (state lives in the library) ----------------------------------------------------------------------
    @property
    def slots(self):
        """per layer int32 [world, n_slot]: the group id held by each slot of each shard (committed state)"""
        out = []
        for l in range(self.L):
            tab = np.empty((self.world, self.n_slot), dtype=np.int32)
            _lib.check(self._lib.spif_balancer_slots(self._h, l, tab.ctypes.data), "spif_balancer_slots")
            out.append(tab)
        return out

    def _state(self):
        lam, swaps, pend = ctypes.c_float(0), ctypes.c_int64(0), ctypes.c_int(0)
        _lib.check(self._lib.spif_balancer_state(self._h, ctypes.byref(lam), ctypes.byref(swaps), ctypes.byref(pend)),
                   "spif_balancer_state")
        return lam.value, swaps.value, pend.value

    @property
    def lam(self):
        return self._state()[0]

    @property
    def swaps_done(self):
        return self._state()[1]

    def shard_of_group(self, l):
        tab = self.slots[l]
        sog = np.empty(tab.size, dtype=np.int32)
        for r in range(self.world):
            sog[tab[r]] = r
        return sog

    def local_groups(self, l):
        return self.slots[l][self.rank]

    # -- observe: decayed firing rate per slot ----------------------------------------------------------------
    def observe(self, masks):
        """masks: [L, n_tok, m] activation probabilities of this rank's rows (FfnStack.p).  One launch per layer."""
        from . import ops
        lam = self.lam
        for l in range(self.L):
            ops.dfr_update(masks[l], self.dfr[l], self.g, lam)
        self.tokens_seen += int(masks.shape[1])
        return self.L

    def slot_scores(self):
        """-> float32 [world, L, n_slot]: every shard's per-slot scores (all-gathered)"""
        if self.world > 1:
            parts = [torch.empty_like(self.dfr) for _ in range(self.world)]
            dist.all_gather(parts, self.dfr, group=self.group)
            return np.ascontiguousarray(torch.stack(parts).cpu().numpy(), dtype=np.float32)
        return np.ascontiguousarray(self.dfr.cpu().numpy()[None], dtype=np.float32)

    def scores(self, allv=None):
        """-> per layer float64 [n_g]: the score of every group, assembled from every shard's slots"""
        allv = self.slot_scores() if allv is None else allv
        out, slots = [], self.slots
        for l in range(self.L):
            d = np.zeros(self.n_g, dtype=np.float64)
            for r in range(self.world):
                d[slots[l][r]] = allv[r, l]
            out.append(d)
        return out

    def imbalance(self, scores=None):
        """per layer max_s / mean_s of the summed scores (placement.imbalance)"""
        scores = scores if scores is not None else self.scores()
        return [placement.imbalance(scores[l], self.shard_of_group(l), self.world) for l in range(self.L)]

    # -- execute ----------------------------------------------------------------------------------------------------
    def _pieces(self, l):
        """the per-neuron arrays of layer l that travel with a group: (flat byte view, bytes per neuron)"""
        w = self.layers[l]
        out = []
        for t in (w.up, w.gate, w.down_t, w.pred_down, w.pred_down_b):
            if t is None:
                continue
            b = t.view(torch.uint8).reshape(-1)
            out.append((b, _row_bytes(b, w.m)))
        return out

    def _pack(self, l, slot):
        parts = [b[slot * self.g * rb:(slot + 1) * self.g * rb] for b, rb in self._pieces(l)]
        parts.append(self.dfr[l, slot:slot + 1].view(torch.uint8).reshape(-1))  # the score travels with the group
        return torch.cat(parts)

    def _unpack(self, l, slot, buf):
        o = 0
        for b, rb in self._pieces(l):
            n = self.g * rb
            b[slot * self.g * rb:(slot + 1) * self.g * rb].copy_(buf[o:o + n])
            o += n
        self.dfr[l, slot:slot + 1].view(torch.uint8).reshape(-1).copy_(buf[o:o + 4])

    # -- windowed form: plan -> exec in small windows between decode steps -> anchor ------------------------------------
    # (The first cut ran the whole exchange on a side stream next to the decode graphs: NCCL's send/recv kernels then sit on
    # SMs waiting for the peer while the persistent layer kernel needs one CTA on EVERY SM -- rounds with many moves stalled
    # for seconds, profiles/r2_balancer_async.txt.  Windows in stream order between two steps cannot interlock.)
    def rebalance_begin(self, max_swaps=8, min_gain=0.0):
        """plan (spif_balancer_plan); nothing moves yet.  Then call rebalance_step() after every decode step until it returns
        0, and rebalance_finish().  The reference's plan -> exec windows (SPIF_RELOAD_WINDOW_SIZE) -> anchor,
        ggml-cuda.cu:2499-2606."""
        assert getattr(self, "_staged", None) is None, "finish the previous round first"
        allv = self.slot_scores()
        cap = max(1, 2 * int(max_swaps) * self.L)
        moves = (Move * cap)()
        n, tot = ctypes.c_int(0), ctypes.c_int(0)
        before, after = ctypes.c_float(0), ctypes.c_float(0)
        _lib.check(self._lib.spif_balancer_plan(self._h, allv.ctypes.data, int(max_swaps), float(min_gain), moves, cap,
                                                ctypes.byref(n), ctypes.byref(before), ctypes.byref(after), ctypes.byref(tot)),
                   "spif_balancer_plan")
        self._moves, self._next = [moves[i] for i in range(n.value)], 0
        self._staged = {}   # (layer, slot) -> the group that will live there after the anchor
        # windows are cut by (round, layer chunk) -- a global rule, so that both ends of every move agree -- and every rank
        # walks ALL windows, also the ones in which it has nothing to move
        self._windows = [(k, c) for k in range(int(max_swaps)) for c in range(0, self.L, self.LAYER_CHUNK)]
        self._w = 0
        self._info = {"imbalance_before": float(before.value), "imbalance_after": float(after.value), "swaps": int(tot.value),
                      "bytes_sent_this_rank": 0}
        return dict(self._info)

    def rebalance_step(self):
        """one window of the exchange on the CURRENT stream (call it between two decode steps): outgoing groups are packed from
        the live slots, incoming groups land in staging buffers.  -> windows left"""
        assert getattr(self, "_staged", None) is not None
        # exactly ONE window per call, empty or not: with more than two shards every rank must be in the same window after the
        # same decode step (a rank that ran ahead would wait in a receive for a peer whose next decode step waits for it)
        if self._w < len(self._windows):
            k, c0 = self._windows[self._w]
            self._w += 1
            ops_, batch = [], []
            while (self._next < len(self._moves) and self._moves[self._next].round == k
                   and c0 <= self._moves[self._next].layer < c0 + self.LAYER_CHUNK):
                mv = self._moves[self._next]
                self._next += 1
                key = (mv.layer, mv.slot)
                out = self._staged[key] if key in self._staged else self._pack(mv.layer, mv.slot)  # arrived in an earlier round
                inc = torch.empty_like(out)
                ops_.append(dist.P2POp(dist.isend, out, mv.peer, group=self.group))
                ops_.append(dist.P2POp(dist.irecv, inc, mv.peer, group=self.group))
                batch.append((key, inc, out))
                self._info["bytes_sent_this_rank"] += out.numel()
            if ops_ and self.world > 1:
                for r in dist.batch_isend_irecv(ops_):
                    r.wait()  # stream-level for NCCL: the next decode step is ordered behind the window, the host is not blocked
                for key, inc, _keep in batch:
                    self._staged[key] = inc
        return len(self._windows) - self._w

    def rebalance_finish(self):
        """the anchor: remaining windows, then the staged groups are copied into their slots (device-to-device, in stream order
        between two decode steps) and the library's pending tables become current."""
        assert getattr(self, "_staged", None) is not None
        while self.rebalance_step() > 0:
            pass
        assert self._next == len(self._moves)
        for (l, slot), inc in self._staged.items():
            self._unpack(l, slot, inc)
        _lib.check(self._lib.spif_balancer_commit(self._h), "spif_balancer_commit")
        self.bytes_moved += self._info["bytes_sent_this_rank"]
        info = dict(self._info, **{"lambda": self.lam})
        self._staged = None
        return info

    def apply(self, moves):
        """moves: this rank's spif_move list, sorted by (round, layer) by the library.  One send/recv batch = the moves of one
        round within LAYER_CHUNK consecutive layers (bounds the ops per NCCL group and the staging memory; the chunking is by
        (round, layer), so both ends of every pair cut their op lists at the same places)."""
        moved, i = 0, 0
        while i < len(moves):
            k, c0 = moves[i].round, moves[i].layer // self.LAYER_CHUNK
            ops_, post = [], []
            while i < len(moves) and moves[i].round == k and moves[i].layer // self.LAYER_CHUNK == c0:
                mv = moves[i]
                out = self._pack(mv.layer, mv.slot)
                inc = torch.empty_like(out)
                ops_.append(dist.P2POp(dist.isend, out, mv.peer, group=self.group))
                ops_.append(dist.P2POp(dist.irecv, inc, mv.peer, group=self.group))
                post.append((mv.layer, mv.slot, inc, out))
                moved += out.numel()
                i += 1
            for r in dist.batch_isend_irecv(ops_):
                r.wait()
            for l, slot, inc, _keep in post:
                self._unpack(l, slot, inc)
        self.bytes_moved += moved
        return moved

    def rebalance(self, max_swaps=8, min_gain=0.0):
        """gather scores -> spif_balancer_plan (plans of every layer, adaptive decay, pending tables) -> row exchange ->
        spif_balancer_commit.  -> dict(lambda, imbalance_before, imbalance_after (planned), swaps, bytes)"""
        allv = self.slot_scores()
        cap = max(1, 2 * int(max_swaps) * self.L)
        moves = (Move * cap)()
        n, tot = ctypes.c_int(0), ctypes.c_int(0)
        before, after = ctypes.c_float(0), ctypes.c_float(0)
        _lib.check(self._lib.spif_balancer_plan(self._h, allv.ctypes.data, int(max_swaps), float(min_gain), moves, cap,
                                                ctypes.byref(n), ctypes.byref(before), ctypes.byref(after), ctypes.byref(tot)),
                   "spif_balancer_plan")
        try:
            moved = self.apply([moves[i] for i in range(n.value)]) if self.world > 1 else 0
        except Exception:
            self._lib.spif_balancer_abort(self._h)
            raise
        _lib.check(self._lib.spif_balancer_commit(self._h), "spif_balancer_commit")
        return {"lambda": self.lam, "imbalance_before": float(before.value), "imbalance_after": float(after.value),
                "swaps": int(tot.value), "bytes_sent_this_rank": int(moved)}
