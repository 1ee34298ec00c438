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
  scores()     all_gather of the per-slot scores -> global score per group (torch.distributed)
  rebalance()  spif_rebalance_plan per layer (host code in the library, identical on every rank) + the row exchange
The plan only depends on gathered data, so every rank derives the same one without a coordinator.
"""
import numpy as np
import torch
import torch.distributed as dist

from . import placement


def _row_bytes(t, rows):
    n = t.numel() * t.element_size()
    assert n % rows == 0
    return n // rows


class ShardBalancer:
    """keeps slot -> group maps of every layer and shard, the decayed firing rates, and moves groups between shards"""

    LAYER_CHUNK = 16

    def __init__(self, layers, shard_of_group, rank, world, group=None, g=16, lam=0.9, lam_step=0.05, drift_tol=0.02):
        """layers: this rank's LayerWeights (m rows each); shard_of_group: per layer int array [n_g] (same on all ranks).
        Slot s of shard r initially holds the s-th smallest group id placed on r (placement.shard_rows order)."""
        self.layers, self.rank, self.world, self.group = layers, rank, world, group
        self.g, self.lam = g, lam
        # adaptive decay (ggml-sparkinfer.hpp:169-173: lambda *= 1 +- SPIF_DX_DFR_DECAY/1000, clamped to [0.05, 0.95], raised
        # when the compute had to wait for reloads, lowered otherwise).  Re-targeted: raised when the balance the last plan
        # promised did not hold until this round (the scores were too jumpy and the swaps were wasted), lowered otherwise.
        self.lam_step, self.drift_tol = lam_step, drift_tol
        self._promised = None
        self.L = len(layers)
        self.n_slot = layers[0].m // g
        self.slots = []  # per layer: int32 [world, n_slot] = group id held by each slot of each shard
        for sog in shard_of_group:
            sog = np.asarray(sog)
            tab = np.stack([np.nonzero(sog == r)[0] for r in range(world)]).astype(np.int32)
            assert tab.shape == (world, self.n_slot), "equal group count per shard"
            self.slots.append(tab)
        dev = layers[0].up.device
        self.dfr = torch.zeros((self.L, self.n_slot), dtype=torch.float32, device=dev)  # per local slot
        self.tokens_seen = 0
        self.swaps_done = 0
        self.bytes_moved = 0

    # -- bookkeeping --------------------------------------------------------------------------------------------
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
        for l in range(self.L):
            ops.dfr_update(masks[l], self.dfr[l], self.g, self.lam)
        self.tokens_seen += int(masks.shape[1])
        return self.L

    def scores(self):
        """-> per layer float64 [n_g]: the score of every group, assembled from every shard's slots"""
        if self.world > 1:
            parts = [torch.empty_like(self.dfr) for _ in range(self.world)]
            dist.all_gather(parts, self.dfr, group=self.group)
            allv = torch.stack(parts).cpu().numpy()  # [world, L, n_slot]
        else:
            allv = self.dfr.cpu().numpy()[None]
        out = []
        for l in range(self.L):
            d = np.zeros(self.slots[l].size, dtype=np.float64)
            for r in range(self.world):
                d[self.slots[l][r]] = allv[r, l]
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

    def apply(self, plans):
        """plans: per layer list of (group_a, group_b, hot_shard, cold_shard) from placement.rebalance_plan.
        Swap k of every layer forms one round (a group may move again in a later swap of the same layer, so the
        swaps of one layer are ordered; different layers are independent and share the round's send/recv batch)."""
        n_round = max((len(p) for p in plans), default=0)
        moved = 0
        # one send/recv batch = swap k of LAYER_CHUNK consecutive layers (bounds the ops per NCCL group and the staging memory;
        # the chunking is by layer index, so both ends of every pair cut their op lists at the same places)
        work = [(k, c) for k in range(n_round) for c in range(0, self.L, self.LAYER_CHUNK)]
        for k, c0 in work:
            ops_, post = [], []
            for l in range(c0, min(c0 + self.LAYER_CHUNK, self.L)):
                plan = plans[l]
                if k >= len(plan):
                    continue
                a, b, hot, cold = plan[k]
                tab = self.slots[l]
                sa = int(np.nonzero(tab[hot] == a)[0][0])
                sb = int(np.nonzero(tab[cold] == b)[0][0])
                if self.rank in (hot, cold):
                    mine, peer = (sa, cold) if self.rank == hot else (sb, hot)
                    out = self._pack(l, mine)
                    inc = torch.empty_like(out)
                    ops_.append(dist.P2POp(dist.isend, out, peer, group=self.group))
                    ops_.append(dist.P2POp(dist.irecv, inc, peer, group=self.group))
                    post.append((l, mine, inc, out))
                    moved += out.numel()
                tab[hot, sa], tab[cold, sb] = b, a
                self.swaps_done += 1
            if ops_:
                for r in dist.batch_isend_irecv(ops_):
                    r.wait()
                for l, slot, inc, _keep in post:
                    self._unpack(l, slot, inc)
        self.bytes_moved += moved
        return moved

    def rebalance(self, max_swaps=8, min_gain=0.0):
        """gather scores, plan per layer, execute.  -> dict(imbalance_before, imbalance_after (planned), swaps, bytes)"""
        sc = self.scores()
        before = self.imbalance(sc)
        if self.lam_step > 0.0 and self._promised is not None:
            drifted = float(np.mean(before)) > self._promised * (1.0 + self.drift_tol)
            self.lam = float(np.clip(self.lam * (1.0 + (self.lam_step if drifted else -self.lam_step)), 0.05, 0.95))
        from . import ops
        plans = []
        for l in range(self.L):  # planner = host code inside the library (spif_rebalance_plan); placement.py mirrors it
            plan, _ = ops.rebalance_plan(sc[l], self.shard_of_group(l), self.world, max_swaps, min_gain)
            plans.append(plan)
        moved = self.apply(plans) if self.world > 1 else 0
        after = [placement.imbalance(sc[l], self.shard_of_group(l), self.world) for l in range(self.L)]
        self._promised = float(np.mean(after))
        return {"lambda": self.lam,
                "imbalance_before": float(np.mean(before)), "imbalance_after": float(np.mean(after)),
                "swaps": int(sum(len(p) for p in plans)), "bytes_sent_this_rank": int(moved)}
