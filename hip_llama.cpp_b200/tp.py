"""Tensor-parallel split of a llama2 model over the GPUs of one box (SURVEY.md 8e "Optional TP"; north_star:
Llama-2-7B 2/4/8-way over NVLink), over the C-ABI's tha_tp_* entry points (include/tha_b200.h).

Megatron-style shards of the reference's weight tensors (layout of include/models.hpp:25-36):
  wq/wk/wv, w1/w3   split by OUTPUT rows  (heads H/tp, kv heads KV/tp, hidden units)
  wo, w2            split by INPUT columns (so each rank's matmul yields partial sums of the residual update)
  embedding, norms, classifier   replicated
The hidden units are dealt in whole Q8_0 groups (`unit`, 64): 11008 = 172 groups -> tp=8 gives 22,22,22,22,21,21,21,21
groups, so the Q8_0 activation quantiser (runq.c:145-171) sees exactly the groups of the unsplit model.

The all-reduce after wo and w2 is fused into its consumer kernel over peer memory (csrc/tp.h); torch.distributed
(NCCL or gloo) only carries the 64-byte IPC handles at start-up.
"""
from __future__ import annotations

import ctypes as C
import dataclasses

import numpy as np

from . import cabi
from .checkpoint import Config


def split_hidden(hidden: int, tp: int, unit: int = 64) -> list[int]:
    """Local hidden_dim of every rank: whole groups of `unit`, the first (groups % tp) ranks get one more."""
    if hidden % unit:
        raise ValueError(f"hidden_dim {hidden} is not a multiple of {unit}")
    g = hidden // unit
    if g < tp:
        raise ValueError("more ranks than hidden groups")
    return [(g // tp + (1 if r < g % tp else 0)) * unit for r in range(tp)]


def local_config(cfg: Config, tp: int, rank: int, unit: int = 64) -> Config:
    if cfg.n_heads % tp or cfg.n_kv_heads % tp:
        raise ValueError(f"n_heads {cfg.n_heads} / n_kv_heads {cfg.n_kv_heads} not divisible by tp={tp}")
    return dataclasses.replace(cfg, hidden_dim=split_hidden(cfg.hidden_dim, tp, unit)[rank],
                               n_heads=cfg.n_heads // tp, n_kv_heads=cfg.n_kv_heads // tp)


def _ranges(cfg: Config, tp: int, rank: int, unit: int):
    hs = cfg.dim // cfg.n_heads
    da, kvd = hs * cfg.n_heads // tp, hs * cfg.n_kv_heads // tp
    hid = split_hidden(cfg.hidden_dim, tp, unit)
    h0 = sum(hid[:rank])
    return slice(rank * da, (rank + 1) * da), slice(rank * kvd, (rank + 1) * kvd), slice(h0, h0 + hid[rank])


def shard_f32(cfg: Config, w: dict, tp: int, rank: int, unit: int = 64):
    """(local Config, local fp32 weight dict) of rank `rank`."""
    qa, kv, hh = _ranges(cfg, tp, rank, unit)
    out = {k: w[k] for k in ("tok_emb", "rms_att", "rms_ffn", "rms_final")}
    out["wq"] = np.ascontiguousarray(w["wq"][:, qa, :])
    out["wk"] = np.ascontiguousarray(w["wk"][:, kv, :])
    out["wv"] = np.ascontiguousarray(w["wv"][:, kv, :])
    out["wo"] = np.ascontiguousarray(w["wo"][:, :, qa])
    out["w1"] = np.ascontiguousarray(w["w1"][:, hh, :])
    out["w3"] = np.ascontiguousarray(w["w3"][:, hh, :])
    out["w2"] = np.ascontiguousarray(w["w2"][:, :, hh])
    if "wcls" in w:
        out["wcls"] = w["wcls"]
    return local_config(cfg, tp, rank, unit), out


def shard_q8(cfg: Config, gs: int, qt: dict, tp: int, rank: int):
    """Local Q8_0 tensors {name: (int8 [L, out, in], scales [L, out*in/gs])}: rows / whole groups of columns."""
    qa, kv, hh = _ranges(cfg, tp, rank, gs)
    L = cfg.n_layers

    def rows(name, sl):
        q, s = qt[name]
        out_dim, in_dim = q.shape[1], q.shape[2]
        s3 = np.asarray(s).reshape(L, out_dim, in_dim // gs)
        return np.ascontiguousarray(q[:, sl, :]), np.ascontiguousarray(s3[:, sl, :]).reshape(L, -1)

    def cols(name, sl):
        q, s = qt[name]
        out_dim, in_dim = q.shape[1], q.shape[2]
        if sl.start % gs or sl.stop % gs:
            raise ValueError(f"{name}: column split {sl} cuts a Q8_0 group of {gs}")
        s3 = np.asarray(s).reshape(L, out_dim, in_dim // gs)
        return (np.ascontiguousarray(q[:, :, sl]),
                np.ascontiguousarray(s3[:, :, sl.start // gs:sl.stop // gs]).reshape(L, -1))

    out = {"tok_emb": qt["tok_emb"], "wq": rows("wq", qa), "wk": rows("wk", kv), "wv": rows("wv", kv), "wo": cols("wo", qa),
           "w1": rows("w1", hh), "w3": rows("w3", hh), "w2": cols("w2", hh)}
    out["wcls"] = out["tok_emb"] if qt["wcls"] is qt["tok_emb"] else qt["wcls"]
    return local_config(cfg, tp, rank, gs), out


def local_shapes(cfg: Config, tp: int, rank: int, unit: int = 64) -> dict:
    """Shapes of rank `rank`'s tensors (embedding first: it is the replicated one)."""
    lc = local_config(cfg, tp, rank, unit)
    hs = cfg.dim // cfg.n_heads
    da, kvd, L, d, h = hs * lc.n_heads, hs * lc.n_kv_heads, cfg.n_layers, cfg.dim, lc.hidden_dim
    return {"tok_emb": (cfg.vocab_size, d), "wq": (L, da, d), "wk": (L, kvd, d), "wv": (L, kvd, d), "wo": (L, d, da),
            "w1": (L, h, d), "w2": (L, d, h), "w3": (L, h, d)}


def device_random_shard(cfg: Config, tp: int, rank: int, q8: bool, seed: int = 1337, device="cuda", gs: int = 64):
    """Random-init LOCAL weights generated on the device (benchmarks of Llama-2-7B: no checkpoint on the box).  The
    replicated tensors are the same on every rank (same seed), the shards are seeded per rank."""
    from .engine import device_random_weights
    lc = local_config(cfg, tp, rank, gs)
    return device_random_weights(lc, q8=q8, seed=seed, device=device, gs=gs, shapes=local_shapes(cfg, tp, rank, gs),
                                 shard_seed=seed * 1000 + rank)


def xchg_floats(cfg: Config, rows: int) -> int:
    """Capacity of one exchange buffer: the widest split-K factor the planner uses (16) x rows x dim."""
    return 16 * rows * cfg.dim


class TpRank:
    """One rank of a tensor-parallel replica: context, local weights (engine.DeviceWeights / DeviceWeightsQ8 built from
    the shards), local state."""

    def __init__(self, cfg_global: Config, weights_local, tp: int, rank: int, batch: int, device="cuda", unit: int = 64,
                 max_rows: int | None = None):
        import torch
        self.lib = cabi.lib()
        self.cfg_g, self.tp, self.rank, self.B = cfg_global, tp, rank, batch
        self.w = weights_local
        self.q8 = hasattr(weights_local, "gs")
        unit = weights_local.gs if self.q8 else unit
        self.cfg = local_config(cfg_global, tp, rank, unit)
        self.hs = cfg_global.dim // cfg_global.n_heads
        self.da, self.kvd = self.hs * self.cfg.n_heads, self.hs * self.cfg.n_kv_heads
        hid = split_hidden(cfg_global.hidden_dim, tp, unit)
        z = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=device)
        c = self.cfg
        self.x, self.xb, self.q = z(batch, c.dim), z(batch, self.da), z(batch, self.da)
        self.hb = z(batch, c.hidden_dim)
        self.logits = z(batch, c.vocab_size)
        self.key_cache = z(batch, c.n_layers, c.seq_len, self.kvd)
        self.value_cache = z(batch, c.n_layers, c.seq_len, self.kvd)
        p = lambda t: C.c_void_p(t.data_ptr())
        self.state = cabi.State(p(self.x), p(self.xb), None, p(self.hb), None, p(self.q), None, None, None, p(self.logits),
                                p(self.key_cache), p(self.value_cache), None, None, None, None)
        self.ccfg = cabi.Config(c.dim, c.hidden_dim, c.n_layers, c.n_heads, c.n_kv_heads, c.vocab_size, c.seq_len)
        self.ctx = C.c_void_p()
        hid_arr = (C.c_int * tp)(*hid)
        rows = max(batch, max_rows or 0)
        self.lib.check(self.lib.c.tha_tp_create(C.byref(self.ctx), tp, rank, self.hs, hid_arr, xchg_floats(cfg_global, rows)),
                       "tha_tp_create")

    # ---- wiring -------------------------------------------------------------------------------------
    def export_handle(self) -> bytes:
        buf = C.create_string_buffer(64)
        self.lib.check(self.lib.c.tha_tp_export(self.ctx, buf), "tha_tp_export")
        return buf.raw

    def connect_ipc(self, handles: list[bytes]) -> None:
        blob = b"".join(handles)
        assert len(blob) == 64 * self.tp
        self.lib.check(self.lib.c.tha_tp_connect_ipc(self.ctx, blob), "tha_tp_connect_ipc")

    def local_ptr(self) -> int:
        p = C.c_void_p()
        self.lib.check(self.lib.c.tha_tp_local_ptr(self.ctx, C.byref(p)), "tha_tp_local_ptr")
        return p.value

    def connect_ptr(self, bases: list[int]) -> None:
        arr = (C.c_void_p * self.tp)(*bases)
        self.lib.check(self.lib.c.tha_tp_connect_ptr(self.ctx, arr), "tha_tp_connect_ptr")

    def connect_distributed(self, group=None) -> None:
        """Across processes (one per GPU, torchrun): all-gather the IPC handles, map the peers."""
        import torch.distributed as dist
        hs = [None] * self.tp
        dist.all_gather_object(hs, self.export_handle(), group=group)
        self.connect_ipc(hs)

    # ---- the steps ------------------------------------------------------------------------------------
    def set_mode(self, prepare_only: bool) -> None:
        """prepare_only: the next step / prefill call grows scratch, loads kernels and builds its CUDA graph but does
        not take part in an exchange (tha_tp_set_mode) -- for hosts that drive several ranks of one device."""
        self.lib.check(self.lib.c.tha_tp_set_mode(self.ctx, 1 if prepare_only else 0), "tha_tp_set_mode")

    def decode_step(self, token_d, pos_d, logits_d=None, next_d=None, advance=False, stream=None):
        import torch
        s = stream if stream is not None else torch.cuda.current_stream()
        fn = self.lib.c.tha_tp_decode_step_q8 if self.q8 else self.lib.c.tha_tp_decode_step_f32
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        st = fn(C.c_void_p(s.cuda_stream), self.ctx, self.B, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state),
                p(token_d), p(pos_d), p(logits_d), p(next_d), 1 if advance else 0)
        self.lib.check(st, "tha_tp_decode_step")

    def prefill(self, seqs, logits_d=None, next_d=None, stream=None):
        import torch
        s = stream if stream is not None else torch.cuda.current_stream()
        slot = np.ascontiguousarray([q[0] for q in seqs], dtype=np.int32)
        pos0 = np.ascontiguousarray([q[1] for q in seqs], dtype=np.int32)
        ln = np.ascontiguousarray([len(q[2]) for q in seqs], dtype=np.int32)
        toks = np.ascontiguousarray(np.concatenate([np.asarray(q[2], dtype=np.int32) for q in seqs]), dtype=np.int32)
        vp = C.c_void_p
        p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
        fn = self.lib.c.tha_tp_prefill_q8 if self.q8 else self.lib.c.tha_tp_prefill_f32
        st = fn(C.c_void_p(s.cuda_stream), self.ctx, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state), len(seqs),
                slot.ctypes.data_as(vp), pos0.ctypes.data_as(vp), ln.ctypes.data_as(vp), toks.ctypes.data_as(vp),
                p(logits_d), p(next_d))
        self.lib.check(st, "tha_tp_prefill")

    def check_error(self) -> bool:
        return bool(self.lib.c.tha_tp_check_error(self.ctx))

    def last_step_launches(self) -> int:
        return self.lib.c.tha_last_step_launches()

    def weight_bytes_per_step(self) -> int:
        """Bytes of weights this rank streams per decode step (shards once + the replicated classifier)."""
        return self.w.nbytes()

    def close(self):
        if self.ctx:
            self.lib.c.tha_tp_destroy(self.ctx)
            self.ctx = C.c_void_p()


def connect_in_process(ranks: list[TpRank]) -> None:
    """All ranks live in this process (one host thread / stream per GPU, or -- in the single-GPU test -- several ranks
    on one device): hand every rank the others' block pointers."""
    bases = [r.local_ptr() for r in ranks]
    for r in ranks:
        r.connect_ptr(bases)
