"""Python mirror of the reference's model/state helpers, over the C-ABI.

  DeviceWeights      ~ copy_weight_to_device            (src/models.cpp:86-125)
  DeviceWeightsQ8    ~ the v2 loader of runq.c:174-251, uploaded
  BatchState         ~ alloc_state_to_device_batch      (src/models.cpp:155-179)
  forward_batch      ~ thaDNN_s_forward_batch           (src/thaDNN.cpp:13-81)

torch is used for device memory, streams and pinned buffers only; every kernel that runs
is one of ours, reached through hip_llama.cpp_b200/cabi.py.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import cabi
from .checkpoint import Config


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def c_config(cfg: Config) -> cabi.Config:
    return cabi.Config(cfg.dim, cfg.hidden_dim, cfg.n_layers, cfg.n_heads, cfg.n_kv_heads, cfg.vocab_size, cfg.seq_len)


class DeviceWeights:
    """fp32 weights resident in HBM, layer-major contiguous exactly as in the v0 file."""
    NAMES = ("tok_emb", "rms_att", "rms_ffn", "wq", "wk", "wv", "wo", "w1", "w2", "w3", "rms_final")

    def __init__(self, cfg: Config, w: dict, device="cuda", shared_classifier=True):
        self.cfg = cfg
        self.t = {}
        for n in self.NAMES:
            self.t[n] = torch.from_numpy(np.ascontiguousarray(w[n], dtype=np.float32)).to(device)
        if shared_classifier or "wcls" not in w or w["wcls"] is w["tok_emb"]:
            self.t["wcls"] = self.t["tok_emb"]
        else:
            self.t["wcls"] = torch.from_numpy(np.ascontiguousarray(w["wcls"], dtype=np.float32)).to(device)
        t = self.t
        self.c = cabi.WeightsF32(_ptr(t["tok_emb"]), _ptr(t["rms_att"]), _ptr(t["rms_ffn"]), _ptr(t["wq"]),
                                 _ptr(t["wk"]), _ptr(t["wv"]), _ptr(t["wo"]), _ptr(t["w1"]), _ptr(t["w2"]),
                                 _ptr(t["w3"]), _ptr(t["rms_final"]), _ptr(t["wcls"]))

    def nbytes(self) -> int:
        seen, n = set(), 0
        for t in self.t.values():
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                n += t.numel() * 4
        return n


class DeviceWeightsQ8:
    """Q8_0 weights (runq.c:34-37): int8 [L, out*in] + fp32 scales [L, out*in/GS]; fp32 norms;
    dequantised fp32 embedding table (runq.c:205-206)."""

    def __init__(self, cfg: Config, gs: int, norms: dict, qt: dict, device="cuda"):
        self.cfg, self.gs = cfg, gs
        self.t = {}
        for n in ("rms_att", "rms_ffn", "rms_final"):
            self.t[n] = torch.from_numpy(np.ascontiguousarray(norms[n], dtype=np.float32)).to(device)
        for n in ("wq", "wk", "wv", "wo", "w1", "w2", "w3", "tok_emb"):
            q, s = qt[n]
            self.t[n + ".q"] = torch.from_numpy(np.ascontiguousarray(q, dtype=np.int8)).to(device)
            self.t[n + ".s"] = torch.from_numpy(np.ascontiguousarray(s, dtype=np.float32)).to(device)
        if qt["wcls"] is qt["tok_emb"]:
            self.t["wcls.q"], self.t["wcls.s"] = self.t["tok_emb.q"], self.t["tok_emb.s"]
        else:
            q, s = qt["wcls"]
            self.t["wcls.q"] = torch.from_numpy(np.ascontiguousarray(q, dtype=np.int8)).to(device)
            self.t["wcls.s"] = torch.from_numpy(np.ascontiguousarray(s, dtype=np.float32)).to(device)
        # dequantise the embedding table: x = q * s[group]  (int8 -> fp32 product is exact)
        q = self.t["tok_emb.q"].reshape(-1, gs).float()
        self.t["tok_emb"] = (q * self.t["tok_emb.s"].reshape(-1, 1)).reshape(cfg.vocab_size, cfg.dim).contiguous()
        t = self.t

        def qten(n):
            return cabi.QTensor(_ptr(t[n + ".q"]), _ptr(t[n + ".s"]))

        self.c = cabi.WeightsQ8(_ptr(t["tok_emb"]), _ptr(t["rms_att"]), _ptr(t["rms_ffn"]), _ptr(t["rms_final"]),
                                qten("wq"), qten("wk"), qten("wv"), qten("wo"), qten("w1"), qten("w2"), qten("w3"),
                                qten("wcls"), gs)

    def nbytes(self) -> int:
        """Bytes one decode step must read: every quantised tensor once (classifier == embedding)."""
        n = 0
        for k, t in self.t.items():
            if k == "tok_emb" or k.startswith("wcls"):
                continue
            n += t.numel() * t.element_size()
        return n


class DeviceWeightsBF16:
    """bf16 weights: the fp32 checkpoint rounded to nearest-even bf16 on upload (matrices and embedding table; norm
    gains stay fp32).  Layer-major contiguous like the fp32 tensors."""
    MATS = ("tok_emb", "wq", "wk", "wv", "wo", "w1", "w2", "w3")

    def __init__(self, cfg: Config, w: dict, device="cuda", shared_classifier=True):
        self.cfg = cfg
        self.t = {}
        for n in ("rms_att", "rms_ffn", "rms_final"):
            self.t[n] = torch.from_numpy(np.ascontiguousarray(w[n], dtype=np.float32)).to(device)
        for n in self.MATS:
            self.t[n] = torch.from_numpy(np.ascontiguousarray(w[n], dtype=np.float32)).to(device).to(torch.bfloat16)
        if shared_classifier or "wcls" not in w or w["wcls"] is w["tok_emb"]:
            self.t["wcls"] = self.t["tok_emb"]
        else:
            self.t["wcls"] = torch.from_numpy(np.ascontiguousarray(w["wcls"], dtype=np.float32)).to(device).to(torch.bfloat16)
        self._bind()

    def _bind(self):
        t = self.t
        self.c = cabi.WeightsBF16(_ptr(t["tok_emb"]), _ptr(t["rms_att"]), _ptr(t["rms_ffn"]), _ptr(t["rms_final"]), _ptr(t["wq"]),
                                  _ptr(t["wk"]), _ptr(t["wv"]), _ptr(t["wo"]), _ptr(t["w1"]), _ptr(t["w2"]), _ptr(t["w3"]),
                                  _ptr(t["wcls"]))

    @classmethod
    def random(cls, cfg: Config, seed: int = 1337, device="cuda", std: float = 0.02):
        """Random-init N(0, std) generated on the device (Llama-2-7B: 13.5 GB), rounded to bf16."""
        g = torch.Generator(device=device)
        g.manual_seed(seed)
        L, d, h, kvd, V = cfg.n_layers, cfg.dim, cfg.hidden_dim, cfg.kv_dim, cfg.vocab_size
        shapes = {"tok_emb": (V, d), "wq": (L, d, d), "wk": (L, kvd, d), "wv": (L, kvd, d), "wo": (L, d, d),
                  "w1": (L, h, d), "w2": (L, d, h), "w3": (L, h, d)}
        w = cls.__new__(cls)
        w.cfg = cfg
        w.t = {"rms_att": torch.ones(L, d, device=device), "rms_ffn": torch.ones(L, d, device=device),
               "rms_final": torch.ones(d, device=device)}
        for n, sh in shapes.items():
            t = torch.empty(sh, dtype=torch.bfloat16, device=device)
            lead = sh[0] if len(sh) == 3 else 1
            tv = t.view(lead, -1)
            for l in range(lead):   # one layer at a time: bounded fp32 scratch
                x = torch.empty(tv.shape[1], dtype=torch.float32, device=device)
                x.normal_(0.0, std, generator=g)
                tv[l] = x.to(torch.bfloat16)
                del x
            w.t[n] = t
        w.t["wcls"] = w.t["tok_emb"]
        w._bind()
        return w

    def nbytes(self) -> int:
        seen, n = set(), 0
        for t in self.t.values():
            if t.data_ptr() not in seen:
                seen.add(t.data_ptr())
                n += t.numel() * t.element_size()
        return n


def device_random_weights(cfg: Config, q8: bool = False, seed: int = 1337, device="cuda", std: float = 0.02, gs: int = 64,
                          shapes: dict | None = None, shard_seed: int | None = None):
    """Random-init weights of the named architecture generated ON the device (benchmarks of models whose
    host-side generation would take minutes: Llama-2-7B is 26 GB in fp32).  Same N(0, std) law as
    checkpoint.random_weights; Q8_0 tensors are quantised on the device with the exporter's rule
    (train/export.py:57-63: scale = max|w| / 127 per group, round half to even)."""
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    L, d, h, kvd, V = cfg.n_layers, cfg.dim, cfg.hidden_dim, cfg.kv_dim, cfg.vocab_size
    if shapes is None:   # tensor-parallel shards pass their local shapes (tp.py); "tok_emb" stays first: it is replicated
        shapes = {"tok_emb": (V, d), "wq": (L, d, d), "wk": (L, kvd, d), "wv": (L, kvd, d), "wo": (L, d, d),
                  "w1": (L, h, d), "w2": (L, d, h), "w3": (L, h, d)}
    norms = {"rms_att": torch.ones(L, d, device=device), "rms_ffn": torch.ones(L, d, device=device),
             "rms_final": torch.ones(d, device=device)}

    def gen(shape):
        t = torch.empty(shape, dtype=torch.float32, device=device)
        t.normal_(0.0, std, generator=g)
        return t

    def reseed(n):   # everything after the replicated embedding table differs per shard
        if n == "tok_emb" and shard_seed is not None:
            g.manual_seed(shard_seed)

    if not q8:
        w = DeviceWeights.__new__(DeviceWeights)
        w.cfg = cfg
        w.t = {}
        for n, sh in shapes.items():
            w.t[n] = gen(sh)
            reseed(n)
        w.t.update(norms)
        w.t["wcls"] = w.t["tok_emb"]
        t = w.t
        w.c = cabi.WeightsF32(_ptr(t["tok_emb"]), _ptr(t["rms_att"]), _ptr(t["rms_ffn"]), _ptr(t["wq"]), _ptr(t["wk"]),
                              _ptr(t["wv"]), _ptr(t["wo"]), _ptr(t["w1"]), _ptr(t["w2"]), _ptr(t["w3"]),
                              _ptr(t["rms_final"]), _ptr(t["wcls"]))
        return w
    w = DeviceWeightsQ8.__new__(DeviceWeightsQ8)
    w.cfg, w.gs = cfg, gs
    w.t = dict(norms)
    for n, sh in shapes.items():
        lead = sh[0] if len(sh) == 3 else 1
        per = int(np.prod(sh)) // lead
        q = torch.empty(lead, per, dtype=torch.int8, device=device)
        sc = torch.empty(lead, per // gs, dtype=torch.float32, device=device)
        for l in range(lead):   # one layer at a time: bounded fp32 scratch
            x = gen((per // gs, gs))
            m = x.abs().amax(dim=1)
            scale = m / 127.0
            sc[l] = scale
            q[l] = torch.round(x / scale[:, None]).to(torch.int8).reshape(-1)
            del x
        w.t[n + ".q"], w.t[n + ".s"] = q.reshape(sh), sc
        reseed(n)
    w.t["wcls.q"], w.t["wcls.s"] = w.t["tok_emb.q"], w.t["tok_emb.s"]
    w.t["tok_emb"] = (w.t["tok_emb.q"].reshape(-1, gs).float() * w.t["tok_emb.s"].reshape(-1, 1)).reshape(V, d).contiguous()
    t = w.t

    def qten(n):
        return cabi.QTensor(_ptr(t[n + ".q"]), _ptr(t[n + ".s"]))

    w.c = cabi.WeightsQ8(_ptr(t["tok_emb"]), _ptr(t["rms_att"]), _ptr(t["rms_ffn"]), _ptr(t["rms_final"]), qten("wq"),
                         qten("wk"), qten("wv"), qten("wo"), qten("w1"), qten("w2"), qten("w3"), qten("wcls"), gs)
    return w


class PagedKV:
    """A block pool for the KV cache of `batch` slots (what BatchManager::create_pool builds in the C++ layer): K and V pools
    [n_blocks][L][block_size][kv_dim] and a device table [batch][blocks per slot] of block ids, registered with the library
    (tha_kv_pool_register).  The caller decides which block holds which positions of which slot (set_table)."""

    def __init__(self, cfg: Config, batch: int, n_blocks: int, block_size: int = 16, device="cuda"):
        self.cfg, self.B, self.n_blocks, self.block_size = cfg, batch, n_blocks, block_size
        self.nrow = (cfg.seq_len + block_size - 1) // block_size
        self.k = torch.zeros(n_blocks, cfg.n_layers, block_size, cfg.kv_dim, dtype=torch.float32, device=device)
        self.v = torch.zeros_like(self.k)
        self.table = torch.zeros(batch, self.nrow, dtype=torch.int32, device=device)
        self.lib = cabi.lib()
        self.lib.check(self.lib.c.tha_kv_pool_register(_ptr(self.k), _ptr(self.v), _ptr(self.table), self.nrow, block_size, n_blocks),
                       "tha_kv_pool_register")

    def set_table(self, rows):
        self.table.copy_(torch.from_numpy(np.ascontiguousarray(rows, dtype=np.int32)).to(self.table.device))

    def rows(self, which: str, slot: int, n: int) -> np.ndarray:
        """Cache rows 0..n-1 of a slot, every layer, gathered through the table: [L, n, kv_dim]."""
        pool = (self.k if which == "key_cache" else self.v).cpu().numpy()
        tab = self.table[slot].cpu().numpy()
        out = np.empty((self.cfg.n_layers, n, self.cfg.kv_dim), np.float32)
        for t in range(n):
            out[:, t] = pool[tab[t // self.block_size], :, t % self.block_size]
        return out

    def close(self):
        if self.k is not None:
            self.lib.c.tha_kv_pool_unregister(_ptr(self.k))
            self.k = self.v = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BatchState:
    """Batched RunState: every buffer x B (src/models.cpp:155-179); KV [B][L][S][kv_dim] fp32, or a PagedKV pool."""

    def __init__(self, cfg: Config, batch: int, device="cuda", with_att=True, paged: "PagedKV | None" = None):
        self.cfg, self.B = cfg, batch
        z = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=device)
        self.x, self.xb, self.xb2, self.q = z(batch, cfg.dim), z(batch, cfg.dim), z(batch, cfg.dim), z(batch, cfg.dim)
        self.hb, self.hb2 = z(batch, cfg.hidden_dim), z(batch, cfg.hidden_dim)
        self.att = z(batch, cfg.n_heads, cfg.seq_len) if with_att else None
        self.logits = z(batch, cfg.vocab_size)
        self.paged = paged
        if paged is not None:
            self.key_cache, self.value_cache = paged.k, paged.v
        else:
            self.key_cache = z(batch, cfg.n_layers, cfg.seq_len, cfg.kv_dim)
            self.value_cache = z(batch, cfg.n_layers, cfg.seq_len, cfg.kv_dim)
        self.c = cabi.State(_ptr(self.x), _ptr(self.xb), _ptr(self.xb2), _ptr(self.hb), _ptr(self.hb2), _ptr(self.q),
                            None, None, _ptr(self.att), _ptr(self.logits), _ptr(self.key_cache),
                            _ptr(self.value_cache), None, None, None, None)


class Engine:
    """One model replica on one GPU, B slots."""

    def __init__(self, weights, batch: int, device="cuda", with_att=True, paged: "PagedKV | None" = None):
        self.lib = cabi.lib()
        self.w = weights
        self.cfg = weights.cfg
        self.q8 = isinstance(weights, DeviceWeightsQ8)
        self.bf16 = isinstance(weights, DeviceWeightsBF16)
        self.B = batch
        self.state = BatchState(self.cfg, batch, device, with_att, paged)
        self.ccfg = c_config(self.cfg)
        self.device = device

    def _fn(self, stem):
        return getattr(self.lib.c, stem + ("_q8" if self.q8 else ("_bf16" if self.bf16 else "_f32")))

    def forward_batch(self, tokens, pos, logits_host=None, stream=None):
        """The reference call: HOST token/pos in, [B, vocab] logits out in pinned host memory."""
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        ps = np.ascontiguousarray(pos, dtype=np.int32)
        if logits_host is None:
            logits_host = torch.empty(self.B, self.cfg.vocab_size, dtype=torch.float32).pin_memory()
        fn = self._fn("tha_forward_batch")
        st = fn(_stream_ptr(stream), self.B, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state.c),
                tok.ctypes.data_as(C.c_void_p), ps.ctypes.data_as(C.c_void_p), C.c_void_p(logits_host.data_ptr()))
        self.lib.check(st, "tha_forward_batch")
        return logits_host

    def forward_greedy(self, tokens, pos, next_host=None, stream=None):
        """thaDNN_s_forward_batch + sample_argmax (src/llama.cpp:275-286) in one call: HOST token/pos in, HOST next
        token per slot out (argmax on the device; B ints cross PCIe instead of B*vocab floats)."""
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        ps = np.ascontiguousarray(pos, dtype=np.int32)
        if next_host is None:
            next_host = np.zeros(self.B, np.int32)
        fn = self._fn("tha_forward_greedy")
        st = fn(_stream_ptr(stream), self.B, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state.c),
                tok.ctypes.data_as(C.c_void_p), ps.ctypes.data_as(C.c_void_p), next_host.ctypes.data_as(C.c_void_p))
        self.lib.check(st, "tha_forward_greedy")
        return next_host

    def forward_sample(self, tokens, pos, force=None, reseed=None, temperature=1.0, topp=0.9, next_host=None, stream=None):
        """thaDNN_s_forward_batch + sample() per slot (src/llama.cpp:1025-1050) in one call: HOST token/pos, forced prompt
        token per slot (-1 = sample) and sampler reseed per slot (0 = keep) in, HOST next token out; softmax / top-p / coin on
        the device."""
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        ps = np.ascontiguousarray(pos, dtype=np.int32)
        fc = None if force is None else np.ascontiguousarray(force, dtype=np.int32)
        rs = None if reseed is None else np.ascontiguousarray(reseed, dtype=np.uint64)
        if next_host is None:
            next_host = np.zeros(self.B, np.int32)
        fn = self._fn("tha_forward_sample")
        vp = C.c_void_p
        st = fn(_stream_ptr(stream), self.B, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state.c),
                tok.ctypes.data_as(vp), ps.ctypes.data_as(vp), fc.ctypes.data_as(vp) if fc is not None else None,
                rs.ctypes.data_as(vp) if rs is not None else None, C.c_float(temperature), C.c_float(topp),
                next_host.ctypes.data_as(vp))
        self.lib.check(st, "tha_forward_sample")
        return next_host

    def decode_step(self, token_d, pos_d, logits_d=None, next_d=None, stream=None):
        """Resident step: device token/pos, optional device logits and greedy next tokens."""
        fn = self._fn("tha_decode_step")
        st = fn(_stream_ptr(stream), self.B, C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state.c),
                _ptr(token_d), _ptr(pos_d), _ptr(logits_d), _ptr(next_d))
        self.lib.check(st, "tha_decode_step")

    def decode_greedy(self, token_d, pos_d, next_d, logits_d=None, stream=None):
        """Resident greedy step: the step, the argmax and the slot bookkeeping (token = next,
        pos += 1) without leaving the device -- one launch where the persistent kernel applies."""
        if self.q8 or self.bf16:
            self.decode_step(token_d, pos_d, logits_d, next_d, stream)
            self.advance(token_d, pos_d, next_d, stream)
            return
        st = self.lib.c.tha_decode_greedy_f32(_stream_ptr(stream), self.B, C.byref(self.ccfg), C.byref(self.w.c),
                                              C.byref(self.state.c), _ptr(token_d), _ptr(pos_d), _ptr(logits_d),
                                              _ptr(next_d))
        self.lib.check(st, "tha_decode_greedy")

    def prefill(self, seqs, logits_d=None, next_d=None, stream=None):
        """Batched prompt ingestion: seqs = [(slot, pos0, tokens), ...]; every position's K/V goes to the cache in
        one pass of tensor-core matmuls; logits_d [n_seqs, vocab] / next_d [n_seqs] get the last position of each."""
        slot = np.ascontiguousarray([s[0] for s in seqs], dtype=np.int32)
        pos0 = np.ascontiguousarray([s[1] for s in seqs], dtype=np.int32)
        ln = np.ascontiguousarray([len(s[2]) for s in seqs], dtype=np.int32)
        toks = np.ascontiguousarray(np.concatenate([np.asarray(s[2], dtype=np.int32) for s in seqs]), dtype=np.int32)
        if slot.size and (slot.min() < 0 or slot.max() >= self.B):
            raise ValueError("slot out of range")
        vp = C.c_void_p
        fn = self._fn("tha_prefill")
        st = fn(_stream_ptr(stream), C.byref(self.ccfg), C.byref(self.w.c), C.byref(self.state.c),
                                        len(seqs), slot.ctypes.data_as(vp), pos0.ctypes.data_as(vp), ln.ctypes.data_as(vp),
                                        toks.ctypes.data_as(vp), _ptr(logits_d), _ptr(next_d))
        self.lib.check(st, "tha_prefill")

    def advance(self, token_d, pos_d, next_d, stream=None):
        """Device-side greedy bookkeeping: token = next, pos += 1 (src/llama.cpp:1027-1050)."""
        st = self.lib.c.tha_advance_slots(_stream_ptr(stream), self.B, _ptr(token_d), _ptr(pos_d), _ptr(next_d),
                                          self.cfg.seq_len)
        self.lib.check(st, "tha_advance_slots")

    def last_step_launches(self) -> int:
        return self.lib.c.tha_last_step_launches()
