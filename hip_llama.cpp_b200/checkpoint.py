"""llama2.c checkpoint files: the two on-disk formats the reference reads.

* legacy v0 (fp32): reader `/root/reference/src/utils.cpp:119-170` (== run.cc:110-161),
  writer `/root/reference/train/export.py:75-127`.
* version 2 (Q8_0): reader `/root/reference/runq.c:174-251`, writer
  `/root/reference/train/export.py:182-260`, group quantiser `export.py:45-70`.

The reference's exporter expands every tensor through `struct.pack(*d)`
(`export.py:34-38`), which is unusable at 7B scale; this module writes the
identical byte layout with numpy.  `tests/test_checkpoint.py` checks byte
identity against files produced by the reference exporter (fixtures made by
`tests/golden/make_golden.py`).

Random-init follows `/root/reference/train/model.py:232-247`: N(0, 0.02) for
embeddings and linear layers, N(0, 0.02/sqrt(2*n_layers)) for wo and w3, ones
for the norms, classifier tied to the embedding.
"""
from __future__ import annotations

import math
import struct
from dataclasses import dataclass

import numpy as np

V2_MAGIC = 0x616B3432
V2_HEADER_BYTES = 256


@dataclass
class Config:
    """Field order == `Config` in include/models.hpp:10-18 (it is fread as 7 ints)."""
    dim: int
    hidden_dim: int
    n_layers: int
    n_heads: int
    n_kv_heads: int
    vocab_size: int
    seq_len: int

    @property
    def head_size(self) -> int:
        return self.dim // self.n_heads

    @property
    def kv_dim(self) -> int:
        return self.dim * self.n_kv_heads // self.n_heads

    def pack(self, shared_classifier: bool = True) -> bytes:
        v = self.vocab_size if shared_classifier else -self.vocab_size
        return struct.pack("iiiiiii", self.dim, self.hidden_dim, self.n_layers, self.n_heads,
                           self.n_kv_heads, v, self.seq_len)


def hidden_dim_for(dim: int, multiple_of: int) -> int:
    """train/model.py:170-173."""
    h = int(2 * (4 * dim) / 3)
    return multiple_of * ((h + multiple_of - 1) // multiple_of)


STORIES110M = Config(768, 2048, 12, 12, 12, 32000, 1024)
STORIES15M = Config(288, 768, 6, 6, 6, 32000, 256)
LLAMA2_7B = Config(4096, 11008, 32, 32, 32, 32000, 2048)

# tensor order of the v0 file after the header
V0_ORDER = ("tok_emb", "rms_att", "wq", "wk", "wv", "wo", "rms_ffn", "w1", "w2", "w3", "rms_final")
# quantised tensor order of the v2 file after the fp32 norms
V2_QORDER = ("tok_emb", "wq", "wk", "wv", "wo", "w1", "w2", "w3")


def tensor_shapes(c: Config) -> dict:
    L, d, h, kvd = c.n_layers, c.dim, c.hidden_dim, c.kv_dim
    return {
        "tok_emb": (c.vocab_size, d), "rms_att": (L, d), "rms_ffn": (L, d), "rms_final": (d,),
        "wq": (L, d, d), "wk": (L, kvd, d), "wv": (L, kvd, d), "wo": (L, d, d),
        "w1": (L, h, d), "w2": (L, d, h), "w3": (L, h, d),
    }


def random_weights(c: Config, seed: int = 1337, std: float = 0.02, chunk: int = 1 << 24) -> dict:
    """Seeded random-init fp32 weights of the given architecture (tied classifier)."""
    rng = np.random.default_rng(seed)
    out = {}
    res_std = std / math.sqrt(2 * c.n_layers)
    for name, shape in tensor_shapes(c).items():
        if name.startswith("rms"):
            out[name] = np.ones(shape, dtype=np.float32)
            continue
        s = res_std if name in ("wo", "w3") else std
        n = int(np.prod(shape))
        a = np.empty(n, dtype=np.float32)
        for i in range(0, n, chunk):  # bounded temporaries at 7B scale
            j = min(n, i + chunk)
            a[i:j] = rng.standard_normal(j - i, dtype=np.float32) * np.float32(s)
        out[name] = a.reshape(shape)
    return out


def round_bf16(a: np.ndarray) -> np.ndarray:
    """fp32 -> nearest bf16 (ties to even), returned as fp32 (what tha_convert_bf16 and torch's .to(bfloat16) produce)."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u + np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))) & np.uint32(0xFFFF0000)).astype(np.uint32)
    return r.view(np.float32).reshape(np.shape(a))


def bf16_bits(a: np.ndarray) -> np.ndarray:
    """The 16-bit patterns of round_bf16(a)."""
    return (round_bf16(a).view(np.uint32) >> np.uint32(16)).astype(np.uint16)


def bf16_weights(w: dict) -> dict:
    """The checkpoint as the bf16 model path sees it: matrices and embedding rounded to bf16 (norm gains stay fp32)."""
    return {k: (v if k.startswith("rms") else round_bf16(v)) for k, v in w.items()}


def rope_tables(c: Config):
    """The two dead tables the v0 file still carries (train/model.py:40-47).  Readers skip
    them (src/utils.cpp:145-146); torch is used when present so the bytes equal the
    reference exporter's (numpy's cos/sin differ from torch's in the last ulp)."""
    hs = c.head_size
    try:
        import torch
        freqs = 1.0 / (10000.0 ** (torch.arange(0, hs, 2)[: (hs // 2)].float() / hs))
        ang = torch.outer(torch.arange(c.seq_len), freqs).float()
        return torch.cos(ang).numpy(), torch.sin(ang).numpy()
    except ImportError:  # pragma: no cover
        freqs = (1.0 / (10000.0 ** (np.arange(0, hs, 2, dtype=np.float32)[: hs // 2] / np.float32(hs)))).astype(np.float32)
        ang = np.outer(np.arange(c.seq_len, dtype=np.float32), freqs).astype(np.float32)
        return np.cos(ang).astype(np.float32), np.sin(ang).astype(np.float32)


def write_v0(path: str, c: Config, w: dict, wcls: np.ndarray | None = None) -> None:
    with open(path, "wb") as f:
        f.write(c.pack(shared_classifier=wcls is None))
        for name in V0_ORDER:
            np.ascontiguousarray(w[name], dtype=np.float32).tofile(f)
        cos, sin = rope_tables(c)
        cos.tofile(f)
        sin.tofile(f)
        if wcls is not None:
            np.ascontiguousarray(wcls, dtype=np.float32).tofile(f)


def quantize_q80(w: np.ndarray, group_size: int):
    """export.py:45-70: symmetric int8 per group, round-half-to-even (torch.round)."""
    flat = np.ascontiguousarray(w, dtype=np.float32).reshape(-1, group_size)
    wmax = np.abs(flat).max(axis=1)
    scale = (wmax / np.float32(127.0)).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        quant = flat / scale[:, None]
    q = np.rint(quant)
    q = np.where(np.isnan(q), 0.0, q).astype(np.int8)
    return q.reshape(-1), scale


def v2_group_size(c: Config, group_size: int = 64) -> int:
    """export.py:193-195 back-off rule."""
    while c.dim % group_size != 0:
        group_size //= 2
    return group_size


def write_v2(path: str, c: Config, w: dict, group_size: int = 64, wcls: np.ndarray | None = None) -> int:
    gs = v2_group_size(c, group_size)
    with open(path, "wb") as f:
        f.write(struct.pack("I", V2_MAGIC))
        f.write(struct.pack("i", 2))
        f.write(Config(c.dim, c.hidden_dim, c.n_layers, c.n_heads, c.n_kv_heads, c.vocab_size, c.seq_len).pack(True))
        f.write(struct.pack("B", int(wcls is None)))
        f.write(struct.pack("i", gs))
        f.write(b"\0" * (V2_HEADER_BYTES - f.tell()))
        for name in ("rms_att", "rms_ffn", "rms_final"):
            np.ascontiguousarray(w[name], dtype=np.float32).tofile(f)
        tensors = []
        for name in V2_QORDER:
            t = w[name]
            tensors.extend([t] if name == "tok_emb" else [t[l] for l in range(c.n_layers)])
        if wcls is not None:
            tensors.append(wcls)
        for t in tensors:
            q, s = quantize_q80(t, gs)
            q.tofile(f)
            s.tofile(f)
    return gs


def read_v0(path: str):
    """Returns (Config, dict of numpy views on a memmap, shared_classifier)."""
    raw = np.memmap(path, dtype=np.uint8, mode="r")
    vals = struct.unpack("iiiiiii", raw[:28].tobytes())
    shared = vals[5] > 0
    c = Config(vals[0], vals[1], vals[2], vals[3], vals[4], abs(vals[5]), vals[6])
    fl = np.memmap(path, dtype=np.float32, mode="r", offset=28)
    shapes = tensor_shapes(c)
    off = 0
    w = {}
    for name in V0_ORDER:
        n = int(np.prod(shapes[name]))
        w[name] = fl[off:off + n].reshape(shapes[name])
        off += n
    off += 2 * (c.seq_len * c.head_size // 2)
    w["wcls"] = w["tok_emb"] if shared else fl[off:off + c.vocab_size * c.dim].reshape(c.vocab_size, c.dim)
    return c, w, shared


def read_v2(path: str):
    """Returns (Config, group_size, fp32 norms dict, {name: (int8 array, fp32 scales)}, shared)."""
    raw = np.memmap(path, dtype=np.uint8, mode="r")
    magic, version = struct.unpack("Ii", raw[:8].tobytes())
    if magic != V2_MAGIC or version != 2:
        raise ValueError(f"{path}: not a version-2 llama2.c checkpoint")
    vals = struct.unpack("iiiiiii", raw[8:36].tobytes())
    shared = bool(raw[36])
    gs = struct.unpack("i", raw[37:41].tobytes())[0]
    c = Config(*vals)
    shapes = tensor_shapes(c)
    off = V2_HEADER_BYTES
    norms = {}
    for name in ("rms_att", "rms_ffn", "rms_final"):
        n = int(np.prod(shapes[name]))
        norms[name] = np.frombuffer(raw, dtype=np.float32, count=n, offset=off).reshape(shapes[name])
        off += 4 * n
    qt = {}
    names = list(V2_QORDER) + ([] if shared else ["wcls"])
    for name in names:
        shape = shapes["tok_emb"] if name == "wcls" else shapes[name]
        if name in ("tok_emb", "wcls"):
            n = int(np.prod(shape))
            q = np.frombuffer(raw, dtype=np.int8, count=n, offset=off); off += n
            s = np.frombuffer(raw, dtype=np.float32, count=n // gs, offset=off); off += 4 * (n // gs)
            qt[name] = (q.reshape(shape), s)
        else:
            per = int(np.prod(shape[1:]))
            qs, ss = [], []
            for _ in range(c.n_layers):
                qs.append(np.frombuffer(raw, dtype=np.int8, count=per, offset=off)); off += per
                ss.append(np.frombuffer(raw, dtype=np.float32, count=per // gs, offset=off)); off += 4 * (per // gs)
            qt[name] = (np.stack(qs).reshape(shape), np.stack(ss))
    if shared:
        qt["wcls"] = qt["tok_emb"]
    return c, gs, norms, qt, shared
