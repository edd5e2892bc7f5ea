"""ctypes bindings for the CHECKERS: oracle/liboracle.so (the CPU restatement) and,
when present, oracle/_ref/*.so (the reference sources compiled unmodified).

Test infrastructure only; nothing under hip_llama.cpp_b200/ may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

f32p = C.POINTER(C.c_float)
i8p = C.POINTER(C.c_int8)
i32p = C.POINTER(C.c_int)


def _fp(a):
    return a.ctypes.data_as(f32p)


def _ip(a):
    return a.ctypes.data_as(i32p)


def _bp(a):
    return a.ctypes.data_as(i8p)


def build_oracle():
    so = os.path.join(ORACLE_DIR, "liboracle.so")
    src = os.path.join(ORACLE_DIR, "llama2_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])
    return so


class Oracle:
    """The plain-C restatement (oracle/llama2_oracle.c)."""

    def __init__(self):
        lib = C.CDLL(build_oracle())
        self.lib = lib
        lib.orc_open_v0.restype = C.c_void_p
        lib.orc_open_v0.argtypes = [C.c_char_p]
        lib.orc_open_v2.restype = C.c_void_p
        lib.orc_open_v2.argtypes = [C.c_char_p]
        lib.orc_close.argtypes = [C.c_void_p]
        lib.orc_set_bf16_activations.argtypes = [C.c_void_p, C.c_int]
        lib.orc_set_bf16_attention.argtypes = [C.c_void_p, C.c_int]
        lib.orc_get_config.argtypes = [C.c_void_p, i32p]
        lib.orc_seq_new.restype = C.c_void_p
        lib.orc_seq_new.argtypes = [C.c_void_p]
        lib.orc_seq_free.argtypes = [C.c_void_p]
        lib.orc_seq_ptr.restype = f32p
        lib.orc_seq_ptr.argtypes = [C.c_void_p, C.c_int]
        lib.orc_seq_q.restype = i8p
        lib.orc_seq_q.argtypes = [C.c_void_p, C.c_int]
        lib.orc_seq_s.restype = f32p
        lib.orc_seq_s.argtypes = [C.c_void_p, C.c_int]
        lib.orc_forward.restype = f32p
        lib.orc_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        lib.orc_forward_many.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), i32p, i32p, C.c_int, C.c_int]
        lib.orc_rmsnorm.argtypes = [f32p, f32p, f32p, C.c_int]
        lib.orc_softmax.argtypes = [f32p, C.c_int]
        lib.orc_matmul_f32.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int]
        lib.orc_rope.argtypes = [f32p, f32p, C.c_int, C.c_int, C.c_int, C.c_int]
        lib.orc_attention.argtypes = [f32p, f32p, f32p, f32p, f32p] + [C.c_int] * 6
        lib.orc_swiglu.argtypes = [f32p, f32p, C.c_int]
        lib.orc_argmax.restype = C.c_int
        lib.orc_argmax.argtypes = [f32p, C.c_int]
        lib.orc_quantize.argtypes = [i8p, f32p, f32p, C.c_int, C.c_int]
        lib.orc_dequantize.argtypes = [f32p, i8p, f32p, C.c_int, C.c_int]
        lib.orc_matmul_q8.argtypes = [f32p, i8p, f32p, i8p, f32p, C.c_int, C.c_int, C.c_int]

    # ---- single ops, numpy in / numpy out ----
    def rmsnorm(self, x, w):
        x = np.ascontiguousarray(x, np.float32); w = np.ascontiguousarray(w, np.float32)
        o = np.empty_like(x)
        self.lib.orc_rmsnorm(_fp(o), _fp(x), _fp(w), x.size)
        return o

    def softmax(self, x):
        y = np.array(x, np.float32, copy=True)
        self.lib.orc_softmax(_fp(y), y.size)
        return y

    def matmul(self, w, x):
        """w [d, n], x [n] -> [d] (run.cc:231-242)."""
        w = np.ascontiguousarray(w, np.float32); x = np.ascontiguousarray(x, np.float32)
        d, n = w.shape
        o = np.empty(d, np.float32)
        self.lib.orc_matmul_f32(_fp(o), _fp(x), _fp(w), n, d)
        return o

    def rope(self, q, k, head_size, pos):
        q = np.array(q, np.float32, copy=True); k = np.array(k, np.float32, copy=True)
        self.lib.orc_rope(_fp(q), _fp(k), q.size, k.size, head_size, pos)
        return q, k

    def attention(self, q, kc, vc, pos, n_heads, kv_mul, head_size):
        """kc, vc: [seq_len, kv_dim] slabs of one layer.  Returns (out[dim], att[n_heads, seq_len])."""
        q = np.ascontiguousarray(q, np.float32)
        kc = np.ascontiguousarray(kc, np.float32); vc = np.ascontiguousarray(vc, np.float32)
        seq_len, kv_dim = kc.shape
        out = np.zeros(n_heads * head_size, np.float32)
        att = np.zeros((n_heads, seq_len), np.float32)
        self.lib.orc_attention(_fp(out), _fp(att), _fp(q), _fp(kc), _fp(vc), pos, n_heads, kv_mul,
                               head_size, kv_dim, seq_len)
        return out, att

    def swiglu(self, hb, hb2):
        hb = np.array(hb, np.float32, copy=True); hb2 = np.ascontiguousarray(hb2, np.float32)
        self.lib.orc_swiglu(_fp(hb), _fp(hb2), hb.size)
        return hb

    def argmax(self, p):
        p = np.ascontiguousarray(p, np.float32)
        return self.lib.orc_argmax(_fp(p), p.size)

    def quantize(self, x, gs):
        x = np.ascontiguousarray(x, np.float32)
        q = np.zeros(x.size, np.int8); s = np.zeros(x.size // gs, np.float32)
        self.lib.orc_quantize(_bp(q), _fp(s), _fp(x), x.size, gs)
        return q, s

    def matmul_q8(self, wq, ws, xq, xs, gs):
        wq = np.ascontiguousarray(wq, np.int8); ws = np.ascontiguousarray(ws, np.float32)
        xq = np.ascontiguousarray(xq, np.int8); xs = np.ascontiguousarray(xs, np.float32)
        d, n = wq.shape
        o = np.empty(d, np.float32)
        self.lib.orc_matmul_q8(_fp(o), _bp(xq), _fp(xs), _bp(wq), _fp(ws), n, d, gs)
        return o

    # ---- models ----
    def open(self, path, quantized=False, bf16=False):
        """bf16 = True: a v0 file whose weights were rounded to bf16 (checkpoint.round_bf16); matmul inputs are then rounded
        to bf16 as the tensor cores consume them (oracle/llama2_oracle.c, orc_set_bf16_activations)."""
        m = OracleModel(self, path, quantized)
        if bf16:
            self.lib.orc_set_bf16_activations(m.h, 1)
        return m


class OracleModel:
    def __init__(self, orc, path, quantized):
        self.orc, self.lib = orc, orc.lib
        self.h = (self.lib.orc_open_v2 if quantized else self.lib.orc_open_v0)(path.encode())
        if not self.h:
            raise IOError(f"oracle could not open {path}")
        cfg = (C.c_int * 9)()
        self.lib.orc_get_config(self.h, cfg)
        (self.dim, self.hidden_dim, self.n_layers, self.n_heads, self.n_kv_heads, self.vocab_size,
         self.seq_len, self.group_size, self.shared_classifier) = list(cfg)
        self.kv_dim = self.dim * self.n_kv_heads // self.n_heads

    def new_seq(self):
        return OracleSeq(self)

    def set_bf16_attention(self, on):
        """Prompt positions of the bf16 path with head size 64 / 128: the tensor-core attention rule (q, k, v, p rounded to
        bf16, keys in tiles of 128); decode positions keep run.cc's fp32 attention."""
        self.lib.orc_set_bf16_attention(self.h, 1 if on else 0)

    def forward_many(self, seqs, tokens, pos, n_threads=1):
        n = len(seqs)
        arr = (C.c_void_p * n)(*[s.h for s in seqs])
        t = np.ascontiguousarray(tokens, np.int32); p = np.ascontiguousarray(pos, np.int32)
        self.lib.orc_forward_many(self.h, arr, _ip(t), _ip(p), n, n_threads)


class OracleSeq:
    NAMES = dict(x=0, xb=1, xb2=2, hb=3, hb2=4, q=5, att=6, logits=7, key_cache=8, value_cache=9)

    def __init__(self, model):
        self.m = model
        self.h = model.lib.orc_seq_new(model.h)

    def forward(self, token, pos):
        p = self.m.lib.orc_forward(self.m.h, self.h, int(token), int(pos))
        return np.ctypeslib.as_array(p, shape=(self.m.vocab_size,)).copy()

    def buf(self, name):
        m = self.m
        sizes = dict(x=m.dim, xb=m.dim, xb2=m.dim, hb=m.hidden_dim, hb2=m.hidden_dim, q=m.dim,
                     att=m.n_heads * m.seq_len, logits=m.vocab_size,
                     key_cache=m.n_layers * m.seq_len * m.kv_dim, value_cache=m.n_layers * m.seq_len * m.kv_dim)
        p = m.lib.orc_seq_ptr(self.h, self.NAMES[name])
        return np.ctypeslib.as_array(p, shape=(sizes[name],))

    def qbuf(self, which):
        m = self.m
        n = m.dim if which == 0 else m.hidden_dim
        q = np.ctypeslib.as_array(m.lib.orc_seq_q(self.h, which), shape=(n,)).copy()
        s = np.ctypeslib.as_array(m.lib.orc_seq_s(self.h, which), shape=(n // m.group_size,)).copy()
        return q, s


# ---------------------------------------------------------------- oracle/_ref ----

def ref_available(kind="fp32", fast=False):
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", f"libref_{kind}{'_fast' if fast else ''}.so"))


def cpu_has_avx2_fma():
    try:
        flags = open("/proc/cpuinfo").read()
        return " avx2" in flags and " fma" in flags
    except OSError:
        return False


class Ref32:
    """/root/reference/run.cc compiled unmodified (oracle/ref_shim_fp32.cc)."""

    def __init__(self, fast=False):
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", f"libref_fp32{'_fast' if fast else ''}.so"))
        self.lib = lib
        lib.ref32_open.restype = C.c_void_p; lib.ref32_open.argtypes = [C.c_char_p]
        lib.ref32_clone.restype = C.c_void_p; lib.ref32_clone.argtypes = [C.c_void_p]
        lib.ref32_config.argtypes = [C.c_void_p, i32p]
        lib.ref32_forward.restype = f32p; lib.ref32_forward.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.ref32_forward_many.argtypes = [C.POINTER(C.c_void_p), i32p, i32p, C.c_int, C.c_int]
        lib.ref32_state_ptr.restype = f32p; lib.ref32_state_ptr.argtypes = [C.c_void_p, C.c_int]
        lib.ref32_rmsnorm.argtypes = [f32p, f32p, f32p, C.c_int]
        lib.ref32_softmax.argtypes = [f32p, C.c_int]
        lib.ref32_matmul.argtypes = [f32p, f32p, f32p, C.c_int, C.c_int]
        lib.ref32_argmax.restype = C.c_int; lib.ref32_argmax.argtypes = [f32p, C.c_int]
        lib.ref32_tok_open.restype = C.c_void_p; lib.ref32_tok_open.argtypes = [C.c_char_p, C.c_int]
        lib.ref32_encode.restype = C.c_int
        lib.ref32_encode.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_int, i32p]
        lib.ref32_decode.restype = C.c_char_p; lib.ref32_decode.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.ref32_sampler_new.restype = C.c_void_p
        lib.ref32_sampler_new.argtypes = [C.c_int, C.c_float, C.c_float, C.c_ulonglong]
        lib.ref32_sample.restype = C.c_int; lib.ref32_sample.argtypes = [C.c_void_p, f32p]
        lib.ref32_max_threads.restype = C.c_int

    def open(self, path):
        h = self.lib.ref32_open(path.encode())
        cfg = (C.c_int * 7)()
        self.lib.ref32_config(h, cfg)
        return h, list(cfg)

    def forward(self, h, token, pos, vocab):
        return np.ctypeslib.as_array(self.lib.ref32_forward(h, int(token), int(pos)), shape=(vocab,)).copy()

    def state(self, h, which, n):
        return np.ctypeslib.as_array(self.lib.ref32_state_ptr(h, which), shape=(n,))

    def forward_many(self, hs, tokens, pos, n_threads):
        n = len(hs)
        arr = (C.c_void_p * n)(*hs)
        t = np.ascontiguousarray(tokens, np.int32); p = np.ascontiguousarray(pos, np.int32)
        self.lib.ref32_forward_many(arr, _ip(t), _ip(p), n, n_threads)


class RefQ8:
    """/root/reference/runq.c compiled unmodified (oracle/ref_shim_q8.c)."""

    def __init__(self, fast=False):
        lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", f"libref_q8{'_fast' if fast else ''}.so"))
        self.lib = lib
        lib.refq8_open.restype = C.c_void_p; lib.refq8_open.argtypes = [C.c_char_p]
        lib.refq8_clone.restype = C.c_void_p; lib.refq8_clone.argtypes = [C.c_void_p]
        lib.refq8_config.argtypes = [C.c_void_p, i32p]
        lib.refq8_forward.restype = f32p; lib.refq8_forward.argtypes = [C.c_void_p, C.c_int, C.c_int]
        lib.refq8_forward_many.argtypes = [C.POINTER(C.c_void_p), i32p, i32p, C.c_int, C.c_int]
        lib.refq8_state_ptr.restype = f32p; lib.refq8_state_ptr.argtypes = [C.c_void_p, C.c_int]
        lib.refq8_state_q.restype = i8p; lib.refq8_state_q.argtypes = [C.c_void_p, C.c_int]
        lib.refq8_state_s.restype = f32p; lib.refq8_state_s.argtypes = [C.c_void_p, C.c_int]
        lib.refq8_group_size.restype = C.c_int
        lib.refq8_set_group_size.argtypes = [C.c_int]
        lib.refq8_quantize.argtypes = [i8p, f32p, f32p, C.c_int]
        lib.refq8_matmul.argtypes = [f32p, i8p, f32p, i8p, f32p, C.c_int, C.c_int]
        lib.refq8_max_threads.restype = C.c_int

    def open(self, path):
        h = self.lib.refq8_open(path.encode())
        cfg = (C.c_int * 8)()
        self.lib.refq8_config(h, cfg)
        return h, list(cfg)

    def forward(self, h, token, pos, vocab):
        return np.ctypeslib.as_array(self.lib.refq8_forward(h, int(token), int(pos)), shape=(vocab,)).copy()

    def quantize(self, x, gs):
        self.lib.refq8_set_group_size(gs)
        x = np.ascontiguousarray(x, np.float32)
        q = np.zeros(x.size, np.int8); s = np.zeros(x.size // gs, np.float32)
        self.lib.refq8_quantize(_bp(q), _fp(s), _fp(x), x.size)
        return q, s

    def matmul(self, wq, ws, xq, xs, gs):
        self.lib.refq8_set_group_size(gs)
        wq = np.ascontiguousarray(wq, np.int8); ws = np.ascontiguousarray(ws, np.float32)
        xq = np.ascontiguousarray(xq, np.int8); xs = np.ascontiguousarray(xs, np.float32)
        d, n = wq.shape
        o = np.empty(d, np.float32)
        self.lib.refq8_matmul(_fp(o), _bp(xq), _fp(xs), _bp(wq), _fp(ws), n, d)
        return o

    def forward_many(self, hs, tokens, pos, n_threads):
        n = len(hs)
        arr = (C.c_void_p * n)(*hs)
        t = np.ascontiguousarray(tokens, np.int32); p = np.ascontiguousarray(pos, np.int32)
        self.lib.refq8_forward_many(arr, _ip(t), _ip(p), n, n_threads)
