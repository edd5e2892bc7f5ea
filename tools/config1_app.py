#!/usr/bin/env python
"""BASELINE.json configs[0]: stories110M (fp32, seeded random-init of that architecture -- no checkpoint on the box) greedy
decode of the 128 prompts of assets/in/gen_in_128.txt through the shipped driver (./build/apps/llama -m test -t 0), with the
app's own throughput line (the reference's definition: sum(pos - 1) / wall, weight upload included, src/llama.cpp:1596-1615)
and a parity sample: the first positions of a few requests against the oracle's greedy loop.

    python tools/config1_app.py [--batch 64] [--check-requests 3] [--check-positions 40]
"""
import argparse, os, re, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from pkgload import load_pkg
from oracle_lib import Oracle
import test_host_cpp as T

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=64)
ap.add_argument("--check-requests", type=int, default=3)
ap.add_argument("--check-positions", type=int, default=40)
a = ap.parse_args()
ck = load_pkg("checkpoint")
cfg = ck.STORIES110M
t0 = time.time()
w = ck.random_weights(cfg, seed=1337, std=0.02)
td = tempfile.mkdtemp(prefix="cfg1_")
path = os.path.join(td, "stories110M_random.bin")
ck.write_v0(path, cfg, w)
print(f"checkpoint written in {time.time() - t0:.1f} s ({os.path.getsize(path)} bytes)", flush=True)
inp = os.path.join(T.ASSETS, "gen_in_128.txt")
out_txt, out_tok = os.path.join(td, "out.txt"), os.path.join(td, "tok.txt")
cmd = [T.LLAMA, path, "-m", "test", "-f", inp, "-o", out_txt, "-k", out_tok, "-b", str(a.batch), "-t", "0", "-z", T.TOKENIZER, "-g", "1"]
t0 = time.time()
r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
wall = time.time() - t0
print(f"app rc={r.returncode} wall {wall:.1f} s", flush=True)
for line in r.stdout.splitlines():
    if re.search(r"achieved|step loop|Load model|total elapsed", line):
        print("  " + line.strip(), flush=True)
if r.returncode != 0:
    print(r.stdout[-2000:])
    sys.exit(1)
lines = open(out_tok).read().splitlines()
n = int(lines[0])
got = [[int(x) for x in l.split()] for l in lines[1:1 + n]]
print(f"{n} requests, {sum(len(g) - 1 for g in got)} positions generated", flush=True)
# parity sample
prompts = open(inp).read().splitlines()[1:1 + n]
gold = T.golden_prompts()
m = Oracle().open(path)
ok = True
for req in range(min(a.check_requests, n)):
    ids = gold[prompts[req]]
    seq = m.new_seq()
    token, pos, want = ids[0], 0, [ids[0]]
    flips = None
    while pos < a.check_positions:
        logits = seq.forward(token, pos)
        nx = ids[pos + 1] if pos < len(ids) - 1 else int(np.argmax(logits))
        pos += 1
        if nx in (1, 2):
            break
        want.append(nx)
        token = nx
        if got[req][:len(want)] != want and flips is None:
            top2 = np.partition(logits, -2)[-2:]
            flips = (pos, float(top2[1] - top2[0]), float(np.abs(logits).max()))
            break
    if flips is None:
        print(f"request {req}: first {len(want)} token ids identical to the oracle's greedy loop", flush=True)
    else:
        pos_, margin, scale = flips
        near_tie = margin <= 1e-4 * max(1.0, scale)
        print(f"request {req}: differs at position {pos_} (oracle top-2 margin {margin:.3e}, max |logit| {scale:.3f}) -> {'near-tie' if near_tie else 'MISMATCH'}", flush=True)
        ok = ok and near_tie
print("CONFIG1", "PASS" if ok else "FAIL")
sys.exit(0 if ok else 1)
