#!/usr/bin/env python
"""Algorithmic integer-op count of the SHA-256d nonce scan, by dataflow (reproduces SURVEY.md §8d).

Builds the dataflow graph of one candidate — block 2 of hash 1 from the midstate, then hash 2 — with
nonce-independent values folded (they are per-job constants), keeps only what the early-reject test needs
(digest word 7 = IV7 + e after round 60 of hash 2), and counts 3-input ALU-class operations:
  SHF   one rotate or shift of a nonce-dependent word
  LOP3  one 3-input boolean (xor3 of Sigma/sigma, Ch, Maj) of nonce-dependent words
  IADD3 an n-term sum costs ceil((terms - 1) / 2), all constant terms of a sum counting as one term
Run: python scripts/count_ops.py
"""
import math

class Node:
    __slots__ = ("kind", "deps", "const", "cost")

    def __init__(self, kind, deps, const, cost):
        self.kind, self.deps, self.const, self.cost = kind, deps, const, cost


def C():
    return Node("c", (), True, (0, 0, 0))


def V():
    return Node("v", (), False, (0, 0, 0))


def shf(x):
    return C() if x.const else Node("shf", (x,), False, (1, 0, 0))


def lop3(*xs):
    return C() if all(x.const for x in xs) else Node("lop3", tuple(x for x in xs if not x.const), False, (0, 1, 0))


def add(*xs):
    var = [x for x in xs if not x.const]
    if not var:
        return C()
    terms = len(var) + (1 if len(var) < len(xs) else 0)
    return Node("add", tuple(var), False, (0, 0, math.ceil((terms - 1) / 2)))


def big_sigma(x):
    return lop3(shf(x), shf(x), shf(x))


small_sigma = big_sigma  # two rotates + one shift, then xor3: same op shape


def total(outputs):
    seen, tot = set(), [0, 0, 0]
    stack = list(outputs)
    while stack:
        n = stack.pop()
        if id(n) in seen:
            continue
        seen.add(id(n))
        for k in range(3):
            tot[k] += n.cost[k]
        stack.extend(n.deps)
    return tot


def t1_t2_shared(h, e, f, g, kw, d, a, b, c):
    """One round with T1 shared between the two outputs, as any implementation does."""
    t1 = add(h, big_sigma(e), lop3(e, f, g), C(), kw)
    return add(d, t1), add(t1, big_sigma(a), lop3(a, b, c))


def compress_shared(state, w, rounds, schedule_out=None):
    w = list(w)
    for i in range(16, rounds):
        w.append(add(small_sigma(w[i - 2]), w[i - 7], small_sigma(w[i - 15]), w[i - 16]))
    if schedule_out is not None:
        schedule_out.extend(w[16:])
    a, b, c, d, e, f, g, h = state
    states = []
    for i in range(rounds):
        new_e, new_a = t1_t2_shared(h, e, f, g, w[i], d, a, b, c)
        a, b, c, d, e, f, g, h = new_a, a, b, c, new_e, e, f, g
        states.append((a, b, c, d, e, f, g, h))
    return states


def scan_graph(fold=True, early=True, schedule_out=None):
    # hash 1, block 2: midstate constant; W0..2 constant, W3 = nonce, W4..15 constant
    mid = [C() for _ in range(8)]
    w1 = [C(), C(), C(), V()] + [C() for _ in range(12)]
    if not fold:  # pretend nothing is constant (reference point without folding)
        mid = [V() for _ in range(8)]
        w1 = [V() for _ in range(16)]
    s1 = compress_shared(mid, w1, 64, schedule_out)[-1]
    digest = [add(x, C()) if fold else add(x, V()) for x in s1]           # + midstate
    iv = [C() for _ in range(8)] if fold else [V() for _ in range(8)]
    w2 = digest + ([C() for _ in range(8)] if fold else [V() for _ in range(8)])
    rounds2 = 61 if early else 64
    st2 = compress_shared(iv, w2, rounds2)
    if early:
        return [st2[60][4]]                                              # e after round 60
    return [add(x, C()) for x in st2[-1]]                                 # all 8 digest words


def version_shared_counts():
    """BIP-320 version instances of one job share block 2 of hash 1 (the version sits in block 1), so that
    block's message schedule W16..W63 is per-NONCE work: returns ((SHF, LOP3, IADD3) per nonce for the schedule,
    (SHF, LOP3, IADD3) per (nonce, version) for everything else).  K + W is one shared term either way."""
    sched = []
    whole = total(scan_graph(fold=True, early=True, schedule_out=sched))
    per_nonce = total(sched)
    return tuple(per_nonce), tuple(w - n for w, n in zip(whole, per_nonce))


if __name__ == "__main__":
    for name, kw in (("midstate, nothing folded, all 8 digest words", dict(fold=False, early=False)),
                     ("nonce-independent work folded, early reject on digest word 7", dict(fold=True, early=True))):
        s, l, a = total(scan_graph(**kw))
        print(f"{name}: SHF {s}  LOP3 {l}  IADD3 {a}  total {s + l + a}   (SHF+LOP3 = {s + l})")
    pn, pv = version_shared_counts()
    print(f"version-shared scan: per nonce (hash-1 schedule) SHF {pn[0]} LOP3 {pn[1]} IADD3 {pn[2]}; "
          f"per version instance SHF {pv[0]} LOP3 {pv[1]} IADD3 {pv[2]}  (SHF+LOP3 = {pv[0] + pv[1]} + {pn[0] + pn[1]}/V)")
    print("SURVEY.md §8d: unfolded 1344 / 704 / 720 = 2768; folded + early reject 1176 / 628 / 646 = 2450")
