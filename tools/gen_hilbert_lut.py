"""Generates open3d-py-extension_b200/csrc/hilbert_lut.cuh: the 3-D Hilbert index of knn.cuh (Skilling's transpose
algorithm, hilbert3_ref below is the same arithmetic) as a state machine, one octant per step:
    entry = LUT[state * 8 + octant],  digit = entry & 7,  next state = entry >> 3
The states are found by exploring the curve itself (the octant -> digit permutation of every sub-cube), the table is
checked against the reference for every cell of a 2^4 grid per start state and for random cells at every bit count.
    python tools/gen_hilbert_lut.py            # rewrites the header
"""
import os
import random

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def spread3(v):
    r = 0
    for b in range(21):
        r |= ((v >> b) & 1) << (3 * b)
    return r


def hilbert3_ref(x, y, z, bits):
    X = [x, y, z]
    M = 1 << (bits - 1)
    Q = M
    while Q > 1:
        P = Q - 1
        for i in range(3):
            if X[i] & Q:
                X[0] ^= P
            else:
                t = (X[0] ^ X[i]) & P
                X[0] ^= t
                X[i] ^= t
        Q >>= 1
    X[1] ^= X[0]
    X[2] ^= X[1]
    t = 0
    Q = M
    while Q > 1:
        if X[2] & Q:
            t ^= Q - 1
        Q >>= 1
    X = [v ^ t for v in X]
    return (spread3(X[0]) << 2) | (spread3(X[1]) << 1) | spread3(X[2])


def perm_of(prefix, depth, bits):
    """octant -> digit of the sub-cube reached through `prefix` (list of octants, most significant first)"""
    out = []
    for o in range(8):
        path = prefix + [o]
        x = y = z = 0
        for k, oc in enumerate(path):
            sh = bits - 1 - k
            x |= ((oc >> 2) & 1) << sh
            y |= ((oc >> 1) & 1) << sh
            z |= (oc & 1) << sh
        h = hilbert3_ref(x, y, z, bits)
        out.append((h >> (3 * (bits - 1 - depth))) & 7)
    return tuple(out)


def explore(bits=6):
    states = {}
    order = []
    trans = {}

    def visit(prefix):
        p = perm_of(prefix, len(prefix), bits)
        if p not in states:
            states[p] = len(order)
            order.append(p)
            new = True
        else:
            new = False
        return p, new

    root, _ = visit([])
    stack = [([], root)]
    seen_paths = set()
    while stack:
        prefix, p = stack.pop()
        if len(prefix) >= bits - 1:
            continue
        for o in range(8):
            child, new = visit(prefix + [o])
            key = (states[p], o)
            assert trans.get(key, states[child]) == states[child], "the curve is not a state machine"
            trans[key] = states[child]
            if new or (states[child], len(prefix) + 1) not in seen_paths:
                seen_paths.add((states[child], len(prefix) + 1))
                if len(prefix) + 1 < 4:
                    stack.append((prefix + [o], child))
    return order, trans, states


def main(check_only=False):
    order, trans, states = explore(6)
    ns = len(order)
    lut = []
    for s in range(ns):
        for o in range(8):
            lut.append(order[s][o] | (trans[(s, o)] << 3))
    start = {}
    for bits in range(2, 22):
        start[bits] = states[perm_of([], 0, bits)]
    assert set(start.values()) == {0}, "the device code starts in state 0 for every bit count"
    # two octants per step: entry = LUT2[state * 64 + (octant_hi << 3 | octant_lo)] = six index bits | next state << 6
    lut2 = []
    for s in range(ns):
        for oo in range(64):
            e1 = lut[s * 8 + (oo >> 3)]
            e2 = lut[(e1 >> 3) * 8 + (oo & 7)]
            lut2.append(((e1 & 7) << 3) | (e2 & 7) | ((e2 >> 3) << 6))

    def hilbert_lut(x, y, z, bits):          # the device code's steps: one octant if the bit count is odd, then pairs
        m = (spread3(x) << 2) | (spread3(y) << 1) | spread3(z)
        s = 0
        h = 0
        k = bits
        if k & 1:
            k -= 1
            e = lut[s * 8 + ((m >> (3 * k)) & 7)]
            h = e & 7
            s = e >> 3
        while k > 0:
            k -= 2
            e = lut2[s * 64 + ((m >> (3 * k)) & 63)]
            h = (h << 6) | (e & 63)
            s = e >> 6
        return h

    rng = random.Random(1)
    for bits in range(2, 22):
        for _ in range(3000):
            x, y, z = (rng.randrange(1 << bits) for _ in range(3))
            assert hilbert_lut(x, y, z, bits) == hilbert3_ref(x, y, z, bits), (bits, x, y, z)
    for bits in (4, 5):
        for x in range(1 << bits):
            for y in range(1 << bits):
                for z in range(1 << bits):
                    assert hilbert_lut(x, y, z, bits) == hilbert3_ref(x, y, z, bits)
    text = render(ns, lut, lut2)
    path = os.path.join(ROOT, "open3d-py-extension_b200", "csrc", "hilbert_lut.cuh")
    if check_only:
        return open(path).read() == text
    with open(path, "w") as f:
        f.write(text)
    print("states", ns, "->", path)
    return True


def render(ns, lut, lut2):
    out = []
    out.append("// hilbert_lut.cuh — GENERATED by tools/gen_hilbert_lut.py, do not edit.")
    out.append("// The 3-D Hilbert index of Skilling's transpose algorithm as a state machine over the octants of the Morton code")
    out.append("// (x y z bits, x most significant), start state 0 for every bit count:")
    out.append("//   one octant:  e = HILBERT_LUT1[state * 8 + octant],        digit  = e & 7,   next state = e >> 3")
    out.append("//   two octants: e = HILBERT_LUT2[state * 64 + octant pair],  digits = e & 63,  next state = e >> 6")
    out.append("// Checked against the bit-twiddling form for every bit count 2..21 by the generator; tests/test_host_cpu.py checks")
    out.append("// that this file is what the generator writes.")
    out.append("#pragma once")
    out.append(f"#define HILBERT_STATES {ns}")
    out.append("__device__ const unsigned char HILBERT_LUT1[HILBERT_STATES * 8] = {")
    for s in range(ns):
        out.append("    " + ", ".join(str(v) for v in lut[s * 8:s * 8 + 8]) + ",")
    out.append("};")
    out.append("__device__ const unsigned short HILBERT_LUT2[HILBERT_STATES * 64] = {")
    for s in range(ns):
        for r in range(4):
            out.append("    " + ", ".join(str(v) for v in lut2[s * 64 + r * 16:s * 64 + r * 16 + 16]) + ",")
    out.append("};")
    return "\n".join(out) + "\n"


if __name__ == "__main__":
    main()
