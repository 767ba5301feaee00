#!/usr/bin/env python
"""Symbolic check of the transpose-reduction network of fullpel_ext.cu (host only, no GPU).

Replays, for every PU class, the sequence of pair_min / shfl_min steps the kernel performs over one tile (cols4 / cols2,
chain<LC, LT>, the trailing shuffles and the lane predicate of the RED) on SETS of contributions instead of keys, and
checks that every PU of the class is issued exactly once per flush, by a lane whose set is exactly the 16 quads' keys of
that PU (column, time step), and that lane_pu() names that PU. Mirrors: lane_pu, cols4, cols2, chain of fullpel_ext.cu.
"""
import itertools

L = [8, 4, 2, 1]


def lbit(lane, i):
    return (lane >> (3 - i)) & 1


def pair_min(a, b, bit, idx):
    """a, b: per-lane sets. Lanes with the bit clear keep a over {lane, lane ^ bit}, the others b."""
    out = []
    for lane in range(32):
        src = b if lbit(lane, idx) else a
        out.append(src[lane] | src[lane ^ bit])
    return out


def shfl_min(v, bit):
    return [v[l] | v[l ^ bit] for l in range(32)]


def chain_run(LC, LT, values_by_t):
    """values_by_t[t] = per-lane sets entering the chain at step t. Yields (t, out) when a value drops out."""
    c = [None] * 3
    for t, v in enumerate(values_by_t):
        done = True
        for i in range(LT):
            if ((t >> i) & 1) == 0:
                c[i] = v
                done = False
                break
            v = pair_min(c[i], v, 8 >> (LC + i), LC + i)
        if done:
            yield t, v


def lane_pu(table, lane):
    H, v0, v1, v2, v3 = (lane >> 4) & 1, (lane >> 3) & 1, (lane >> 2) & 1, (lane >> 1) & 1, lane & 1
    j, q, p = v0 + 2 * v1 + 4 * H, v0 + 2 * H, H
    return {
        'T8A': (21, (v2 + 2 * v3) * 8 + j), 'T8B': (21, (4 + v2 + 2 * v3) * 8 + j),
        'T16x8': (95, (v1 + 2 * v2 + 4 * v3) * 4 + q), 'T32x8': (169, (v0 + 2 * v1 + 4 * v2) * 2 + p),
        'T8x16': (137, (v2 + 2 * v3) * 8 + j), 'T16x16': (5, (v1 + 2 * v2) * 4 + q), 'T32x16': (87, (v0 + 2 * v1) * 2 + p),
        'T64x16': (201, v0 + 2 * v1), 'T8x32': (185, v2 * 8 + j), 'T16x32': (129, v1 * 4 + q), 'T32x32': (1, v0 * 2 + p),
        'T64x32': (85, v0), 'T16x64': (205, q), 'T32x64': (127, p),
    }[table]


# class: (tables per flush, base, columns per thread n (0 = 64-wide, halves already added), time steps T, steps per flush,
#         trailing shuffle bits, lane mask that must be zero to issue)
CLASSES = [
    (['T8A', 'T8B'], 4, 8, 4, [], 0),
    (['T16x8'], 2, 8, 8, [], 0),
    (['T32x8'], 1, 8, 8, [1], 1),
    (['T8x16'], 4, 4, 4, [], 0),
    (['T16x16'], 2, 4, 4, [1], 1),
    (['T32x16'], 1, 4, 4, [2, 1], 3),
    (['T64x16'], 0, 4, 4, [2, 1], 19),
    (['T8x32'], 4, 2, 2, [1], 1),
    (['T16x32'], 2, 2, 2, [2, 1], 3),
    (['T32x32'], 1, 2, 2, [4, 2, 1], 7),
    (['T64x32'], 0, 2, 2, [4, 2, 1], 23),
    (['T16x64'], 2, 1, 1, [4, 2, 1], 7),
    (['T32x64'], 1, 1, 1, [8, 4, 2, 1], 15),
]


def check():
    for tables, n, T, per_flush, trail, mask in CLASSES:
        ncols = {4: 8, 2: 4, 1: 2, 0: 1}[n]
        LC = {4: 2, 2: 1, 1: 0, 0: 0}[n]
        LT = per_flush.bit_length() - 1
        vals = []
        for t in range(T):
            regs = []
            for jj in range(max(n, 1)):
                regs.append([frozenset([((jj + n * (lane >> 4)) if n else 0, t, lane & 15)]) for lane in range(32)])
            if n == 4:
                v = pair_min(pair_min(regs[0], regs[1], 8, 0), pair_min(regs[2], regs[3], 8, 0), 4, 1)
            elif n == 2:
                v = pair_min(regs[0], regs[1], 8, 0)
            else:
                v = regs[0]
            vals.append(v)
        issued = {}
        flush = 0
        for f0 in range(0, T, per_flush):
            for t, out in chain_run(LC, LT, vals[f0:f0 + per_flush]):
                for bit in trail:
                    out = shfl_min(out, bit)
                table = tables[flush]
                for lane in range(32):
                    if lane & mask:
                        continue
                    base, idx = lane_pu(table, lane)
                    col, step = idx % ncols, idx // ncols
                    want = frozenset((col, step, ql) for ql in range(16))
                    assert out[lane] == want, (table, lane, sorted(out[lane])[:4], sorted(want)[:4])
                    assert (table, idx) not in issued, (table, idx)
                    issued[(table, idx)] = lane
            flush += 1
        total = ncols * T
        got = sorted(idx for (_, idx) in issued)
        assert got == list(range(total)), (tables, got)
        print("ok", tables, "PUs", total)


if __name__ == "__main__":
    check()
