#!/usr/bin/env python3
"""Warp-level cost model of the linear-program walk (kernel modes 3 / 4) on the Cornell box, CPU only.

Why: ncu (profiles/r2o_*) says the walk is bound by the issue rate at ~15 of 32 lanes.  Before writing another variant of
trace_linear, this replays real lane populations (32 pixels of a tile, every lane on its own path, regenerated when it
ends -- as the persistent kernel does) through the program and counts warp-level block executions under different
schedules.  Float64 numpy, no claim of bit-exactness: only the statistics of who is where matter.

  P0  today's walk: every iteration each lane does ONE op; the lanes meet by kind of op (switch)
  P1  kind-ordered sweep: an iteration runs the blocks in a fixed order and a lane may pass through several of them
  P2  P0 with the wrappers fused (translate+rotate one op, un-rotate+un-translate one op)
  P3  P1 with the wrappers fused
  P4  majority vote: every iteration runs ONE block, the kind most lanes wait at (P4f: with the wrappers fused)
usage: sim_linear_walk.py [warps=40] [bounce_steps=60]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracing_b200 as rt  # noqa: E402

K_SPHERE, K_BVH, K_TRI, K_QUAD, K_T, K_R, K_TX, K_RX = range(8)
rng = np.random.default_rng(7)


def cornell_quads():
    def box(a, b):
        x0, y0, z0 = a
        x1, y1, z1 = b
        dx, dy, dz = (x1 - x0, 0, 0), (0, y1 - y0, 0), (0, 0, z1 - z0)
        return [((x0, y0, z1), dx, dy), ((x1, y0, z1), (0, 0, -(z1 - z0)), dy), ((x1, y0, z0), (-(x1 - x0), 0, 0), dy),
                ((x0, y0, z0), dz, dy), ((x0, y1, z1), dx, (0, 0, -(z1 - z0))), ((x0, y0, z0), dx, dz)]
    q = [((555, 0, 0), (0, 555, 0), (0, 0, 555)), ((0, 0, 0), (0, 555, 0), (0, 0, 555)), ((343, 554, 332), (-130, 0, 0), (0, 0, -105)),
         ((0, 0, 0), (555, 0, 0), (0, 0, 555)), ((555, 555, 555), (-555, 0, 0), (0, 0, -555)), ((0, 0, 555), (555, 0, 0), (0, 555, 0))]
    q += box((0, 0, 0), (165, 330, 165)) + box((0, 0, 0), (165, 165, 165))
    out = []
    for Q, U, V in q:
        Q, U, V = (np.array(x, dtype=np.float64) for x in (Q, U, V))
        n = np.cross(U, V)
        w = n / n.dot(n)
        nn = n / np.linalg.norm(n)
        out.append(dict(q=Q, u=U, v=V, n=nn, d=nn.dot(Q), w=w))
    return out


def load_program():
    hs = rt.HostScene(1)
    prog = hs.program()
    hs.close()
    ops = []
    for op in prog:
        u = op.view(np.uint32)
        f = op.view(np.float32).astype(np.float64)
        ops.append(dict(kind=int(u[7] & 7), fused=bool(u[7] & 0x200), arg=int(u[6]), f=f))
    return ops


def trace(ops, quads, o, d):
    """-> (list of (kind, phase) per op visited, hit dict | None).  phase for planars: 0 = plane only, 1 = + inside, 2 = + record."""
    o, d = o.copy(), d.copy()
    tmax, hit, seq, pc, in_wrap_hit = np.inf, None, [], 0, False
    stack = []
    while pc < len(ops):
        op = ops[pc]
        k, f = op["kind"], op["f"]
        pc += 1
        if k == K_BVH:
            seq.append((K_BVH, 0))
            t0, t1, ok = 0.001, tmax, True
            for ax, (lo, hi) in enumerate(((f[0], f[1]), (f[2], f[3]), (f[4], f[5]))):
                inv = 1.0 / d[ax] if d[ax] != 0 else np.inf
                ta, tb = (lo - o[ax]) * inv, (hi - o[ax]) * inv
                if inv < 0:
                    ta, tb = tb, ta
                t0, t1 = max(t0, ta), min(t1, tb)
                if t1 <= t0:
                    ok = False
            if not ok:
                pc = op["arg"]
        elif k in (K_QUAD, K_TRI):
            g = quads[op["arg"]]
            den = g["n"].dot(d)
            ph = 0
            if abs(den) > 1e-8:
                t = (g["d"] - g["n"].dot(o)) / den
                if 0.001 <= t <= tmax:
                    ph = 1
                    p = o + t * d
                    pl = p - g["q"]
                    a, b = g["w"].dot(np.cross(pl, g["v"])), g["w"].dot(np.cross(g["u"], pl))
                    if 0 <= a <= 1 and 0 <= b <= 1:
                        ph = 2
                        tmax = t
                        n = g["n"] if den < 0 else -g["n"]
                        hit = dict(p=p, n=n, prim=op["arg"])
                        in_wrap_hit = len(stack) > 0
            seq.append((K_QUAD, ph))
        elif k == K_T:
            seq.append((K_T, 1 if op["fused"] else 0)); in_wrap_hit = False
            o = o - f[0:3]; stack.append(1)
            if op["fused"]:                                    # translate(rotate(x)) as one op: {offset,0} {sin,cos,..}
                sn, cs = f[4], f[5]
                o = np.array([o[0] * cs - o[2] * sn, o[1], o[0] * sn + o[2] * cs])
                d = np.array([d[0] * cs - d[2] * sn, d[1], d[2] * cs + d[0] * sn])
        elif k == K_TX:
            seq.append((K_TX, 0))
            o = o + f[0:3]
            if in_wrap_hit:
                hit["p"] = hit["p"] + f[0:3]
            stack.pop()
        elif k == K_R:
            seq.append((K_R, 0)); in_wrap_hit = False
            sn, cs = f[0], f[1]
            o = np.array([o[0] * cs - o[2] * sn, o[1], o[0] * sn + o[2] * cs])
            d = np.array([d[0] * cs - d[2] * sn, d[1], d[2] * cs + d[0] * sn])
        elif k == K_RX:
            seq.append((K_RX, 0))
            sn, cs = f[0], f[1]
            rot = lambda v: np.array([v[0] * cs + v[2] * sn, v[1], v[2] * cs - v[0] * sn])
            o, d = rot(o), rot(d)
            if in_wrap_hit:
                hit["p"], hit["n"] = rot(hit["p"]), rot(hit["n"])
            if op["fused"]:                                    # ... and the way out: {sin,cos,offset.xy} {offset.z,..}
                seq[-1] = (K_RX, 1)
                o = o + f[2:5]
                if in_wrap_hit:
                    hit["p"] = hit["p"] + f[2:5]
                stack.pop()
        else:
            raise RuntimeError("unexpected op kind %d" % k)
    return seq, hit


def unit_vector():
    while True:
        v = rng.uniform(-1, 1, 3)
        s = v.dot(v)
        if 0 < s <= 1:
            return v / np.sqrt(s)


class Lane:
    def __init__(self, x, y):
        self.x, self.y, self.depth = x, y, 0
        self.new_path()

    def new_path(self):
        # camera of cornell_box_scene: lookfrom (278,278,-800), vfov 40, 600x600
        h = 2 * np.tan(np.radians(40) / 2) * 800.0
        px = (self.x + rng.uniform()) / 600.0 - 0.5
        py = 0.5 - (self.y + rng.uniform()) / 600.0
        self.o = np.array([278.0, 278.0, -800.0])
        self.d = np.array([-px * h, py * h, 800.0])          # the reference's image is mirrored in x (w = lookfrom - lookat)
        self.depth = 0

    def step(self, ops, quads):
        seq, hit = trace(ops, quads, self.o, self.d)
        self.depth += 1
        if hit is None or hit["prim"] == 2 or self.depth >= 50:
            self.new_path()
        else:
            dirn = hit["n"] + unit_vector()
            if np.abs(dirn).max() < 1e-8:
                dirn = hit["n"]
            self.o, self.d = hit["p"], dirn
        return seq


# block costs in warp-instructions (calibrated against the per-line breakdown of the ncu capture, see main)
COST = dict(loop=16.0, box=30.0, qa=30.0, qb=38.0, qrec=16.0, t=8.0, r=26.0, rx=36.0, tx=12.0, next=9.0, skip=2.0)


def fuse_wrappers(seq):
    out, i = [], 0
    while i < len(seq):
        if seq[i][0] == K_T and i + 1 < len(seq) and seq[i + 1][0] == K_R:
            out.append((K_T, 1)); i += 2
        elif seq[i][0] == K_RX and i + 1 < len(seq) and seq[i + 1][0] == K_TX:
            out.append((K_RX, 1)); i += 2
        else:
            out.append(seq[i]); i += 1
    return out


def block_cost(kind, items):
    """cost of one execution of a kind's block for the lanes in `items` (their (kind, phase) tuples)"""
    c = COST
    if kind == K_BVH:
        return c["box"]
    if kind == K_QUAD:
        ph = max(p for _, p in items)
        return c["qa"] + (c["qb"] if ph >= 1 else 0) + (c["qrec"] if ph >= 2 else 0)
    fused = any(p for _, p in items)
    if kind == K_T:
        return c["t"] + (c["r"] if fused else 0)
    if kind == K_R:
        return c["r"]
    if kind == K_RX:
        return c["rx"] + (c["tx"] if fused else 0)
    return c["tx"]


def cost_p0(seqs):
    n = max(len(s) for s in seqs)
    tot, parts = 0.0, {}
    for i in range(n):
        tot += COST["loop"]
        parts["loop"] = parts.get("loop", 0) + COST["loop"]
        present = {}
        for s in seqs:
            if i < len(s):
                present.setdefault(s[i][0], []).append(s[i])
        for k, items in present.items():
            x = block_cost(k, items)
            tot += x
            parts[k] = parts.get(k, 0) + x
    return tot, n, parts


def cost_sweep(seqs, order):
    pos = [0] * len(seqs)
    tot, iters = 0.0, 0
    while any(pos[j] < len(s) for j, s in enumerate(seqs)):
        iters += 1
        tot += COST["loop"] - COST["next"]          # the loop keeps its vote / branch; the op load moves into the blocks
        for k in order:
            items, who = [], []
            for j, s in enumerate(seqs):
                if pos[j] < len(s) and s[pos[j]][0] == k:
                    items.append(s[pos[j]]); who.append(j)
            if items:
                tot += block_cost(k, items) + COST["next"]
                for j in who:
                    pos[j] += 1
            else:
                tot += COST["skip"]
    return tot, iters


def cost_majority(seqs, vote=14.0, weight=None):
    """every iteration ONE block: the kind most lanes are waiting at (ties: the cheaper block); the others hold their op"""
    pos = [0] * len(seqs)
    tot, iters = 0.0, 0
    while True:
        present = {}
        for j, s in enumerate(seqs):
            if pos[j] < len(s):
                present.setdefault(s[pos[j]][0], []).append(j)
        if not present:
            break
        iters += 1
        k = max(present, key=lambda x: (len(present[x]) * (weight[x] if weight else 1.0), -block_cost(x, [seqs[j][pos[j]] for j in present[x]])))
        items = [seqs[j][pos[j]] for j in present[k]]
        tot += COST["loop"] + vote + block_cost(k, items)
        for j in present[k]:
            pos[j] += 1
    return tot, iters


def main():
    warps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    ops, quads = load_program(), cornell_quads()
    res = {k: 0.0 for k in ("P0", "P1", "P1b", "P2", "P3", "P3b", "P4", "P4f", "P4w")}
    iters = {k: 0 for k in res}
    parts_tot, lane_ops, n_steps = {}, 0, 0
    lens = []
    for w in range(warps):
        tx, ty = int(rng.integers(0, 600 // 8)), int(rng.integers(0, 600 // 4))
        lanes = [Lane(tx * 8 + (l & 7), ty * 4 + (l >> 3)) for l in range(32)]
        for s in range(steps):
            seqs = [ln.step(ops, quads) for ln in lanes]
            if s < 8:
                continue                                   # let the lanes' depths decorrelate first
            n_steps += 1
            lane_ops += sum(len(q) for q in seqs)
            lens += [len(q) for q in seqs]
            c, n, parts = cost_p0(seqs)
            res["P0"] += c; iters["P0"] += n
            for k, v in parts.items():
                parts_tot[k] = parts_tot.get(k, 0) + v
            c, n = cost_sweep(seqs, (K_T, K_R, K_BVH, K_QUAD, K_RX, K_TX)); res["P1"] += c; iters["P1"] += n
            c, n = cost_sweep(seqs, (K_BVH, K_T, K_R, K_QUAD, K_RX, K_TX)); res["P1b"] += c; iters["P1b"] += n
            f = [fuse_wrappers(q) for q in seqs]
            c, n, _ = cost_p0(f); res["P2"] += c; iters["P2"] += n
            c, n = cost_sweep(f, (K_T, K_BVH, K_QUAD, K_RX)); res["P3"] += c; iters["P3"] += n
            c, n = cost_sweep(f, (K_BVH, K_T, K_QUAD, K_RX)); res["P3b"] += c; iters["P3b"] += n
            c, n = cost_majority(seqs); res["P4"] += c; iters["P4"] += n
            c, n = cost_majority(f); res["P4f"] += c; iters["P4f"] += n
            c, n = cost_majority(f, weight={K_BVH: 0.5, K_QUAD: 1.0, K_T: 0.7, K_R: 0.7, K_RX: 0.7, K_TX: 0.7}); res["P4w"] += c; iters["P4w"] += n
    print("warp-bounces %d, ops per lane %.1f (min %d, max %d)" % (n_steps, lane_ops / n_steps / 32.0, min(lens), max(lens)))
    names = {K_BVH: "box", K_QUAD: "planar", K_T: "translate", K_R: "rotate", K_RX: "un-rotate", K_TX: "un-translate", "loop": "loop"}
    print("P0 parts per warp-bounce:", {names[k]: round(v / n_steps, 1) for k, v in parts_tot.items()})
    for k in res:
        print("%-4s %7.1f warp-instr per warp-bounce (%.3f of P0), %.1f iterations" % (k, res[k] / n_steps, res[k] / res["P0"], iters[k] / n_steps))


if __name__ == "__main__":
    main()
