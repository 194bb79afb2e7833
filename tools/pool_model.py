#!/usr/bin/env python
"""pool_model.py -- executable model of the path kernel's warp-private pool scheduler (csrc/rtc_pool.cuh).

The kernel evaluates SAMPLES OF A PIXEL OUT OF ORDER.  In the reference every hit-bounce draws exactly 7 random
numbers and a miss draws none (raytracing.c:272-286, moremath.c:89-108), and every sample of a pixel re-traces the
same primary ray (main.c:94-99).  A sample's light is therefore a pure function of the RNG state it starts from, and
every reachable start state is the pixel's seed advanced by 7*j LCG steps: T[j] = (light, h_j), h_j = hit-bounces of
the sample that starts at offset j.  The pixel is the pointer walk  j <- j + h_j,  summed in sample order.

This file models ONE warp's pool (NS evaluation slots, NC pixel contexts) with the same steps, ownership rules and
invariants as the CUDA code -- consume (ordered walk over finished slots, cancelling skipped offsets), refill (new
pixels first, then one slot for every context with nothing outstanding, then speculative children), stage selection
(sweep when 64 rays are ready, else shade) -- against a serial walk over a synthetic T.  It checks that the
accumulated float sums, ray counts and final RNG states are identical and that the pool never stalls.

    python tools/pool_model.py            # randomised self-test
"""
from __future__ import annotations

import random
import sys

import numpy as np

FREE, SHADE, SWEEP, DONE = 0, 1, 2, 3
A7, C7 = None, None


def lcg(s):
    return (s * 747796405 + 2891336453) & 0xFFFFFFFF


def adv7(s):
    for _ in range(7):
        s = lcg(s)
    return s


class Script:
    """What evaluating (pixel, start state) does: a list of stages and the result (c, h, rays, final state)."""

    def __init__(self, state, max_bounce, p_hit, p_rr, known_primary=True):
        rnd = random.Random(state * 2654435761 & 0xFFFFFFFF)
        self.stages = []
        h, rays, s = 0, 0, state
        bounce = 0
        pending_hit = known_primary
        if not known_primary:  # the kernel sweeps the primary ray itself
            self.stages.append(SWEEP)
            rays += 1
            pending_hit = rnd.random() < p_hit
        while pending_hit:
            self.stages.append(SHADE)
            h += 1
            s = adv7(s)
            if rnd.random() < p_rr:
                break
            bounce += 1
            if bounce >= max_bounce:
                break
            self.stages.append(SWEEP)
            rays += 1
            pending_hit = rnd.random() < p_hit
        self.c = np.float32(rnd.random())
        self.h, self.rays, self.final = h, rays, s


def serial_walk(seed, spp, mk):
    acc, rays, s = np.float32(0), 0, seed
    for _ in range(spp):
        sc = mk(s)
        acc = np.float32(acc + sc.c)
        rays += sc.rays
        s = sc.final
    return acc, rays, s


class Pool:
    def __init__(self, NS, NC, pixels, spp, mk, spec, maxctx, rng):
        self.NS, self.NC, self.spp, self.mk, self.spec = NS, NC, spp, mk, spec
        self.pixels = list(pixels)  # seeds
        self.cursor = 0
        self.maxctx = maxctx
        self.rng = rng
        self.st = [FREE] * NS
        self.ctx = [0] * NS
        self.off = [0] * NS
        self.script = [None] * NS
        self.pc = [0] * NS
        self.live = [False] * NC
        self.rem = [0] * NC
        self.j = [0] * NC
        self.top = [0] * NC
        self.tstate = [0] * NC
        self.acc = [np.float32(0)] * NC
        self.rays = [0] * NC
        self.pix = [0] * NC
        self.out = {}
        self.evaluated = 0
        self.consumed = 0

    def consume(self):
        while True:
            progress = False
            for s in range(self.NS):
                if self.st[s] == FREE:
                    continue
                c = self.ctx[s]
                if not self.live[c] or self.off[s] < self.j[c]:
                    self.st[s] = FREE  # a skipped offset, or the pixel is finished
                    continue
                if self.st[s] == DONE and self.off[s] == self.j[c]:
                    sc = self.script[s]
                    self.acc[c] = np.float32(self.acc[c] + sc.c)
                    self.rays[c] += sc.rays
                    self.consumed += 1
                    nj = self.j[c] + (sc.h if self.spec else 1)
                    self.j[c] = nj
                    if nj >= self.top[c]:
                        self.top[c] = nj
                        self.tstate[c] = sc.final
                    self.rem[c] -= 1
                    self.st[s] = FREE
                    progress = True
                    if self.rem[c] == 0:
                        # final state: the state at offset j (== top here or derived)
                        self.out[self.pix[c]] = (self.acc[c], self.rays[c])
                        self.live[c] = False
            if not progress:
                break

    def refill(self):
        free = [s for s in range(self.NS) if self.st[s] == FREE]
        nlive = sum(self.live)
        # class 1 demand: live contexts with nothing outstanding
        d1 = [c for c in range(self.NC) if self.live[c] and self.top[c] == self.j[c]]
        room = len(free) - len(d1)
        freectx = [c for c in range(self.NC) if not self.live[c]]
        nnew = max(0, min(room, self.maxctx - nlive, len(freectx), len(self.pixels) - self.cursor))
        for i in range(nnew):
            c = freectx[i]
            seed = self.pixels[self.cursor]
            self.pix[c] = self.cursor
            self.cursor += 1
            self.live[c], self.rem[c], self.j[c], self.top[c], self.tstate[c] = True, self.spp, 0, 0, seed
            self.acc[c], self.rays[c] = np.float32(0), 0
            d1.append(c)
        grants = {}
        for c in d1[:len(free)]:
            grants[c] = 1
        left = len(free) - len(grants)
        if self.spec and left > 0:
            want = {c: self.j[c] + self.rem[c] - self.top[c] - grants.get(c, 0) for c in range(self.NC) if self.live[c]}
            want = {c: w for c, w in want.items() if w > 0}
            if want:
                q = -(-left // len(want))
                for c, w in want.items():
                    g = min(w, q, left)
                    if g <= 0:
                        break
                    grants[c] = grants.get(c, 0) + g
                    left -= g
        fi = 0
        for c, g in grants.items():
            for _ in range(g):
                s = free[fi]
                fi += 1
                self.ctx[s], self.off[s] = c, self.top[c]
                self.script[s] = self.mk(self.tstate[c])
                self.pc[s] = 0
                self.st[s] = self.script[s].stages[0] if self.script[s].stages else DONE
                self.evaluated += 1
                self.top[c] += 1
                self.tstate[c] = adv7(self.tstate[c]) if self.spec else self.tstate[c]

    def stage(self):
        nsh = [s for s in range(self.NS) if self.st[s] == SHADE]
        nsw = [s for s in range(self.NS) if self.st[s] == SWEEP]
        if len(nsw) >= 64 or (not nsh and nsw):
            pick, kind = nsw[:64], SWEEP
        elif nsh:
            pick, kind = nsh[:32], SHADE
        else:
            return False
        for s in pick:
            assert self.script[s].stages[self.pc[s]] == kind
            self.pc[s] += 1
            self.st[s] = self.script[s].stages[self.pc[s]] if self.pc[s] < len(self.script[s].stages) else DONE
        return True

    def run(self):
        iters = 0
        while True:
            iters += 1
            assert iters < 1000000, "pool stalled"
            self.consume()
            self.refill()
            if not self.stage():
                if any(self.live) or self.cursor < len(self.pixels):
                    # nothing to run although pixels are unfinished: only legal if a DONE slot was just produced
                    assert any(st == DONE for st in self.st), "deadlock"
                    continue
                break
        return iters


def selftest(trials=200, seed=1):
    rnd = random.Random(seed)
    worst_waste = 0.0
    for t in range(trials):
        NS = rnd.choice([32, 64, 96])
        NC = rnd.choice([1, 2, 8, 32, 64])
        NC = min(NC, NS)
        spp = rnd.choice([1, 2, 5, 16, 64])
        mb = rnd.choice([1, 2, 4, 10])
        p_hit = rnd.choice([0.0, 0.05, 0.3, 0.9])
        p_rr = rnd.choice([0.0, 0.14, 0.5])
        spec = rnd.random() < 0.7
        npix = rnd.choice([1, 3, 17, 200])
        maxctx = rnd.choice([1, 2, 8, NC])
        maxctx = min(maxctx, NC)
        cache = {}

        def mk(state, mb=mb, p_hit=p_hit, p_rr=p_rr, spec=spec, cache=cache):
            if state not in cache:
                cache[state] = Script(state, mb, p_hit, p_rr, known_primary=spec)
            return cache[state]

        pixels = [rnd.getrandbits(32) for _ in range(npix)]
        pool = Pool(NS, NC, pixels, spp, mk, spec, maxctx, rnd)
        pool.run()
        for i, sd in enumerate(pixels):
            want = serial_walk(sd, spp, mk)
            got = pool.out[i]
            assert got[0] == want[0] and got[1] == want[1], (t, i, got, want)
        waste = pool.evaluated / max(pool.consumed, 1) - 1.0
        worst_waste = max(worst_waste, waste)
    print(f"pool model: {trials} random configurations identical to the serial walk; worst speculative waste {worst_waste:.2f}x")


if __name__ == "__main__":
    selftest(int(sys.argv[1]) if len(sys.argv) > 1 else 200)
