"""Model check of the fused kernel's halo protocol (DESIGN.md "Halo protocol / hazards").

compute-sanitizer is refused on this GPU pool, so the race argument of the design is checked on a
model instead: ranks, tiles, CTAs, per-direction epoch counters and double-buffered ghost regions
are simulated at the granularity the kernel works at (one tile = wait on flags, read ghost slices,
push edge slices into the neighbours' ghost regions of the OTHER buffer, signal), and a random
scheduler interleaves the CTAs of all ranks.  Every ghost slice carries the time level it holds;
the checker asserts

  RAW  a tile of step s only ever reads ghost slices that hold level s, and
  WAR  a slice is only overwritten after every tile that had to read its previous content has
       done so,

and that no schedule deadlocks.  The same checker must FIND a violation when the protocol is
broken in the ways the design says would be wrong (wait one epoch short, signal before the push,
single-buffered ghosts, corners not waited for).  The reference has no race detection at all, only
one deliberately racy kernel (ref_parallel-dot-product-atomics.cu:26-32).
"""
from __future__ import annotations

import random

import pytest

DIRS = ["NW", "N", "NE", "W", "E", "SW", "S", "SE"]          # order of MPIGridCellID
OFF = {"NW": (-1, -1), "N": (-1, 0), "NE": (-1, 1), "W": (0, -1), "E": (0, 1), "SW": (1, -1), "S": (1, 0), "SE": (1, 1)}
OPP = {d: DIRS[7 - i] for i, d in enumerate(DIRS)}


class Violation(Exception):
    pass


def decode_order(nbx, nby):
    """Boundary ring first, interior x-fastest (decode_tile in stencil_kernel.cuh)."""
    if nby == 1:
        return [(x, 0) for x in range(nbx)]
    out = [(x, 0) for x in range(nbx)] + [(x, nby - 1) for x in range(nbx)]
    out += [(0, y) for y in range(1, nby - 1)]
    if nbx > 1:
        out += [(nbx - 1, y) for y in range(1, nby - 1)]
    out += [(x, y) for y in range(1, nby - 1) for x in range(1, nbx - 1)]
    return out


def edge_dirs(tx, ty, nbx, nby):
    n, s, w, e = ty == 0, ty == nby - 1, tx == 0, tx == nbx - 1
    d = set()
    if n: d.add("N")
    if s: d.add("S")
    if w: d.add("W")
    if e: d.add("E")
    if n and w: d.add("NW")
    if n and e: d.add("NE")
    if s and w: d.add("SW")
    if s and e: d.add("SE")
    return d


class Rank:
    def __init__(self, rid, nbx, nby):
        self.id, self.nbx, self.nby = rid, nbx, nby
        self.flags = {d: 0 for d in DIRS}
        # ghost[buffer][(region, index)] = {"level": int, "readers": set of tiles that read this level}
        self.ghost = [dict(), dict()]
        for b in (0, 1):
            for x in range(nbx):
                self.ghost[b][("N", x)] = {"level": 0 if b == 0 else -1, "readers": set()}
                self.ghost[b][("S", x)] = {"level": 0 if b == 0 else -1, "readers": set()}
            for y in range(nby):
                self.ghost[b][("W", y)] = {"level": 0 if b == 0 else -1, "readers": set()}
                self.ghost[b][("E", y)] = {"level": 0 if b == 0 else -1, "readers": set()}
            for c in ("NW", "NE", "SW", "SE"):
                self.ghost[b][(c, 0)] = {"level": 0 if b == 0 else -1, "readers": set()}

    def slices_read_by(self, tx, ty):
        """Ghost slices inside the TMA box of tile (tx, ty): the box reaches one halo into the
        x/y neighbours, so an edge tile also reads the slices next to its own."""
        nbx, nby = self.nbx, self.nby
        out = []
        if ty == 0:
            out += [("N", x) for x in (tx - 1, tx, tx + 1) if 0 <= x < nbx]
            if tx == 0: out.append(("NW", 0))
            if tx == nbx - 1: out.append(("NE", 0))
        if ty == nby - 1:
            out += [("S", x) for x in (tx - 1, tx, tx + 1) if 0 <= x < nbx]
            if tx == 0: out.append(("SW", 0))
            if tx == nbx - 1: out.append(("SE", 0))
        if tx == 0:
            out += [("W", y) for y in (ty - 1, ty, ty + 1) if 0 <= y < nby]
        if tx == nbx - 1:
            out += [("E", y) for y in (ty - 1, ty, ty + 1) if 0 <= y < nby]
        return out

    def required_readers(self, region, idx):
        """Tiles whose box covers the slice (they must all have read it before it is overwritten)."""
        return {(tx, ty) for (tx, ty) in decode_order(self.nbx, self.nby)
                if (region, idx) in self.slices_read_by(tx, ty)}


class Sim:
    def __init__(self, py, px, nbx, nby, ctas, steps, rng, corners=True, bug=None):
        self.py, self.px, self.nbx, self.nby, self.ctas, self.steps = py, px, nbx, nby, ctas, steps
        self.rng, self.corners, self.bug = rng, corners, bug
        self.ranks = [Rank(r, nbx, nby) for r in range(py * px)]
        self.order = decode_order(nbx, nby)
        # per rank: current step and the per-CTA cursor (index into its tile list, phase)
        self.step = [0] * len(self.ranks)
        self.cta = [self._fresh_ctas() for _ in self.ranks]

    def _fresh_ctas(self):
        lists = [self.order[c::self.ctas] for c in range(self.ctas)]
        return [{"tiles": l, "i": 0, "phase": 0} for l in lists]

    def neighbour(self, r, d):
        y, x = divmod(r, self.px)
        dy, dx = OFF[d]
        return ((y + dy) % self.py) * self.px + (x + dx) % self.px      # periodic

    def contrib(self, d):
        return 1 if len(d) == 2 else (self.nbx if d in "NS" else self.nby)

    def active(self, tx, ty):
        return {d for d in edge_dirs(tx, ty, self.nbx, self.nby) if self.corners or len(d) == 1}

    def runnable(self):
        out = []
        for r, rk in enumerate(self.ranks):
            if self.step[r] >= self.steps:
                continue
            for c, st in enumerate(self.cta[r]):
                if st["i"] >= len(st["tiles"]):
                    continue
                if st["phase"] == 0 and not self._flags_ready(r, st):
                    continue
                out.append((r, c))
        return out

    def _flags_ready(self, r, st):
        tx, ty = st["tiles"][st["i"]]
        s = self.step[r]
        need = s - 1 if self.bug == "wait_one_short" else s
        for d in self.active(tx, ty):
            if self.bug == "skip_corner_wait" and len(d) == 2:
                continue
            if self.ranks[r].flags[d] < need * self.contrib(d):
                return False
        return True

    def advance(self, r, c):
        rk, st, s = self.ranks[r], self.cta[r][c], self.step[r]
        tx, ty = st["tiles"][st["i"]]
        act = self.active(tx, ty)
        inb = 0 if self.bug == "single_buffer" else s % 2
        outb = 0 if self.bug == "single_buffer" else (s + 1) % 2
        phase = st["phase"]
        if phase == 0:                       # flags acquired (checked by runnable())
            st["phase"] = 1
        elif phase == 1:                     # TMA load of the box: reads ghost slices of the input buffer
            for key in rk.slices_read_by(tx, ty):
                if not self.corners and len(key[0]) == 2:
                    continue                 # star stencil: corner cells are loaded but never used
                g = rk.ghost[inb][key]
                if g["level"] != s:
                    raise Violation(f"RAW: rank {r} step {s} tile {(tx, ty)} read {key} at level {g['level']}")
                g["readers"].add((tx, ty))
            st["phase"] = 2
        elif phase in (2, 3):
            # phase 2 = push, phase 3 = signal -- swapped in the "signal_before_push" variant
            do_push = (phase == 2) != (self.bug == "signal_before_push")
            if do_push:                      # my edges -> the neighbours' ghosts (other buffer)
                for d in act:
                    nb = self.ranks[self.neighbour(r, d)]
                    region = OPP[d]
                    idx = 0 if len(d) == 2 else (tx if d in "NS" else ty)
                    g = nb.ghost[outb][(region, idx)]
                    if g["level"] >= 0 and g["level"] != s + 1:
                        missing = nb.required_readers(region, idx) - g["readers"]
                        if not self.corners and len(region) == 2:
                            missing = set()
                        if missing:
                            raise Violation(f"WAR: rank {r} step {s} overwrote {region, idx} of rank {nb.id} "
                                            f"(level {g['level']}) before tiles {sorted(missing)} read it")
                    g["level"] = s + 1
                    g["readers"] = set()
            else:                            # release-increment the neighbours' counters
                self._signal(r, act)
            if phase == 2:
                st["phase"] = 3
                return
            st["phase"] = 0
            st["i"] += 1
            if all(x["i"] >= len(x["tiles"]) for x in self.cta[r]):      # kernel of this step finished
                self.step[r] += 1
                self.cta[r] = self._fresh_ctas()

    def _signal(self, r, act):
        for d in act:
            self.ranks[self.neighbour(r, d)].flags[OPP[d]] += 1

    def run(self):
        while any(s < self.steps for s in self.step):
            ready = self.runnable()
            if not ready:
                raise Violation(f"deadlock at steps {self.step}")
            self.advance(*self.rng.choice(ready))
        # skew between ranks never exceeded one step by construction of the waits; final state:
        for rk in self.ranks:
            for d in DIRS:
                assert rk.flags[d] == self.steps * self.contrib(d) * (1 if (self.corners or len(d) == 1) else 0)


GRIDS = [(1, 1, 3, 2), (1, 2, 3, 3), (2, 2, 2, 2), (2, 3, 4, 1), (3, 3, 1, 1), (2, 1, 5, 3)]


@pytest.mark.parametrize("py,px,nbx,nby", GRIDS)
@pytest.mark.parametrize("corners", [True, False])
def test_protocol_is_hazard_free_under_random_schedules(py, px, nbx, nby, corners):
    for seed in range(40):
        rng = random.Random(seed * 7919 + py * 100 + px * 10 + nbx)
        Sim(py, px, nbx, nby, ctas=rng.choice([1, 2, 3, 5]), steps=5, rng=rng, corners=corners).run()


@pytest.mark.parametrize("bug", ["wait_one_short", "signal_before_push", "single_buffer", "skip_corner_wait"])
def test_checker_finds_the_hazard_when_the_protocol_is_broken(bug):
    found = 0
    for seed in range(200):
        rng = random.Random(seed)
        try:
            Sim(2, 2, 3, 3, ctas=3, steps=4, rng=rng, corners=True, bug=bug).run()
        except Violation:
            found += 1
    assert found > 0, f"the model checker missed the '{bug}' hazard in 200 random schedules"


# ------------------------------------------------------------------------------------------------
# Fused dot + allreduce: {value, epoch} slots, two parities.
class DotSim:
    """Each rank, per epoch e: write {e} into slot[e % parities][me] of EVERY rank, then wait until all
    P slots of that parity in its OWN table carry epoch e, then read them one by one.  A rank starts
    epoch e+1 only when its epoch-e kernel has finished (stream order)."""

    def __init__(self, nranks, epochs, rng, parities=2):
        self.P, self.E, self.rng, self.par = nranks, epochs, rng, parities
        self.table = [[[0] * nranks for _ in range(parities)] for _ in range(nranks)]   # [rank][parity][src] = epoch
        self.state = [{"e": 1, "phase": "write", "k": 0} for _ in range(nranks)]

    def run(self):
        while any(s["e"] <= self.E for s in self.state):
            ready = []
            for r, s in enumerate(self.state):
                if s["e"] > self.E:
                    continue
                if s["phase"] == "wait" and any(v < s["e"] for v in self.table[r][s["e"] % self.par]):
                    continue
                ready.append(r)
            if not ready:
                raise Violation("dot: deadlock")
            r = self.rng.choice(ready)
            s = self.state[r]
            e, p = s["e"], s["e"] % self.par
            if s["phase"] == "write":                  # one peer store per scheduling step
                self.table[s["k"]][p][r] = e
                s["k"] += 1
                if s["k"] == self.P:
                    s["phase"], s["k"] = "wait", 0
            elif s["phase"] == "wait":
                s["phase"] = "read"
            else:                                       # read slot k of my table
                if self.table[r][p][s["k"]] != e:
                    raise Violation(f"dot: rank {r} epoch {e} read slot {s['k']} holding epoch "
                                    f"{self.table[r][p][s['k']]}")
                s["k"] += 1
                if s["k"] == self.P:
                    s["e"], s["phase"], s["k"] = e + 1, "write", 0


@pytest.mark.parametrize("nranks", [1, 2, 3, 8])
def test_dot_slot_protocol_two_parities_are_enough(nranks):
    for seed in range(100):
        DotSim(nranks, epochs=6, rng=random.Random(seed)).run()


def test_dot_slot_protocol_one_parity_is_not_enough():
    found = 0
    for seed in range(200):
        try:
            DotSim(3, epochs=6, rng=random.Random(seed), parities=1).run()
        except Violation:
            found += 1
    assert found > 0


# ------------------------------------------------------------------------------------------------
# Rooted reduce / mixed sequences: only the readers of an epoch wait, so the pushers can run ahead.
# The kernel's rule (csrc/kernels/dot.cu): a reader publishes ack[reader] = e in every table once it
# has read its slots of epoch e; a pusher that targets rank t in epoch e first waits for
# ack[t] >= (last earlier epoch of the same parity in which t read).  The collectives
# (csrc/collectives/device_coll.cu) use the coarser form of the same idea: "consumed" epochs.
class RootedDotSim:
    """roots[e-1] is the root of epoch e (None: allreduce).  Pushers: everybody; readers: root / all."""

    def __init__(self, nranks, roots, rng, with_ack=True):
        self.P, self.roots, self.rng, self.with_ack = nranks, roots, rng, with_ack
        self.table = [[[0] * nranks for _ in range(2)] for _ in range(nranks)]    # [rank][parity][src]
        self.ack = [[0] * nranks for _ in range(nranks)]                           # [table of][reader]
        self.state = [{"e": 1, "phase": "write", "k": 0} for _ in range(nranks)]
        # host-side bookkeeping, identical on every rank: last epoch of each parity in which r read
        self.need = []
        last = [[0, 0] for _ in range(nranks)]
        for e, root in enumerate(roots, start=1):
            self.need.append([last[t][e & 1] for t in range(nranks)])
            for t in range(nranks):
                if root is None or t == root:
                    last[t][e & 1] = e

    def targets(self, e):
        root = self.roots[e - 1]
        return list(range(self.P)) if root is None else [root]

    def reads(self, r, e):
        root = self.roots[e - 1]
        return root is None or r == root

    def run(self):
        E = len(self.roots)
        while any(s["e"] <= E for s in self.state):
            ready = []
            for r, s in enumerate(self.state):
                if s["e"] > E:
                    continue
                e = s["e"]
                if s["phase"] == "write" and self.with_ack:
                    t = self.targets(e)[s["k"]]
                    if self.ack[r][t] < self.need[e - 1][t]:
                        continue
                if s["phase"] == "wait" and any(v < e for v in self.table[r][e & 1]):
                    continue
                ready.append(r)
            if not ready:
                raise Violation("rooted dot: deadlock")
            r = self.rng.choice(ready)
            s = self.state[r]
            e, p = s["e"], s["e"] & 1
            if s["phase"] == "write":
                tg = self.targets(e)
                self.table[tg[s["k"]]][p][r] = e
                s["k"] += 1
                if s["k"] == len(tg):
                    s["k"] = 0
                    if self.reads(r, e):
                        s["phase"] = "wait"
                    else:
                        s["e"] = e + 1
            elif s["phase"] == "wait":
                s["phase"] = "read"
            elif s["phase"] == "read":
                if self.table[r][p][s["k"]] != e:
                    raise Violation(f"rooted dot: rank {r} epoch {e} read slot {s['k']} holding epoch "
                                    f"{self.table[r][p][s['k']]}")
                s["k"] += 1
                if s["k"] == self.P:
                    s["phase"], s["k"] = "ack", 0
            else:                                       # publish ack[r] = e, one table per step
                self.ack[s["k"]][r] = e
                s["k"] += 1
                if s["k"] == self.P:
                    s["e"], s["phase"], s["k"] = e + 1, "write", 0


def _root_sequences(nranks, rng):
    yield [0] * 8                                                  # always the same root
    yield [e % nranks for e in range(9)]                           # rotating root
    yield [rng.choice([None] + list(range(nranks))) for _ in range(10)]   # reduce and allreduce mixed


@pytest.mark.parametrize("nranks", [2, 3, 8])
def test_rooted_reduce_with_acknowledgements_is_hazard_free(nranks):
    for seed in range(60):
        rng = random.Random(seed)
        for roots in _root_sequences(nranks, rng):
            RootedDotSim(nranks, roots, rng).run()


def test_rooted_reduce_without_acknowledgements_is_broken():
    """What ADVICE.md (round 1) found: two parities are only enough when EVERY rank reads."""
    found = 0
    for seed in range(200):
        try:
            RootedDotSim(3, [0] * 8, random.Random(seed), with_ack=False).run()
        except Violation:
            found += 1
    assert found > 0


# ------------------------------------------------------------------------------------------------
# Tail of a sweep inside the temporal strip kernel (csrc/stencil/stencil_tstrip_kernel.cuh, end of
# the consumer path): every CTA writes its residual partial, fences, takes a ticket from a counter
# that wraps at the grid size; the CTA that draws the last ticket sums ALL partials and advances the
# device step counter.  The producers of all CTAs read that step counter during the sweep (epoch of
# the halo flags), a consumer can only finish after its producer's last load.  Sweeps follow each
# other in stream order (a new grid starts when the old one has ended).
class SweepTailSim:
    def __init__(self, ncta, nreads, sweeps, rng, ticket_before_partial=False, no_wrap=False):
        self.G, self.R, self.S, self.rng = ncta, nreads, sweeps, rng
        # consumer program of every CTA ("tail" is skipped by all but the drawer of the last ticket)
        self.program = ["ticket", "tail", "partial"] if ticket_before_partial else ["partial", "ticket", "tail"]
        self.no_wrap = no_wrap
        self.ctr = 0                    # device step counter
        self.ticket = 0
        self.partials = [None] * ncta   # (sweep, value)

    def run(self):
        for sweep in range(self.S):
            start_ctr = self.ctr
            prod = [0] * self.G          # counter reads done by the producer of each CTA
            pc = [0] * self.G            # position of each consumer in self.program
            last = [False] * self.G
            result = None
            while any(p < self.R for p in prod) or any(c < len(self.program) for c in pc):
                ready = [("p", i) for i in range(self.G) if prod[i] < self.R]
                # a consumer reaches the end of its rows only after its producer issued every load
                ready += [("c", i) for i in range(self.G) if pc[i] < len(self.program) and prod[i] == self.R]
                kind, i = self.rng.choice(ready)
                if kind == "p":
                    if self.ctr != start_ctr:
                        raise Violation(f"tail: producer of CTA {i} read the step counter after the bump")
                    prod[i] += 1
                    continue
                op = self.program[pc[i]]
                pc[i] += 1
                if op == "partial":
                    self.partials[i] = (sweep, i + 1)
                elif op == "ticket":
                    t = self.ticket
                    self.ticket = 0 if (t + 1 == self.G and not self.no_wrap) else t + 1
                    last[i] = t == self.G - 1
                elif last[i]:            # "tail": sum every partial, advance the counter
                    for j in range(self.G):
                        if self.partials[j] is None or self.partials[j][0] != sweep:
                            raise Violation(f"tail: sweep {sweep} summed a stale partial of CTA {j}")
                    result = sum(v for _, v in self.partials)
                    self.ctr += 1
            if result is None:
                raise Violation(f"tail: nobody drew the last ticket in sweep {sweep}")
            if result != self.G * (self.G + 1) // 2:
                raise Violation("tail: wrong sum")
        if self.ctr != self.S:
            raise Violation("tail: step counter")


@pytest.mark.parametrize("ncta", [1, 2, 5, 16])
def test_sweep_tail_last_cta_sums_and_bumps_safely(ncta):
    for seed in range(100):
        SweepTailSim(ncta, nreads=3, sweeps=4, rng=random.Random(seed)).run()


@pytest.mark.parametrize("bug", ["ticket_before_partial", "no_wrap"])
def test_sweep_tail_checker_finds_broken_variants(bug):
    found = 0
    for seed in range(200):
        try:
            SweepTailSim(5, nreads=2, sweeps=3, rng=random.Random(seed), **{bug: True}).run()
        except Violation:
            found += 1
    assert found > 0
