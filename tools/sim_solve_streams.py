"""Happens-before check of the stream / event schedule of the grid triangular solve
(dplasma_b200/csrc/potrf_dist.cu: solve_panels_multi).

The schedule issues work on three streams per rank -- compute (chain), update (bulk), communication (panel
prefetch + inverse blocks) -- ordered by events.  A missing wait does not fail a small GPU test reliably (it is a race),
so the issue order of the C++ routine is restated here operation by operation with the buffers each one reads and
writes, and every pair of conflicting accesses (same buffer, at least one write) is required to be ordered by
stream order or by a record -> wait edge.  Development aid and CPU test (tests/test_solve_schedule_cpu.py).
"""
import itertools


class Sched:
    def __init__(self):
        self.ops = []            # (stream, name, reads, writes)
        self.last = {}           # stream -> index of its last op (or None)
        self.deps = []           # per op: set of op indices it directly depends on
        self.pending = {}        # stream -> deps accumulated by waits since its last op
        self.events = {}         # event -> op index captured by the last record (None: nothing before it)

    def op(self, stream, name, reads=(), writes=()):
        i = len(self.ops)
        d = set(self.pending.pop(stream, ()))
        if self.last.get(stream) is not None:
            d.add(self.last[stream])
        self.ops.append((stream, name, frozenset(reads), frozenset(writes)))
        self.deps.append(d)
        self.last[stream] = i
        return i

    def record(self, ev, stream):
        # everything issued on the stream so far: its last op plus whatever that stream is itself waiting for
        cap = set(self.pending.get(stream, ()))
        if self.last.get(stream) is not None:
            cap.add(self.last[stream])
        self.events[ev] = cap

    def wait(self, stream, ev):
        self.pending.setdefault(stream, set()).update(self.events.get(ev, ()))

    def races(self):
        n = len(self.ops)
        before = [set() for _ in range(n)]   # transitive predecessors
        for i in range(n):                   # ops are appended in an order compatible with the dependencies
            for d in self.deps[i]:
                before[i].add(d)
                before[i] |= before[d]
        out = []
        for i, j in itertools.combinations(range(n), 2):
            si, ni, ri, wi = self.ops[i]
            sj, nj, rj, wj = self.ops[j]
            conflict = (wi & (rj | wj)) | (wj & ri)
            if conflict and i not in before[j] and j not in before[i]:
                out.append((ni, nj, sorted(conflict)))
        return out


def first_local_ge(g, me, n):
    return 0 if g <= me else (g - me + n - 1) // n


def solve_schedule(P, Q, prow, pcol, mt, lower, rhs_cols, drop=()):
    """the issue order of solve_panels_multi on rank (prow, pcol); rhs_cols = local RHS tile columns (colEndB).
    drop: waits of the schedule to leave out ("xrow_reuse", "lookahead_after_bulk", "panel_free", "x_ready") -- the test
    uses it to show that the check sees the race each wait is there to prevent"""
    S = Sched()
    s, sc, su = "compute", "comm", "update"
    lmt = first_local_ge(mt, prow, P)
    row_end = lmt
    S.op(s, "caller's work before the solve", writes=["Tm"] + [f"B{r}" for r in range(lmt)])
    S.record("start", s)
    S.wait(sc, "start")
    S.wait(su, "start")

    def rows_of(k):
        a = min(first_local_ge(k, prow, P), row_end) if lower else 0
        b = row_end if lower else min(first_local_ge(k + 1, prow, P), row_end)
        return a, b - a

    def post_panel(kk):
        k = kk if lower else mt - 1 - kk
        st = kk & 1
        kq = k % Q
        _, cnt = rows_of(k)
        if kk >= 2 and "panel_free" not in drop:
            S.wait(sc, f"free{st}")
        owner = pcol == kq
        if Q > 1 and cnt > 0:
            S.op(sc, f"row bcast {kk}", reads=["Tm"] if owner else [], writes=[] if owner else [f"panel{st}"])
        if prow == k % P:
            src = "Tm" if (Q == 1 or owner) else f"panel{st}"
            S.op(sc, f"inverse blocks {kk}", reads=[src], writes=[f"winv{st}", f"wbig{st}", f"y{st}"])
        S.record(f"recv{st}", sc)

    post_panel(0)
    for kk in range(mt):
        k = kk if lower else mt - 1 - kk
        st = kk & 1
        kp, kq = k % P, k % Q
        if kk + 1 < mt:
            post_panel(kk + 1)
        S.wait(s, f"recv{st}")
        panel = "Tm" if (Q == 1 or pcol == kq) else f"panel{st}"
        ilk = k // P
        if prow == kp and rhs_cols > 0:
            S.op(s, f"solve row {kk}", reads=[panel, f"wbig{st}", f"winv{st}", f"B{ilk}"], writes=[f"B{ilk}", "xt"])
        if kk >= 2 and "xrow_reuse" not in drop:
            S.wait(s, f"bulk{st}")
        if prow == kp and rhs_cols > 0:
            S.op(s, f"transpose row {kk}", reads=[f"B{ilk}"], writes=[f"xrow{st}"])
        if P > 1 and rhs_cols > 0:
            root = prow == kp
            S.op(s, f"col bcast {kk}", reads=[f"xrow{st}"] if root else [], writes=[] if root else [f"xrow{st}"])
        S.record(f"x{st}", s)
        u0 = min(first_local_ge(k + 1, prow, P), row_end) if lower else 0
        u1 = row_end if lower else min(first_local_ge(k, prow, P), row_end)
        kn = k + 1 if lower else k - 1
        next_here = kk + 1 < mt and kn % P == prow and u1 > u0
        la0, la1, b0, b1 = u0, u0, u0, u1
        if next_here:
            if lower:
                la0, la1, b0 = u0, u0 + 1, u0 + 1
            else:
                la0, la1, b1 = u1 - 1, u1, u1 - 1
        if la1 > la0:
            if kk >= 1 and "lookahead_after_bulk" not in drop:
                S.wait(s, f"bulk{st ^ 1}")
            if rhs_cols > 0:
                S.op(s, f"look-ahead update {kk}", reads=[panel, f"xrow{st}"] + [f"B{r}" for r in range(la0, la1)],
                     writes=[f"B{r}" for r in range(la0, la1)])
        S.record(f"step{st}", s)
        if "x_ready" not in drop:
            S.wait(su, f"x{st}")
        if b1 > b0 and rhs_cols > 0:
            S.op(su, f"bulk update {kk}", reads=[panel, f"xrow{st}"] + [f"B{r}" for r in range(b0, b1)],
                 writes=[f"B{r}" for r in range(b0, b1)])
        S.record(f"bulk{st}", su)
        S.wait(su, f"step{st}")
        S.record(f"free{st}", su)
    for e in range(2):
        S.wait(s, f"bulk{e}")
    S.record("done", sc)
    S.wait(s, "done")
    S.op(s, "caller's work after the solve", reads=["Tm"] + [f"B{r}" for r in range(lmt)],
         writes=[f"B{r}" for r in range(lmt)] + [f"panel{e}" for e in range(2)] + [f"xrow{e}" for e in range(2)] +
                [f"wbig{e}" for e in range(2)] + ["xt"])
    return S


if __name__ == "__main__":
    bad = 0
    for (P, Q), mt, lower, cols in itertools.product([(1, 1), (1, 2), (2, 1), (2, 2), (2, 4), (3, 2)], [1, 2, 3, 7, 12],
                                                     [True, False], [0, 1, 2]):
        for prow in range(P):
            for pcol in range(Q):
                r = solve_schedule(P, Q, prow, pcol, mt, lower, cols).races()
                if r:
                    bad += 1
                    print(f"P={P} Q={Q} rank=({prow},{pcol}) mt={mt} lower={lower} cols={cols}: {len(r)} unordered pairs, e.g. {r[0]}")
    print("unordered conflicting pairs found in", bad, "configurations")
