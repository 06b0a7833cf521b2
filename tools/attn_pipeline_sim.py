"""Discrete-time model of one SM running the split-row attention pipeline (csrc/attn_split.cu, csrc/attn_pipe.cu).

Not product code: a design aid.  It models, per clock, the in-order tensor pipe, the single MMA-issuing thread with
its blocking mbarrier waits, and the four softmax warps of one SM sub-partition (one from each of the four softmax
warpgroups; the sub-partitions are symmetric) sharing the MUFU budget.  It answers one question: how many clocks does
a 128-key tile (both query tiles) take under a given synchronisation structure, given the measured unit costs
(profiles/r01_attn_v5_full_ncu.txt: 2096 clk per tile pair, tensor pipe 48.8 % busy, 11.6 polls of s_full per tile).

    python tools/attn_pipeline_sim.py            # 'tile' (round-1 structure) vs 'quarter' (half-tile barriers)
"""
from __future__ import annotations

import argparse
from collections import deque


class Bar:
    """mbarrier with `count` arrivals per phase; a waiter sees a completed phase `wake` clocks after the last arrival."""

    def __init__(self, count, wake):
        self.count, self.wake, self.pending, self.done_at = count, wake, 0, []

    def arrive(self, t):
        self.pending += 1
        if self.pending == self.count:
            self.pending = 0
            self.done_at.append(t)

    def ready(self, phase, t):
        return phase < 0 or (phase < len(self.done_at) and t >= self.done_at[phase] + self.wake)


def simulate(mode, n_tiles=64, chunk_work=160.0, solo_clk=300.0, wake=80, ld_lat=45, st_lat=30, mma_issue=12, qk_clk=256,
             pv_clk=256, order="fixed"):
    """mode 'tile': S_w is one 128-column buffer, freed after the second 32-column load; one PV per tile.
    mode 'quarter': QK and PV are issued per 64-key half tile with their own barriers (S_w = two 64-column buffers).
    chunk_work: MUFU-pipe clocks a warp needs for 32 scores per thread (20 MUFU x 8 clk per warp instruction);
    solo_clk: duration of that chunk when the warp has the sub-partition to itself (dependent-issue bound)."""
    NW = 2
    sub = 2 if mode == "quarter" else 1
    mk = lambda c: [[Bar(c, wake) for _ in range(sub)] for _ in range(NW)]
    s_full, s_free, p_full, pv_done = mk(1), mk(2), mk(2), mk(1)
    tensor_q = deque()
    state = {"busy_until": 0, "busy": 0}
    completions = []  # (time, bar)

    def mma_thread():
        if mode == "tile":
            for w in range(NW):
                yield ("issue", qk_clk, s_full[w][0])
            for j in range(n_tiles):
                if j + 1 < n_tiles:
                    for w in range(NW):
                        yield ("wait", s_free[w][0], j)
                        yield ("issue", qk_clk, s_full[w][0])
                for w in range(NW):
                    yield ("wait", p_full[w][0], j)
                    yield ("issue", pv_clk, pv_done[w][0])
        else:
            for s in range(2):
                for w in range(NW):
                    yield ("issue", qk_clk // 2, s_full[w][s])
            for j in range(n_tiles):
                for s in range(2):
                    if j + 1 < n_tiles:
                        for w in range(NW):
                            yield ("wait", s_free[w][s], j)
                            yield ("issue", qk_clk // 2, s_full[w][s])
                    for w in range(NW):
                        yield ("wait", p_full[w][s], j)
                        yield ("issue", pv_clk // 2, pv_done[w][s])

    def softmax_warp(w, log):
        for j in range(n_tiles):
            if mode == "tile":
                yield ("wait", s_full[w][0], j)
                yield ("delay", ld_lat)
                yield ("compute", chunk_work)
                yield ("delay", ld_lat)
                yield ("arrive", s_free[w][0])
                yield ("compute", chunk_work)
                yield ("wait", pv_done[w][0], j - 1)
                yield ("delay", st_lat)
                yield ("arrive", p_full[w][0])
            else:
                for s in range(2):
                    yield ("wait", s_full[w][s], j)
                    yield ("delay", ld_lat)
                    yield ("arrive", s_free[w][s])
                    yield ("compute", chunk_work)
                    yield ("wait", pv_done[w][s], j - 1)
                    yield ("delay", st_lat)
                    yield ("arrive", p_full[w][s])
            log.append(None)

    class Actor:
        def __init__(self, gen):
            self.gen, self.cur, self.t_ready, self.work, self.wait_clk, self.done = gen, None, 0, 0.0, 0, False
            self.advance()

        def advance(self):
            try:
                self.cur = next(self.gen)
                if self.cur[0] == "compute":
                    self.work = self.cur[1]
            except StopIteration:
                self.done, self.cur = True, None

    tile_end = []
    logs = [[] for _ in range(4)]
    warps = [Actor(softmax_warp(g // 2, logs[g])) for g in range(4)]
    mma = Actor(mma_thread())
    t = 0
    n_logged = 0
    while not all(x.done for x in warps) and t < 20_000_000:
        if t >= state["busy_until"] and tensor_q:
            dur, bar = tensor_q.popleft()
            state["busy_until"] = t + dur
            state["busy"] += dur
            completions.append((t + dur, bar))
        for c in [c for c in completions if c[0] <= t]:
            c[1].arrive(t)
            completions.remove(c)
        while not mma.done and t >= mma.t_ready:
            op = mma.cur
            if op[0] == "wait":
                if not op[1].ready(op[2], t):
                    break
            else:
                tensor_q.append((op[1], op[2]))
                mma.t_ready = t + mma_issue
            mma.advance()
        active = [x for x in warps if not x.done and x.cur[0] == "compute" and t >= x.t_ready]
        if active:
            rate = min(chunk_work / solo_clk, 1.0 / len(active))
            for x in active:
                x.work -= rate
        for x in warps:
            while not x.done and t >= x.t_ready:
                op = x.cur
                if op[0] == "wait":
                    if not op[1].ready(op[2], t):
                        x.wait_clk += 1
                        break
                elif op[0] == "delay":
                    x.t_ready = t + op[1]
                elif op[0] == "arrive":
                    op[1].arrive(t)
                elif op[0] == "compute":
                    if x.work > 0:
                        break
                x.advance()
        if len(logs[0]) > n_logged:
            n_logged = len(logs[0])
            tile_end.append(t)
        t += 1
    lo, hi = n_tiles // 4, 3 * n_tiles // 4
    per_tile = (tile_end[hi] - tile_end[lo]) / (hi - lo)
    return {"mode": mode, "clk_per_tile": round(per_tile, 1), "tensor_busy_frac": round((qk_clk + pv_clk) * 2 / per_tile, 3),
            "softmax_wait_frac": round(sum(x.wait_clk for x in warps) / 4 / t, 3)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--solo", type=float, default=300.0)
    ap.add_argument("--wake", type=int, default=80)
    ap.add_argument("--work", type=float, default=160.0)
    a = ap.parse_args()
    for mode in ("tile", "quarter"):
        print(simulate(mode, solo_clk=a.solo, wake=a.wake, chunk_work=a.work))
