#!/usr/bin/env python
"""Discrete-event model of the fused scoring kernel's pipeline (csrc/score_fused_f16.cu), fed with the constants the probes
measured, to see which of the measured periods a plain "agents + FIFO tensor queue + hand-over latency" picture can explain.

Agents: three MMA-issuing threads (layer 1, bottleneck, output layer), the ReLU warp group (hops 1a and 1b, polled in that
order), the squared-error warps, and the tensor pipe as a FIFO server of bounded depth.  Dependencies and ring depths are
the kernel's.  Everything is in SM cycles per (tile, model) pair.

  python scripts/score_pipeline_model.py            # the measured configurations side by side

Result (recorded in profiles/score_fused_f16_tuning_history.txt): with the probes' constants -- nominal MMA times, ~150 cycles
from a commit to its waiter, 145-225 cycles per ReLU hop, the squared-error warps' 348 + 408 cycles, 80 cycles of thread time
per commit -- this picture gives 760-870 cycles per pair for every variant, the tensor pipe 82-94 % busy.  The GPU runs them
all at 1,170-1,255.  The ~450 cycles per pair between the two are not in any term this model has.
"""
import argparse
import heapq
import itertools


class Sim:
    def __init__(self):
        self.t = 0.0
        self.q = []
        self.n = itertools.count()

    def at(self, t, fn):
        heapq.heappush(self.q, (t, next(self.n), fn))

    def run(self, until_fn):
        while self.q and not until_fn():
            self.t, _, fn = heapq.heappop(self.q)
            fn()


class Flag:
    """A monotone event counter: wait(k, fn) calls fn once at least k+1 events have been posted (visible)."""

    def __init__(self, sim):
        self.sim, self.count, self.waiters = sim, 0, []

    def post(self, delay=0.0):
        def vis():
            self.count += 1
            w, self.waiters = self.waiters, []
            for k, fn in w:
                self.wait(k, fn)
        self.sim.at(self.sim.t + delay, vis)

    def wait(self, k, fn):
        if self.count > k:
            fn()
        else:
            self.waiters.append((k, fn))


def simulate(n_pairs=400, nb=2, nd2=2, qcap=32, mma1=24 * 16.4, mma1b=4 * 16.4, mma2=4 * 64.5, c_commit=150.0, c_arrive=60.0,
             e1=225.0, e1b=200.0, poll=60.0, b_hold=348.0, b_rest=408.0, t_issue_mma=6.0, t_commit_thread=80.0, t_loop=60.0,
             split_a=False, hollow=False, tensor_only=False, throttle=0, commits=(1, 1, 1)):
    sim = Sim()
    if hollow:
        e1 = e1b = 60.0
        b_hold, b_rest = 40.0, 40.0
    # ---- tensor pipe: FIFO of (duration, callback) with bounded depth (entries) ----
    fifo, busy, space_waiters = [], [False], []
    pipe_busy_time = [0.0]

    def pipe_kick():
        if busy[0] or not fifo:
            return
        dur, cb = fifo.pop(0)
        busy[0] = True
        pipe_busy_time[0] += dur

        def done():
            busy[0] = False
            if cb:
                cb()
            w, space_waiters[:] = list(space_waiters), []
            for fn in w:
                fn()
            pipe_kick()
        sim.at(sim.t + dur, done)
        if space_waiters:
            w, space_waiters[:] = list(space_waiters), []
            for fn in w:
                fn()

    t_m2 = {}

    def push(items, then, per_item):
        """An issuing thread pushes items one by one (thread time per item), blocking while the queue is full."""
        items = list(items)

        def step():
            if not items:
                then()
                return
            if len(fifo) >= qcap:
                space_waiters.append(step)
                return
            fifo.append(items.pop(0))
            pipe_kick()
            sim.at(sim.t + per_item, step)
        step()

    m1, m1b, m2 = Flag(sim), Flag(sim), Flag(sim)         # "GEMM x of pair q retired" (visible after c_commit)
    h1, h2, d2free, thr = Flag(sim), Flag(sim), Flag(sim), Flag(sim)
    done_pairs = [0]
    t_first = {}

    def issuer(name, q, work, n_mma, out_flag, waits, n_commits):
        """waits: list of (flag, index) that must be visible before pair q is issued."""
        if q >= n_pairs:
            return
        pend = [w for w in waits(q) if w[1] >= 0 and not tensor_only_skip(name, w)]

        def ready():
            if name == "T1":
                t_first.setdefault(q, sim.t)
            per = work / n_mma
            chunks = [(per, None)] * n_mma
            if name == "T1" and throttle:
                # at most `throttle` chunks of 6 MMAs in the queue: wait for chunk c - throttle before issuing chunk c
                def issue_chunk(c):
                    if c == 4:
                        return finish()
                    idx = q * 4 + c

                    def go():
                        push([(per, None)] * 5 + [(per, lambda: thr.post(c_commit))], lambda: issue_chunk(c + 1), t_issue_mma)
                    if idx >= throttle:
                        thr.wait(idx - throttle, go)
                    else:
                        go()

                def finish():
                    push([(0.0, lambda: out_flag.post(c_commit))], lambda: sim.at(sim.t + t_commit_thread * n_commits + t_loop,
                                                                              lambda: issuer(name, q + 1, work, n_mma, out_flag, waits, n_commits)), 0.0)
                return issue_chunk(0)
            chunks[-1] = (per, None)
            push(chunks + [(0.0, lambda: out_flag.post(c_commit))],
                 lambda: sim.at(sim.t + t_commit_thread * n_commits + t_loop,
                                lambda: issuer(name, q + 1, work, n_mma, out_flag, waits, n_commits)), t_issue_mma)

        def wait_all(i=0):
            if i == len(pend):
                return ready()
            pend[i][0].wait(pend[i][1], lambda: wait_all(i + 1))
        wait_all()

    def tensor_only_skip(name, w):
        return tensor_only            # no epilogue dependencies at all

    issuer_T1 = lambda: issuer("T1", 0, mma1, 24, m1, lambda q: [(m1b, q - nb)], commits[0])
    issuer_T2 = lambda: issuer("T2", 0, mma1b, 4, m1b, lambda q: [(h1, q), (m2, q - nb)], commits[1])
    class M2(Flag):
        def post(self, delay=0.0):
            t_m2[len(t_m2)] = sim.t
            Flag.post(self, delay)
    m2 = M2(sim)
    issuer_T3 = lambda: issuer("T3", 0, mma2, 4, m2, lambda q: [(h2, q), (d2free, q - nd2)], commits[2])

    # ---- ReLU warps: one group polling 1a then 1b, or two groups with blocking waits ----
    if not tensor_only:
        if split_a:
            def hop(flag_in, flag_out, dur, q=0):
                if q >= n_pairs:
                    return
                flag_in.wait(q, lambda: sim.at(sim.t + dur, lambda: (flag_out.post(c_arrive), hop(flag_in, flag_out, dur, q + 1))))
            sim.at(0, lambda: hop(m1, h1, e1))
            sim.at(0, lambda: hop(m1b, h2, e1b))
        else:
            st = {"q1": 0, "q1b": 0}

            def a_loop():
                if st["q1b"] >= n_pairs:
                    return
                if st["q1"] < n_pairs and m1.count > st["q1"]:
                    st["q1"] += 1
                    sim.at(sim.t + e1 + poll, lambda: (h1.post(c_arrive), a_loop()))
                elif st["q1b"] < st["q1"] and m1b.count > st["q1b"]:
                    st["q1b"] += 1
                    sim.at(sim.t + e1b + poll, lambda: (h2.post(c_arrive), a_loop()))
                else:
                    sim.at(sim.t + poll, a_loop)
            sim.at(0, a_loop)

        def b_loop(q=0):
            if q >= n_pairs:
                return

            def got():
                def released():
                    d2free.post(c_arrive)

                    def fin():
                        done_pairs[0] += 1
                        b_loop(q + 1)
                    sim.at(sim.t + b_rest, fin)
                sim.at(sim.t + b_hold, released)
            m2.wait(q, got)
        sim.at(0, b_loop)
    sim.at(0, issuer_T1)
    sim.at(0, issuer_T2)
    sim.at(0, issuer_T3)
    target = n_pairs
    if tensor_only:
        sim.run(lambda: m2.count >= target and m1.count >= target)
    else:
        sim.run(lambda: done_pairs[0] >= target)
    lo, hi = n_pairs // 4, 3 * n_pairs // 4
    period = (t_m2[hi] - t_m2[lo]) / (hi - lo)          # pace of the output GEMM's completions
    return period, pipe_busy_time[0] / sim.t


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qcap", type=int, default=32)
    a = ap.parse_args()
    rows = [
        ("v8: rings of 2, one ReLU group, 8 commits", dict(commits=(2, 3, 3)), 1237),
        ("v13: one commit per group", dict(), 1215),
        ("hollow epilogues (v8)", dict(hollow=True, commits=(2, 3, 3)), 1170),
        ("v9: own group for hop 1b", dict(split_a=True, commits=(2, 3, 3)), 1240),
        ("v10: rings of 3 (D2 1.5 pairs)", dict(nb=3, nd2=1, commits=(2, 3, 3)), 1245),
        ("v12: layer-1 throttle, <= 2 chunks of 6", dict(throttle=2, commits=(2, 3, 3)), 1255),
        ("v14: rings of 3 + throttle, one commit", dict(nb=3, nd2=1, throttle=2), 1220),
    ]
    print(f"queue depth {a.qcap} entries; cycles per (tile, model) pair: model vs measured")
    for name, kw, meas in rows:
        p, util = simulate(qcap=a.qcap, **kw)
        print(f"  {name:46s} model {p:7.0f}  (tensor pipe {100 * util:3.0f} % busy)   measured {meas}")


if __name__ == "__main__":
    main()
