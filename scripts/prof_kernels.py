"""Stand-alone driver for ncu captures: runs one kernel family a few times on bench-shaped data.

  python scripts/prof_kernels.py train|score|round [--clients 125] [--iters 3]
"""
import argparse
import os
import random
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.data import NBAIOT_SPEC, ShardStore, make_clients, prepare_from_raw
from fedmse_decentralized_b200.ops.backend import RoundControl
from fedmse_decentralized_b200.trainer import FederationEngine

ap = argparse.ArgumentParser()
ap.add_argument("what", choices=["train", "score", "round"])
ap.add_argument("--clients", type=int, default=125)
ap.add_argument("--iters", type=int, default=3)
ap.add_argument("--epochs", type=int, default=5)
ap.add_argument("--patience", type=int, default=1)
ap.add_argument("--score-precision", default="fp32")
ap.add_argument("--phase-cycles", action="store_true")
ap.add_argument("--aggregation", default="elected", choices=["elected", "gossip"])
ap.add_argument("--participation", type=float, default=0.5)
ap.add_argument("--dev-scale", type=float, default=None, help="multiply the dev matrix (timing experiments: 0 = all zeros, 1024 = fp16 remainders out of the subnormal range)")
ap.add_argument("--dev-rows-mult", type=int, default=1,
                help="repeat the pooled dev matrix: 8 reproduces on one GPU the per-GPU scoring work of an 8-GPU run")
a = ap.parse_args()
dev = torch.device("cuda", 0)
random.seed(0); np.random.seed(0); torch.manual_seed(0)
clients, devset = prepare_from_raw(make_clients(a.clients, NBAIOT_SPEC, seed=1234), True, seed=1)
if a.dev_rows_mult > 1:
    devset = np.tile(devset, (a.dev_rows_mult, 1))
if a.dev_scale is not None:
    devset = (devset * a.dev_scale).astype(devset.dtype)
sh = ShardStore.build(clients, devset).to(dev)
cfg = ExperimentConfig(epoch=a.epochs, network_size=a.clients, backend="cuda", write_latents=False,
                       save_client_models=False, local_patience=a.patience, score_precision=a.score_precision,
                       aggregation=a.aggregation, topology="random-k", topology_k=4, num_participants=a.participation)
eng = FederationEngine(cfg, sh, "hybrid", "mse_avg")
be, st = eng.backend, eng.state
if a.phase_cycles:
    be.phase_cycles = torch.zeros(16, dtype=torch.int64, device=dev)
    be.mm_cycles = torch.zeros(2048, dtype=torch.int64, device=dev)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(a.iters):
    ctl = eng.make_control(it)
    be.begin_round(ctl)
    torch.cuda.synchronize()
    ev[0].record()
    if a.what == "train":
        be.train(st, sh, eng._sel_local(ctl.sel), eng.hp)
    elif a.what == "score":
        be.merge_scores(st, sh, ctl)
    else:
        eng.run_round(it, ctl)
    ev[1].record()
    torch.cuda.synchronize()
    extra = ""
    if a.what == "train":
        ep = st.epochs_run[st.epochs_run > 0]
        extra = f" epochs max {int(ep.max())} mean {float(ep.float().mean()):.2f}"
    print(f"{a.what} iter {it}: {ev[0].elapsed_time(ev[1]):.3f} ms{extra}")
    if a.phase_cycles and a.what == "train":
        names = ["wait-batch", "P1 fwd 115->27", "bottleneck fwd", "P4 out+loss", "B1 dgrad", "bottleneck bwd", "update", "loop/epoch"]
        cyc = be.phase_cycles.tolist()
        work = cyc[8:11]
        pre = cyc[11]
        cyc = cyc[:8]
        cyc[1] += work[0] + pre; cyc[3] += work[1]; cyc[4] += work[2]
        print(f"  P1 split: remainder prepass + barrier {pre/1e3:.0f}k")
        tot = sum(cyc)
        print("  phase cycles (CTA 0): " + ", ".join(f"{n} {c/1e3:.0f}k ({100*c/tot:.0f}%)" for n, c in zip(names, cyc)))
        print(f"  of which warp 0's own work (rest = waiting for other warps): P1 {work[0]/1e3:.0f}k, P4 {work[1]/1e3:.0f}k, B1 {work[2]/1e3:.0f}k")
    if a.phase_cycles and a.what == "score" and getattr(be, "fused_f16", False):
        cyc = be.mm_cycles.tolist()
        names = ["1:wait tile", "1:wait w1", "1:wait D1 free", "1:issue", "1b:wait h1", "1b:wait w32", "1b:wait D1b free", "1b:issue",
                 "2:wait h2", "2:wait w4", "2:wait D2 free", "2:issue", "A:park", "A:e1", "A:e1b", "A:idle",
                 "B:wait D2", "B:hold D2", "B:rest", "B:unit entry", "total", "A:e1 wait", "A:e1b wait"]
        tot = cyc[20] or 1
        print("  fused f16 kernel, CTA 0 (k cycles, % of the MMA1 thread's loop): " +
              ", ".join(f"{n} {c / 1e3:.0f}k ({100 * c / tot:.0f}%)" for n, c in zip(names, cyc[:23]) if c))
        TN, TB = 96, 128
        if len(cyc) >= TB + 13 * TN and cyc[TB + 40]:
            import numpy as _np
            tl = _np.array(cyc[TB:TB + 13 * TN], dtype=_np.int64).reshape(13, TN)
            names_ev = ["1 issue start", "1 issued", "A sees D1", "h1 ready", "1b woke", "1b issued", "A sees D1b", "h2 ready",
                        "2 woke", "2 issued", "B woke", "B freed D2", "B done"]
            q0, q1 = 16, 56                     # steady state, inside the first unit (62 pairs)
            print(f"  time line, CTA 0, pairs {q0}..{q1 - 1}: period {(tl[0, q1] - tl[0, q0]) / (q1 - q0):.0f} cycles; mean gap to the previous event of the same pair:")
            for e in range(1, 13):
                d = tl[e, q0:q1] - tl[e - 1, q0:q1]
                print(f"    {names_ev[e - 1]:>14} -> {names_ev[e]:<14} {d.mean():7.0f}  (min {d.min()}, max {d.max()})")
            print(f"    whole chain (1 issue start -> B done): {(tl[12, q0:q1] - tl[0, q0:q1]).mean():.0f} cycles = {(tl[12, q0:q1] - tl[0, q0:q1]).mean() / ((tl[0, q1] - tl[0, q0]) / (q1 - q0)):.2f} pairs in flight")
            for q in range(24, 28):
                print("    pair %d: " % q + " ".join("%d" % (tl[e, q] - tl[0, 24]) for e in range(13)))
        for k in range(4):
            w0, w1, p0, p1, e_last, xl = cyc[32 + 8 * k: 38 + 8 * k]
            if w0:
                print(f"  boundary {k + 1}: MMA1 thread waits {w1 - w0} cycles for the tile; relative to its arrival: last e1 of the unit "
                      f"{e_last - w0:+d}, parking starts {p0 - w0:+d}, ends {p1 - w0:+d}, TMA load of this tile was issued {xl - w0:+d}")
    elif a.phase_cycles and a.what == "score" and be.tc:
        names = ["issue MMA1(g+1)", "wait MMA2(g-1)", "wait epi1(g)", "wait w4 image", "wait D2 free", "issue MMA2(g)",
                 "wait X in TMEM", "wait X staging free",
                 "A: wait X smem", "A: X->TMEM", "A: wait MMA1", "A: epilogue 1",
                 "B: unit entry", "B: wait MMA2", "B: D2->regs", "B: rest"]
        cyc = be.mm_cycles.tolist()
        tot = sum(cyc[:8]) or 1
        print("  producer thread (CTA 0): " + ", ".join(f"{n} {c/1e3:.0f}k ({100*c/tot:.0f}%)" for n, c in zip(names[:8], cyc[:8])))
        print("  epilogue warps (CTA 0):  " + ", ".join(f"{n} {c/1e3:.0f}k" for n, c in zip(names[8:], cyc[8:])))
        for k in range(4):
            w0, w1, a0, a1 = cyc[16 + 4 * k: 20 + 4 * k]
            if w0:
                print(f"  boundary {k}: producer waits {w1 - w0} cycles; A reaches the boundary {a0 - w0:+d}, X in TMEM {a1 - w0:+d} (relative to the producer's arrival)")
