"""tools/tail_model.py -- list-scheduling model of the end-of-launch tail of the MPC kernel on config 2 (4096 robots, 296 CTA slots
on a B200: 2 per SM), from the per-robot ADMM iteration counts tools/iters_dump.py writes on the GPU box.  A robot's solve costs
(iterations + 35) units; CTAs pull robots from a queue in the given order.  Prints the makespan over the ideal (sum / slots) for
the index order, for longest-first orders built from what is known BEFORE the launch, and for the unattainable longest-first by
the true counts.  usage: python tools/tail_model.py iters.npz   (no GPU)"""
import heapq
import sys

import numpy as np

d = np.load(sys.argv[1])
its, speed = d["iters"].astype(float), d["speed"]
T, B = its.shape
SLOTS = 296


def makespan(times, order):
    h = [0.0] * SLOTS
    heapq.heapify(h)
    for i in order:
        heapq.heappush(h, heapq.heappop(h) + times[i])
    return max(h)


def ev(name, pred):
    out = []
    for t in range(4, T):
        cur = its[t] + 35
        out.append(makespan(cur, np.argsort(-pred(t), kind="stable")) / (cur.sum() / SLOTS) - 1)
    print(f"{name:50s}" + " ".join(f"{100 * x:5.1f}" for x in out) + f" | mean +{100 * np.mean(out):.1f} %")


print(f"{B} robots, ticks 4..{T - 1}: makespan over ideal, per tick [%]")
ev("index order", lambda t: -np.arange(float(B)))
ev("longest first by the previous tick's count", lambda t: its[t - 1])
ev("... by the max of the last 3 ticks", lambda t: its[t - 3:t].max(0))
ev("... previous + 5 x its rise", lambda t: its[t - 1] + 5 * np.maximum(0, its[t - 1] - its[t - 2]))
ev("... commanded speed only", lambda t: speed)
for g in (1000, 2000, 3000):
    ev(f"... max(previous, {g} x commanded speed)", lambda t: np.maximum(its[t - 1], g * speed))
ev("longest first by the TRUE count (unattainable)", lambda t: its[t])
c = [float(np.corrcoef(its[t - 1], its[t])[0, 1]) for t in range(3, T)]
j = [int(((its[t] > 1200) & (its[t - 1] < 700)).sum()) for t in range(3, T)]
print(f"correlation of consecutive ticks {min(c):.2f}..{max(c):.2f}; robots that jump from < 700 to > 1200 iterations per tick: {min(j)}..{max(j)}")
print("one such robot at the end of a longest-first queue costs its whole solve (~1,600 of ~8,500 units); they are the robots commanded to move fast")
