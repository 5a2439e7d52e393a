"""Discrete-event model of the paired backward sweep's pipeline (tensor pipe in-order, blocking issue thread).
Used to reason about schedules; not part of the product."""
import sys

def sim(T=64, S=3072, G=1536, E1=2500, E2=2500, C=600, order="cur", sbuf=1):
    INF = 1e18
    # per CTA c: list of ops in program order
    def program(c):
        ops = []
        if order == "cur":
            ops.append(("S", c))
            for t in range(T):
                if t % 2 == c and t + 2 < T: ops.append(("S", t + 2))
                ops.append(("G", t))
        elif order == "g_first":
            ops.append(("S", c))
            for t in range(T):
                ops.append(("G", t))
                if t % 2 != c and t + 1 < T and (t + 1) % 2 == c and t + 1 >= 2: ops.append(("S", t + 1))
        return ops
    progs = [program(0), program(1)]
    pc = [0, 0]
    pipe_free = [0.0, 0.0]          # when tensor pipe of CTA c is free
    s_done = {}                     # S(t) completion time
    g_done = [dict(), dict()]       # G(t) completion in CTA c
    epi_start, s_free, g_ready_local, g_ready_peer = {}, {}, {}, {}
    epi_free = [0.0, 0.0]           # epilogue warps availability
    issue_time = [0.0, 0.0]
    busy = [0.0, 0.0]
    progress = True
    # iterate until all ops are scheduled; resolve dependencies lazily
    def try_epi(t):
        c = t % 2
        if t in epi_start or t not in s_done: return
        need = s_done[t]
        if t - 2 >= 0:
            if (t - 2) not in g_done[0] or (t - 2) not in g_done[1]: return
            need = max(need, g_done[0][t - 2], g_done[1][t - 2])
        if t - 2 >= 0 and (t - 2) not in epi_start: return
        need = max(need, epi_free[c])
        epi_start[t] = need
        s_free[t] = need + E1
        g_ready_local[t] = need + E1 + E2
        g_ready_peer[t] = need + E1 + E2 + C
        epi_free[c] = need + E1 + E2
    while progress:
        progress = False
        for c in (0, 1):
            while pc[c] < len(progs[c]):
                kind, t = progs[c][pc[c]]
                if kind == "S":
                    # needs the S buffer: epilogue of this CTA's previous duty tile has read it
                    prev = t - 2 * sbuf
                    if prev >= 0:
                        try_epi(prev)
                        if prev not in s_free: break
                        ready = s_free[prev]
                    else:
                        ready = 0.0
                    start = max(ready, issue_time[c], pipe_free[c])
                    issue_time[c] = max(ready, issue_time[c])
                    pipe_free[c] = start + S
                    busy[c] += S
                    s_done[t] = start + S
                    try_epi(t)
                else:
                    try_epi(t)
                    if t not in g_ready_local: break
                    ready = g_ready_local[t] if t % 2 == c else g_ready_peer[t]
                    start = max(ready, issue_time[c], pipe_free[c])
                    issue_time[c] = max(ready, issue_time[c])
                    pipe_free[c] = start + G
                    busy[c] += G
                    g_done[c][t] = start + G
                    if t + 2 < T: try_epi(t + 2)
                pc[c] += 1
                progress = True
    end = max(pipe_free)
    return end / (T / 2), busy[0] / end

if __name__ == "__main__":
    for E in (2000, 3000, 4000, 5500, 8000, 12000):
        for C in (500, 2000):
            for sb in (1, 2):
                per, util = sim(E1=E / 2, E2=E / 2, C=C, sbuf=sb)
                print(f"E={E:6d} C={C:5d} sbuf={sb}: clk per tile pair {per:8.0f}  pipe util {util:.2f}")
