"""Slot-occupancy model of matchingPursuitTeamsKernel's block scheduling (csrc/mp_kernel_teams.cu): 148 CTAs x 2 teams x 32
slots, blocks handed out in order from one global counter per colour channel, one greedy iteration per slot and trip, a team
leaves a channel only when ALL its slots are idle ("drain"). Compared with the bound sum(iterations) / slots and with a
hypothetical scheduler that never drains between channels ("mixed"). Input: the per-block-channel counts of an encode as a
.npy of shape (nblocks, 3) — `np.save(path, mn.lastSymbols(ctx, w, h)[0])` after an encode on a GPU (for the figures in
DESIGN.md §3.1 the counts came from the CPU checker, which is symbol-identical).

    python tools/mp_schedule_sim.py counts.npy
"""
import sys, heapq, numpy as np
counts = np.load(sys.argv[1])           # (nblocks, 3)
K = 32
TEAMS = 148 * 2
SLOTS = 32
FIFO = 128
iters = np.where(counts < K, counts + 1, K).astype(np.int64)   # greedy iterations a block-channel runs (the q == 0 one included)
nblocks = counts.shape[0]
print("blocks", nblocks, "mean iterations per bc", iters.mean(), "per channel", iters.mean(axis=0), "max", iters.max(axis=0))
ideal = iters.sum() / (TEAMS * SLOTS)
def simulate(mode):
    # event-driven in units of trips, all teams advance in lock step (same trip time); a team's clock = its trip count
    # mode "drain": a team leaves a channel only when ALL its slots are idle (today's kernel)
    # mode "mixed": a slot that finds its channel exhausted takes a block of the next channel at once (no drain between channels)
    ctr = [0, 0, 0]
    def take(ch):
        if ctr[ch] < nblocks:
            b = ctr[ch]; ctr[ch] += 1; return b
        return None
    tclock = np.zeros(TEAMS, np.int64)
    busy_slot_trips = 0
    # per team: channel, remaining iterations per slot
    team_ch = [0] * TEAMS
    rem = np.zeros((TEAMS, SLOTS), np.int64)
    slot_ch = np.zeros((TEAMS, SLOTS), np.int64)
    # process teams round-robin by trip to share the global counters fairly
    active = set(range(TEAMS))
    # initial fill
    for t in range(TEAMS):
        for s in range(SLOTS):
            b = take(0)
            if b is not None: rem[t, s] = iters[b, 0]
    trip = 0
    done_at = np.zeros(TEAMS, np.int64)
    while active:
        trip += 1
        for t in list(active):
            r = rem[t]
            act = r > 0
            busy_slot_trips += int(act.sum())
            r[act] -= 1
            # refill finished slots
            for s in np.nonzero(act & (r == 0))[0]:
                if mode == "drain":
                    b = take(team_ch[t])
                    if b is not None: r[s] = iters[b, team_ch[t]]
                else:
                    ch = int(slot_ch[t, s])
                    while ch < 3:
                        b = take(ch)
                        if b is not None:
                            r[s] = iters[b, ch]; slot_ch[t, s] = ch; break
                        ch += 1
            if not (r > 0).any():
                if mode == "drain":
                    nxt = team_ch[t] + 1
                    filled = False
                    while nxt < 3 and not filled:
                        team_ch[t] = nxt
                        for s in range(SLOTS):
                            b = take(nxt)
                            if b is not None: r[s] = iters[b, nxt]; filled = True
                        if not filled: nxt += 1
                    if not filled:
                        active.discard(t); done_at[t] = trip
                else:
                    active.discard(t); done_at[t] = trip
    return trip, done_at, busy_slot_trips
for mode in ("drain", "mixed"):
    trips, done_at, busy = simulate(mode)
    print(mode, "kernel length", trips, "trips; ideal", round(ideal, 1), "; efficiency", round(ideal / trips, 3),
          "; team finish min/mean/max", done_at.min(), round(done_at.mean(), 1), done_at.max(),
          "; slot occupancy", round(busy / (trips * TEAMS * SLOTS), 3))
