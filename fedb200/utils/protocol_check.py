"""Exhaustive interleaving check of the peer-memory flag protocol (csrc/peer_flags.cuh: start_barrier / end_barrier).

``compute-sanitizer --tool racecheck`` cannot see cross-GPU ordering, and it is closed on the GPU pool this repository was
developed on, so the protocol that every fused collective relies on is checked here as a small explicit-state model instead:
R ranks, C thread blocks (CTAs) per launch, E consecutive launches per rank, at most K CTAs of a rank resident at a time
(K < C models a grid that is NOT co-resident — the case the claim-based start barrier was designed for).  Every
interleaving of the atomic steps is explored (breadth-first over the reachable states) and four properties are asserted:

* **start safety** — a CTA leaves the start barrier of epoch e only after every rank has entered launch e (its input is
  complete) — and the weight words it then reads are the ones published for epoch e;
* **end safety** — a rank's launch e completes only after every CTA of every rank has issued its pushes of epoch e;
* **no reset races** — the finished-block counter is back at zero whenever a launch starts;
* **progress** — no reachable state is stuck (spins are modelled as blocking waits, so a protocol that could hang shows
  up as a deadlocked state instead of a timeout).

The steps mirror the device code one to one: ``atomicMax(claim, e)`` decides who signals, the signalling threads (one per
peer) store the weight and then the epoch with release order into the PEER's buffer, every CTA polls its OWN buffer, the
last CTA to finish (``atomicAdd(done)``) resets the counter and runs the cross-rank end barrier.
"""
from __future__ import annotations

from collections import deque
from typing import Dict, List, Tuple

# per-CTA program counters
PENDING, CLAIM, START, WORK, COUNT, END, EXITED = range(7)


def check(ranks: int = 2, ctas: int = 2, launches: int = 2, resident: int = 1, max_states: int = 5_000_000,
          mutation: str = "") -> Dict[str, int]:
    """Explore every interleaving; raises AssertionError with a trace summary when a property fails.

    ``mutation`` breaks the protocol on purpose (the tests use it to show that each property can fail): ``"no_end_wait"`` —
    the last block signals the end of the epoch but does not wait for the peers; ``"epoch_before_weight"`` — the signalling
    thread stores the epoch first and the weight afterwards."""
    R, C, E, K = ranks, ctas, launches, resident
    # ---- state -----------------------------------------------------------------------------------------------
    # launch[r]                      current launch (epoch) of rank r, 1..E, E+1 = finished
    # pc[r][c]                       program counter of CTA c of rank r in the current launch
    # first[r][c], last[r][c]        outcome of the claim / the finished-block count
    # sig[r][c][p]                   START/END: 0 = thread p has not signalled, 1 = weight stored, 2 = epoch stored (or not its job)
    # got[r][c][p]                   START/END: thread p has seen the peer's epoch
    # flags: start[r][p], end[r][p], weight[r][p] (epoch tag), claim[r], done[r]; pushed[r] = CTAs of the launch past WORK
    def initial():
        z = tuple(tuple(0 for _ in range(R)) for _ in range(R))
        cta = (PENDING, 0, 0, (0,) * R, (0,) * R)
        return (tuple(1 for _ in range(R)), tuple(tuple(cta for _ in range(C)) for _ in range(R)), z, z, z,
                tuple(0 for _ in range(R)), tuple(0 for _ in range(R)), tuple(0 for _ in range(R)))

    def set2(m, i, j, v):
        row = list(m[i]); row[j] = v
        out = list(m); out[i] = tuple(row)
        return tuple(out)

    def set1(t, i, v):
        out = list(t); out[i] = v
        return tuple(out)

    def successors(s):
        launch, ctas_s, start, end, weight, claim, done, pushed = s
        for r in range(R):
            e = launch[r]
            if e > E:
                continue
            n_res = sum(1 for c in ctas_s[r] if c[0] not in (PENDING, EXITED))
            for ci, (pc, first, last, sig, got) in enumerate(ctas_s[r]):
                def put(new_cta, **kw):
                    d = dict(launch=launch, start=start, end=end, weight=weight, claim=claim, done=done, pushed=pushed)
                    d.update(kw)
                    row = list(ctas_s[r]); row[ci] = new_cta
                    cs = list(ctas_s); cs[r] = tuple(row)
                    return (d["launch"], tuple(cs), d["start"], d["end"], d["weight"], d["claim"], d["done"], d["pushed"])
                if pc == PENDING:
                    if n_res < K:                       # a slot is free: the block scheduler starts this CTA
                        assert done[r] == 0 or any(c[0] not in (PENDING,) for c in ctas_s[r]), "finished-block counter not reset at launch"
                        yield put((CLAIM, 0, 0, (0,) * R, (0,) * R))
                elif pc == CLAIM:                       # s_first = atomicMax(claim, e) < e
                    f = 1 if claim[r] < e else 0
                    yield put((START, f, 0, tuple(0 if f else 2 for _ in range(R)), (0,) * R), claim=set1(claim, r, max(claim[r], e)))
                elif pc in (START, END):
                    signalling = first if pc == START else last
                    flagm = start if pc == START else end
                    for p in range(R):                  # thread p of the block
                        swap = mutation == "epoch_before_weight" and pc == START
                        w_at, e_at = (1, 0) if swap else (0, 1)
                        if signalling and pc == START and sig[p] == w_at:      # weight word first (plain store) ...
                            yield put((pc, first, last, set1(sig, p, 2 if swap else 1), got), weight=set2(weight, p, r, e))
                        elif signalling and sig[p] == (e_at if pc == START else 0):   # ... then the epoch, release order
                            nf = set2(flagm, p, r, e)
                            yield put((pc, first, last, set1(sig, p, 1 if swap else 2), got), **({"start": nf} if pc == START else {"end": nf}))
                        elif sig[p] == 2 and not got[p] and (flagm[r][p] >= e or (mutation == "no_end_wait" and pc == END)):
                            yield put((pc, first, last, sig, set1(got, p, 1)))      # poll of the rank's OWN buffer succeeds
                    if all(got) and all(x == 2 for x in sig):               # __syncthreads(): the block leaves the barrier
                        if pc == START:
                            for p in range(R):
                                assert launch[p] >= e, f"start safety: rank {r} passed epoch {e} before rank {p} entered it"
                                assert weight[r][p] == e, f"rank {r} read rank {p}'s weight of epoch {weight[r][p]} in epoch {e}"
                            yield put((WORK, first, last, (0,) * R, (0,) * R))
                        else:
                            yield put((EXITED, 0, 0, (0,) * R, (0,) * R))
                elif pc == WORK:                        # reduce + push to the peers, then threadfence
                    yield put((COUNT, first, last, sig, got), pushed=set1(pushed, r, pushed[r] + 1))
                elif pc == COUNT:                       # is_last = atomicAdd(done, 1) == gridDim.x - 1
                    if done[r] == C - 1:
                        yield put((END, first, 1, (0,) * R, (0,) * R), done=set1(done, r, 0))          # last block resets the counter
                    else:
                        yield put((EXITED, 0, 0, (0,) * R, (0,) * R), done=set1(done, r, done[r] + 1))
            if all(c[0] == EXITED for c in ctas_s[r]):                      # the launch has completed on this rank's stream
                for p in range(R):
                    assert launch[p] > e or (launch[p] == e and pushed[p] == C), \
                        f"end safety: rank {r} finished epoch {e} while rank {p} had pushed {pushed[p]}/{C} blocks"
                fresh = tuple((PENDING, 0, 0, (0,) * R, (0,) * R) for _ in range(C))
                cs = list(ctas_s); cs[r] = fresh
                yield (set1(launch, r, e + 1), tuple(cs), start, end, weight, claim, done, set1(pushed, r, 0))

    s0 = initial()
    seen = {s0}
    todo = deque([s0])
    finals = 0
    while todo:
        s = todo.popleft()
        nxt = list(successors(s))
        if not nxt:
            if all(l > E for l in s[0]):
                finals += 1
                continue
            raise AssertionError(f"progress: stuck state with launches {s[0]} and CTA states "
                                 f"{[[c[0] for c in row] for row in s[1]]}")
        for t in nxt:
            if t not in seen:
                if len(seen) >= max_states:
                    raise RuntimeError(f"state space exceeds {max_states} states")
                seen.add(t)
                todo.append(t)
    return {"states": len(seen), "final_states": finals, "ranks": R, "ctas": C, "launches": E, "resident": K}


def check_broken_variant() -> str:
    """The same exploration with the protocol's key property removed — a start barrier in which only block 0 signals and the
    other blocks wait for IT (the round-1 design) — under one resident CTA: must be reported as stuck.  Used by the test to
    show that the checker can fail."""
    # blocks: 0 signals + releases the others through a local GO word; block 1 waits for GO.  With one resident slot and
    # block 1 scheduled first, nobody ever writes GO.
    R, C = 2, 2
    # state: (running block or None, go, finished blocks)
    frontier, seen = deque([(None, 0, frozenset())]), set()
    while frontier:
        run, go, fin = frontier.popleft()
        if (run, go, fin) in seen:
            continue
        seen.add((run, go, fin))
        nxt: List[Tuple] = []
        if run is None:
            nxt += [(b, go, fin) for b in range(C) if b not in fin]
        elif run == 0:
            nxt.append((None, 1, fin | {0}))
        elif go:
            nxt.append((None, go, fin | {run}))
        if not nxt and len(fin) < C:
            return f"stuck: block {run} resident, GO={go}, finished {sorted(fin)}"
        frontier.extend(nxt)
    return ""


def check_ring(ring: int = 2, chunks: int = 2, versions: int = 4, invalidate: bool = True) -> Dict[str, int]:
    """The asynchronous outbox ring (p2p/observed.py + csrc/gossip.cu): a producer publishes versions 1..V into slot v % R,
    chunk by chunk (bytes, then the chunk's landing flag = v, release order); a reader picks a published version, and per chunk
    waits for flag >= v, reads the bytes, and re-reads the flag: a value other than v counts the chunk as torn (the mix is then
    discarded).  Property: a chunk the reader ACCEPTS holds exactly the bytes of version v.  With ``invalidate`` the producer
    first retires the slot's flags (0) before it overwrites the bytes of version v - R, as the runtime does; without it the
    exploration finds the lapped read that goes unnoticed."""
    R, C, V = ring, chunks, versions
    # producer: (version being published, phase, chunk): phase 0 = retire flags (per chunk), 1 = write bytes, 2 = set flag
    # memory:   data[slot][chunk] = version of the bytes, "mid" while half written (tuple (old, new)); flag[slot][chunk]
    # reader:   (version chosen or 0, chunk, phase): 0 = choose, 1 = wait flag, 2 = read bytes (value latched), 3 = re-check
    init = ((1, 0 if invalidate else 1, 0), tuple(tuple(0 for _ in range(C)) for _ in range(R)),
            tuple(tuple(0 for _ in range(C)) for _ in range(R)), 0, (0, 0, 0, None))
    seen, todo = {init}, deque([init])
    accepted = torn = 0
    while todo:
        prod, data, flag, pub, rd = todo.popleft()
        nxt = []
        v, ph, c = prod
        if v <= V:
            sl = v % R
            if ph == 0:                                   # retire the landing flag of chunk c (only if the slot was used before)
                nf = tuple(tuple(0 if (i == sl and j == c and v > R) else flag[i][j] for j in range(C)) for i in range(R))
                nxt.append((((v, 0, c + 1) if c + 1 < C else (v, 1, 0)), data, nf, max(pub, v), rd))      # `pub` = v was stored first
            elif ph == 1:                                 # the bytes of chunk c change (first half, then second half)
                cur = data[sl][c]
                new = (cur, v) if not isinstance(cur, tuple) else v
                nd = tuple(tuple(new if (i == sl and j == c) else data[i][j] for j in range(C)) for i in range(R))
                nxt.append(((v, 1, c) if isinstance(new, tuple) else (v, 2, c), nd, flag, max(pub, v), rd))
            else:                                         # flag[c] = v, release
                nf = tuple(tuple(v if (i == sl and j == c) else flag[i][j] for j in range(C)) for i in range(R))
                nxt.append((((v, 1, c + 1) if c + 1 < C else (v + 1, 0 if invalidate else 1, 0)), data, nf, pub, rd))
        rv, rc, rph, latched = rd
        if rph == 0:
            if pub > 0:
                nxt.append((prod, data, flag, pub, (pub, 0, 1, None)))          # the version decision (logged for replay)
        elif rph == 1:
            if flag[rv % R][rc] >= rv:
                nxt.append((prod, data, flag, pub, (rv, rc, 2, None)))
        elif rph == 2:
            nxt.append((prod, data, flag, pub, (rv, rc, 3, data[rv % R][rc])))
        elif rph == 3:
            if flag[rv % R][rc] == rv:                    # accepted
                assert latched == rv, (f"undetected torn read: version {rv} chunk {rc} accepted with bytes {latched} "
                                       f"(ring {R}, producer at version {v})")
                accepted += 1
            else:
                torn += 1
            nxt.append((prod, data, flag, pub, (rv, rc + 1, 1, None) if rc + 1 < C else (0, 0, 0, None)))
        for t in nxt:
            if t not in seen:
                seen.add(t)
                todo.append(t)
    return {"states": len(seen), "accepted_reads": accepted, "torn_reads_detected": torn, "ring": R, "chunks": C, "versions": V}


if __name__ == "__main__":
    import json
    for cfg in ((2, 2, 2, 1), (2, 2, 2, 2), (3, 1, 2, 1), (2, 3, 1, 1)):
        print(json.dumps(check(*cfg)))
    print("round-1 design under one resident CTA:", check_broken_variant())
    print(json.dumps(check_ring(2, 2, 4, True)))
    try:
        check_ring(2, 2, 4, False)
    except AssertionError as ex:
        print("ring without retiring the flags first:", ex)
