"""Randomised replay of the barrier protocol of fast_attn_ws.cu (producer, two MMA issuers, helper, two softmax groups)
with mbarrier PHASE-PARITY semantics and asynchronous (optionally out-of-order) TMA landings / in-order MMA commits.
Reports false positives (a waiter a phase ahead), missed completions and deadlocks.  python profiles/scripts/attn_ws_protocol_sim.py"""
# randomized simulation with mbarrier PARITY semantics and asynchronous TMA landings / MMA commits
import random, sys
def sim(nt, ring, n_it, seed, NS=8, ooo_tma=False):
    rnd = random.Random(seed)
    C = {}                       # barrier -> completions so far
    def comp(name): return C.get(name, 0)
    tma_q, mma_q = [], []        # in-flight async completions (barrier names)
    problems = []
    def wait_ok(name, k, who):
        # wants completion #k (1-based); parity p = (k-1)&1; HW: true iff p != (current phase index & 1)
        cur = comp(name)
        hw = ((k - 1) & 1) != (cur & 1)
        truth = cur >= k
        if hw and not truth: problems.append(('FALSE-POSITIVE', who, name, k, cur))
        if truth and not hw: problems.append(('MISSED', who, name, k, cur))
        return hw
    def producer():
        cnt = 0
        for it in range(n_it):
            for ps in range(2):
                for t in range(nt):
                    slot, use = cnt % ring, cnt // ring
                    if use > 0: yield ('wait', 'empty%d' % slot, use)
                    yield ('tma', 'full%d' % slot)
                    cnt += 1
    def s_issuer():
        for it in range(n_it):
            yield ('wait', 'qfull%d' % (it & 3), (it >> 2) + 1)
            for t in range(nt):
                cnt = it * 2 * nt + t; slot = cnt % ring
                n = it * nt + t; buf = n % NS
                if cnt // ring > 0: yield ('wait', 'empty%d' % slot, cnt // ring)
                yield ('wait', 'full%d' % slot, cnt // ring + 1)
                if n >= NS: yield ('wait', 'sempty%d' % buf, n // NS)
                yield ('mma', 'sfull%d' % buf); yield ('mma', 'empty%d' % slot)
    def pv_issuer():
        for it in range(n_it):
            g, my = it & 1, it >> 1
            for t in range(nt):
                cnt = it * 2 * nt + nt + t; slot = cnt % ring
                pb_cnt = my * nt + t; pb = pb_cnt & 1
                yield ('wait', 'pfull%d_%d' % (g, pb), (pb_cnt >> 1) + 1)
                yield ('wait', 'full%d' % slot, cnt // ring + 1)
                if t == 0 and my > 0: yield ('wait', 'oempty%d' % g, my)
                yield ('mma', 'empty%d' % slot); yield ('mma', 'pempty%d_%d' % (g, pb))
                if t == nt - 1: yield ('mma', 'ofull%d' % g)
    def helper():
        for it in range(n_it):
            if it >= 4: yield ('wait', 'done%d' % (it & 3), it >> 2)
            yield ('fire', 'qfull%d' % (it & 3))
    def group(g):
        my = 0
        for it in range(g, n_it, 2):
            yield ('wait', 'qfull%d' % (it & 3), (it >> 2) + 1)
            for t in range(nt):
                n = it * nt + t; buf = n % NS
                yield ('wait', 'sfull%d' % buf, n // NS + 1)
                yield ('fire', 'sempty%d' % buf)
            for t in range(nt):
                pb_cnt = my * nt + t; pb = pb_cnt & 1
                if pb_cnt >= 2: yield ('wait', 'pempty%d_%d' % (g, pb), pb_cnt >> 1)
                yield ('fire', 'pfull%d_%d' % (g, pb))
            if nt > 0:
                yield ('wait', 'ofull%d' % g, my + 1)
                yield ('fire', 'oempty%d' % g)
            yield ('fire', 'done%d' % (it & 3))
            my += 1
    agents = {'prod': producer(), 'S': s_issuer(), 'PV': pv_issuer(), 'help': helper(), 'g0': group(0), 'g1': group(1)}
    pending = {k: None for k in agents}
    alive = set(agents)
    steps = 0
    while alive or tma_q or mma_q:
        steps += 1
        if steps > 200000: return ('LIVELOCK', problems[:3])
        choices = list(alive) + (['_tma'] if tma_q else []) + (['_mma'] if mma_q else [])
        k = rnd.choice(choices)
        if k == '_tma':
            i = rnd.randrange(len(tma_q)) if ooo_tma else 0
            n = tma_q.pop(i); C[n] = comp(n) + 1; continue
        if k == '_mma':
            n = mma_q.pop(0); C[n] = comp(n) + 1; continue
        if pending[k] is None:
            try: pending[k] = next(agents[k])
            except StopIteration:
                alive.discard(k); continue
        ev = pending[k]
        if ev[0] == 'fire': C[ev[1]] = comp(ev[1]) + 1; pending[k] = None
        elif ev[0] == 'tma': tma_q.append(ev[1]); pending[k] = None
        elif ev[0] == 'mma': mma_q.append(ev[1]); pending[k] = None
        elif wait_ok(ev[1], ev[2], k): pending[k] = None
        else:
            # blocked: is anything else able to move?  detect global deadlock
            if not tma_q and not mma_q:
                stuck = True
                for a in alive:
                    e = pending[a]
                    if e is None or e[0] != 'wait' or ((e[2] - 1) & 1) != (comp(e[1]) & 1): stuck = False; break
                if stuck: return ('DEADLOCK', {a: pending[a] for a in alive}, problems[:3])
        if problems: return ('PARITY', problems[:3])
    return None
bad = 0
for ring in (4, 6, 10, 12, 24):
    for nt in list(range(1, 13)) + [15, 21, 22, 29]:
        if not (nt < 30 and ring >= nt + 2): continue
        for n_it in (1, 2, 6, 7, 9, 40):
            for seed in range(40):
                r = sim(nt, ring, n_it, seed, NS=30, ooo_tma=(seed % 2 == 1))
                if r:
                    print("ring", ring, "nt", nt, "items", n_it, "seed", seed, r); bad += 1; break
print("violations:", bad)
