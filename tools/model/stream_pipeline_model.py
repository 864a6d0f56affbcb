"""Model check (random interleavings) of stream_ops_kernel's producer / consumer protocol
(csrc/stream_ops.cu): 5 stages, full[s] completed by the producer's bulk loads, empty[s] completed
by one arrival per consumer warp -- the elected thread's one tile later, when the bulk store has
read the stage.  Every agent is a generator that yields whenever it would block or be preempted;
the scheduler picks agents at random (optionally starving one warp for long stretches).  Checked:
no deadlock, and every consumer warp sees every tile exactly once and in order.

    python tools/model/stream_pipeline_model.py            # the protocol as shipped
    python tools/model/stream_pipeline_model.py --old      # empty[s] with ONE arrival: a starved warp misses a phase

CPU only; not part of the product."""
import random
import sys

STAGES = 5


class Mbar:
    def __init__(self, count):
        self.count, self.pending, self.phase, self.tx = count, count, 0, 0

    def arrive(self, tx=0):
        self.tx += tx
        self.pending -= 1
        self._maybe()

    def complete_tx(self, n):
        self.tx -= n
        self._maybe()

    def _maybe(self):
        assert self.pending >= 0, "more arrivals than the barrier expects in one phase"
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count

    def passed(self, parity):  # mbarrier.try_wait.parity
        return (self.phase & 1) != parity


def run(ntiles, fix_prob, old, starve, rng):
    full = [Mbar(1) for _ in range(STAGES)]
    empty = [Mbar(1 if old else 3) for _ in range(STAGES)]
    job = [None] * STAGES
    inflight = []           # bulk loads: (stage,)
    stores = []             # elected thread's bulk groups, oldest first: [stage or None, read_done]
    seen = {w: [] for w in (1, 2, 3)}
    bar = {"n": 0, "gen": 0}  # named barrier of the three consumer warps

    def producer():
        for k in range(ntiles + 1):
            s = k % STAGES
            while not empty[s].passed(((k // STAGES) & 1) ^ 1):
                yield
            if k == ntiles:
                job[s] = -1
                full[s].arrive()
                return
            job[s] = (k, rng.random() < fix_prob)
            full[s].arrive(tx=1)
            inflight.append(s)
            yield

    def consumer(w):
        prev = None
        k = 0
        while True:
            s = k % STAGES
            while not full[s].passed((k // STAGES) & 1):
                yield
            d = job[s]
            yield  # the warp may be preempted between the wait and anything else
            if d == -1:
                return
            seen[w].append(d[0])
            if d[1]:  # fix-up tile: all consumer warps meet at the named barrier
                gen = bar["gen"]
                bar["n"] += 1
                if bar["n"] == 3:
                    bar["n"], bar["gen"] = 0, gen + 1
                while bar["gen"] == gen:
                    yield
            if w != 1:
                if not old:
                    empty[s].arrive()
            else:
                stores.append([s, False])
                if prev is not None:
                    while not all(g[1] for g in stores[:-1]):  # cp.async.bulk.wait_group.read 1
                        yield
                    empty[prev].arrive()
                prev = s
            k += 1
            yield

    agents = {"P": producer(), 1: consumer(1), 2: consumer(2), 3: consumer(3)}
    starving, until = None, 0
    step = 0
    while agents:
        step += 1
        if step > 400 * (ntiles + 10):
            return "deadlock"
        # hardware progress: loads land, stores get read, in random order
        if inflight and rng.random() < 0.5:
            full[inflight.pop(0)].complete_tx(1)
        for g in stores:
            if not g[1] and rng.random() < 0.5:
                g[1] = True
                break
        if starve and step >= until:
            starving = rng.choice([2, 3, None])
            until = step + rng.randint(1, 60)
        names = [n for n in agents if n != starving] or list(agents)
        n = rng.choice(names)
        try:
            next(agents[n])
        except StopIteration:
            del agents[n]
        if not agents:
            break
    for w in (1, 2, 3):
        if seen[w] != list(range(ntiles)):
            return f"warp {w} saw {len(seen[w])} of {ntiles} tiles (first gap at {next((i for i, v in enumerate(seen[w]) if v != i), len(seen[w]))})"
    return "ok"


def main():
    old = "--old" in sys.argv
    rng = random.Random(20190501)
    bad = {}
    runs = 3000
    for i in range(runs):
        r = run(rng.randint(0, 60), rng.choice([0.0, 0.0, 0.3, 1.0]), old, starve=True, rng=rng)
        if r != "ok":
            bad[r.split(" (")[0].split(" saw")[0]] = bad.get(r.split(" (")[0].split(" saw")[0], 0) + 1
    print(f"{'one arrival on empty[s] (before)' if old else 'one arrival per consumer warp (shipped)'}: {runs} random schedules, failures: {bad or 'none'}")
    return 1 if (bad and not old) else 0


if __name__ == "__main__":
    sys.exit(main())
