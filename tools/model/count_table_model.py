"""Model check (random interleavings, sequentially consistent memory) of the global count table's
lock-free insert (csrc/vocab.cu: table_add): a slot is claimed by CAS on word a; the claimer copies the
term's bytes to the arena, stores strref, adds its count and only then publishes word b; whoever finds
its a waits for b, moves on if b differs and -- for a fingerprint key -- compares the BYTES with the
arena copy and moves on if they differ.  Terms are drawn so that many share (a, b) (truncated
fingerprints, like txf_debug_set_identity_bits) and many share the home slot.

Checked after every run: one slot per distinct term, its count = the occurrences inserted, no two
terms merged, every returned slot holds the caller's term.

    python tools/model/count_table_model.py [--no-verify]     # --no-verify: drop the byte compare -> terms merge

CPU only; not part of the product."""
import random
import sys


def run(nthreads, nslots, terms, rng, verify=True):
    """terms: list of (a, b, bytes) each thread inserts in turn (with repetition across threads)."""
    A, B, REF, CNT = [0] * nslots, [0] * nslots, [None] * nslots, [0] * nslots
    arena = {}
    returned = []

    def insert(a, b, s):
        idx = a % nslots
        for _ in range(nslots):
            cur = A[idx]
            yield
            if cur == 0:
                cur = A[idx]          # atomicCAS(&slot.a, 0, a)
                if cur == 0:
                    A[idx] = a
                    yield
                    ref = len(arena) + 1
                    arena[ref] = bytearray()
                    for ch in s:      # byte-by-byte copy, preemptible
                        arena[ref].append(ch)
                        yield
                    REF[idx] = ref
                    yield
                    CNT[idx] += 1
                    yield             # __threadfence()
                    B[idx] = b
                    returned.append((idx, s))
                    return
            if cur != a:
                idx = (idx + 1) % nslots
                continue
            while B[idx] == 0:
                yield
            if B[idx] != b:
                idx = (idx + 1) % nslots
                continue
            if verify:
                ref = REF[idx]
                yield
                if bytes(arena[ref]) != s:
                    idx = (idx + 1) % nslots
                    continue
            CNT[idx] += 1             # atomicAdd
            returned.append((idx, s))
            return
        raise AssertionError("table full")

    def thread(work):
        for a, b, s in work:
            yield from insert(a, b, s)
            yield

    work = [[rng.choice(terms) for _ in range(rng.randint(1, 12))] for _ in range(nthreads)]
    agents = [thread(w) for w in work]
    steps = 0
    while agents:
        steps += 1
        assert steps < 2_000_000, "livelock"
        i = rng.randrange(len(agents))
        try:
            next(agents[i])
        except StopIteration:
            agents.pop(i)
    want = {}
    for w in work:
        for _, _, s in w:
            want[s] = want.get(s, 0) + 1
    got = {}
    for i in range(nslots):
        if A[i]:
            s = bytes(arena[REF[i]])
            if s in got:
                return f"two slots for {s!r}"
            got[s] = CNT[i]
    if got != want:
        return "counts differ (terms merged or occurrences lost)"
    for idx, s in returned:
        if bytes(arena[REF[idx]]) != s:
            return "a caller was handed another term's slot"
    return "ok"


def make_terms(rng, n, id_space, home_space):
    out = []
    for i in range(n):
        s = bytes(rng.randrange(97, 123) for _ in range(rng.randint(16, 24))) + bytes([i])
        a = 1 + rng.randrange(id_space) * home_space + rng.randrange(home_space)
        out.append((a, 1 + rng.randrange(id_space), s))
    return out


def main():
    verify = "--no-verify" not in sys.argv
    rng = random.Random(20190501)
    bad = {}
    runs = 400
    for _ in range(runs):
        terms = make_terms(rng, rng.randint(2, 14), rng.choice([1, 2, 8]), rng.choice([1, 4]))
        r = run(rng.randint(2, 8), 64, terms, rng, verify)
        if r != "ok":
            bad[r] = bad.get(r, 0) + 1
    print(f"{'with' if verify else 'WITHOUT'} the byte compare: {runs} random schedules, failures: {bad or 'none'}")
    return 1 if (bad and verify) else 0


if __name__ == "__main__":
    sys.exit(main())
