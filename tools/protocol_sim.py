"""Discrete-event check of the mbarrier protocols of the warp-specialised FlashAttention kernels.

The kernels coordinate a TMA producer warp, an MMA-issuing warp (whose tcgen05 work executes asynchronously and in
order on the tensor pipe, signalling mbarriers through tcgen05.commit) and softmax warps through phase-parity
mbarriers.  A protocol mistake shows up on hardware as a hang or as silent corruption; this model replays the exact
wait / arrive / commit sequences of a kernel variant with randomised latencies and checks

  * progress: every agent finishes (no deadlock),
  * hazards: a buffer version is only read after its write completed, and only overwritten after every read of the
    previous version completed (shared-memory tiles, tensor-memory regions).

It knows nothing about data, only about ordering.  ``python tools/protocol_sim.py`` runs all modelled variants;
``tests/test_protocol_sim.py`` does the same under pytest.
"""
from __future__ import annotations

import heapq
import random


class Barrier:
    def __init__(self, name, count):
        self.name, self.count, self.pending, self.phase, self.tx = name, count, count, 0, 0

    def done(self, parity):  # mbarrier.try_wait.parity: has the phase with this parity completed?
        return (self.phase & 1) != parity

    def arrive(self, n=1):
        self.pending -= n
        assert self.pending >= 0, f"{self.name}: more arrivals than the barrier's count"
        self._maybe_flip()

    def expect_tx(self, nbytes):  # arrive.expect_tx
        self.tx += nbytes
        self.arrive()

    def complete_tx(self, nbytes):
        self.tx -= nbytes
        self._maybe_flip()

    def _maybe_flip(self):
        if self.pending == 0 and self.tx == 0:
            self.phase += 1
            self.pending = self.count


class Buffer:
    """A storage location that is rewritten over time; versions are integers chosen by the protocol."""

    def __init__(self, name):
        self.name = name
        self.written = -1          # last version whose write completed
        self.writing = None        # version being written
        self.reads_open = {}       # version -> number of reads in flight
        self.reads_done = {}       # version -> completed reads

    def begin_write(self, v):
        assert self.writing is None, f"{self.name}: overlapping writes"
        for ver, n in self.reads_open.items():
            assert n == 0 or ver >= v, f"{self.name}: version {v} overwrites version {ver} while it is still being read"
        self.writing = v

    def end_write(self, v):
        assert self.writing == v
        self.writing, self.written = None, v

    def begin_read(self, v):
        assert self.written == v and self.writing is None, \
            f"{self.name}: read of version {v} but buffer holds {self.written} (writing {self.writing})"
        self.reads_open[v] = self.reads_open.get(v, 0) + 1

    def end_read(self, v):
        self.reads_open[v] -= 1
        self.reads_done[v] = self.reads_done.get(v, 0) + 1


class Sim:
    def __init__(self, seed=0):
        self.rng = random.Random(seed)
        self.now = 0
        self.events = []   # (time, seq, callable)
        self.seq = 0
        self.agents = []   # [name, generator, blocked_on]
        self.tensor_queue = []  # in-order async tensor pipe: list of ops
        self.tensor_busy_until = 0

    # ---- scheduling
    def at(self, dt, fn):
        self.seq += 1
        heapq.heappush(self.events, (self.now + dt, self.seq, fn))

    def spawn(self, name, gen):
        self.agents.append([name, gen, None])

    # ---- the asynchronous, in-order tensor pipe
    def mma(self, dur, reads=(), writes=()):
        """Issue an MMA: starts when the pipe is free, reads/writes are held for its duration."""
        start = max(self.now, self.tensor_busy_until)
        end = start + dur
        self.tensor_busy_until = end

        def s():
            for b, v in reads:
                b.begin_read(v)
            for b, v in writes:
                b.begin_write(v)

        def e():
            for b, v in reads:
                b.end_read(v)
            for b, v in writes:
                b.end_write(v)

        self.at(start - self.now, s)
        self.at(end - self.now, e)

    def commit(self, bar, n=1):
        """tcgen05.commit: arrive on `bar` once everything issued so far has completed."""
        self.at(max(self.now, self.tensor_busy_until) - self.now, lambda: bar.arrive(n))

    # ---- run
    def run(self, limit=10_000_000):
        live = True
        while live:
            progressed = True
            while progressed:
                progressed = False
                for a in self.agents:
                    if a[1] is None:
                        continue
                    while True:
                        if a[2] is not None:
                            kind = a[2][0]
                            if kind == "wait":
                                if not a[2][1].done(a[2][2]):
                                    break
                            elif kind == "sleep":
                                if self.now < a[2][1]:
                                    break
                            a[2] = None
                        try:
                            req = next(a[1])
                        except StopIteration:
                            a[1] = None
                            progressed = True
                            break
                        progressed = True
                        if req[0] == "wait":
                            a[2] = req
                        elif req[0] == "delay":
                            a[2] = ("sleep", self.now + req[1])
                            self.at(req[1], lambda: None)  # make time advance to the wake-up
            if all(a[1] is None for a in self.agents) and not self.events:
                return
            if not self.events:
                stuck = [(a[0], a[2][1].name if a[2] and a[2][0] == "wait" else a[2]) for a in self.agents if a[1] is not None]
                raise AssertionError(f"deadlock at t={self.now}: {stuck}")
            t, _, fn = heapq.heappop(self.events)
            assert t <= limit, "simulation ran away"
            self.now = t
            fn()

    def jitter(self, lo, hi):
        return self.rng.randint(lo, hi)


# ------------------------------------------------------------------------------------------------
# FlashAttention backward, re-ordered variant (csrc/fa2_bwd_sm100.cu: fa2_bwd_wgr_kernel)
# ------------------------------------------------------------------------------------------------
def sim_bwd_reordered(n_iter, n_warps=8, stages=2, seed=0, mutate=None):
    """``mutate`` breaks the protocol on purpose (checker self-test): "no_dq_full_wait" lets the softmax warps overwrite
    the shared P/dS tiles without waiting for the previous tile's dQ MMA; "no_dq_empty_wait" lets the MMA warp
    overwrite the dQ partial before it has been drained."""
    S = Sim(seed)
    kv_full = Barrier("kv_full", 1)
    qdo_full = [Barrier(f"qdo_full{s}", 1) for s in range(stages)]
    qdo_empty = [Barrier(f"qdo_empty{s}", 1) for s in range(stages)]
    sdp_full = Barrier("sdp_full", 1)
    pds_full = Barrier("pds_full", n_warps)
    dq_full = Barrier("dq_full", 1)
    dq_empty = Barrier("dq_empty", n_warps)
    sKV = Buffer("sKV")
    stage_buf = [Buffer(f"stage{s}") for s in range(stages)]      # Q_i, dO_i
    sPdS = Buffer("sP/sdS")
    tS = Buffer("tmem S/dP")
    tDQ = Buffer("tmem dQ")
    TILE = 1

    def tma():
        if n_iter == 0:   # the kernel's producer and MMA warps are guarded by n_iter > 0
            return
        kv_full.expect_tx(TILE)
        sKV.begin_write(0)
        S.at(S.jitter(5, 40), lambda: (sKV.end_write(0), kv_full.complete_tx(TILE)))
        stage, phase = 0, 0
        for it in range(n_iter):
            yield ("wait", qdo_empty[stage], phase ^ 1)
            qdo_full[stage].expect_tx(TILE)
            stage_buf[stage].begin_write(it)
            S.at(S.jitter(5, 60), lambda b=stage_buf[stage], v=it, f=qdo_full[stage]: (b.end_write(v), f.complete_tx(TILE)))
            stage += 1
            if stage == stages:
                stage, phase = 0, phase ^ 1
            yield ("delay", S.jitter(0, 3))

    def mma():
        if n_iter == 0:
            return
        yield ("wait", kv_full, 0)

        def issue_sdp(t):
            st = t % stages
            yield ("wait", qdo_full[st], (t // stages) & 1)
            S.mma(S.jitter(8, 12), reads=[(stage_buf[st], t), (sKV, 0)], writes=[(tS, t)])
            S.commit(sdp_full)

        yield from issue_sdp(0)
        for it in range(n_iter):
            st = it % stages
            yield ("wait", pds_full, it & 1)
            if it + 1 < n_iter:
                yield from issue_sdp(it + 1)
            S.mma(S.jitter(8, 12), reads=[(sPdS, it), (stage_buf[st], it)])  # dV, dK
            if it > 0 and mutate != "no_dq_empty_wait":
                yield ("wait", dq_empty, (it - 1) & 1)
            S.mma(S.jitter(4, 6), reads=[(sPdS, it), (sKV, 0)], writes=[(tDQ, it)])  # dQ
            S.commit(qdo_empty[st])
            S.commit(dq_full)
            yield ("delay", S.jitter(0, 2))

    def softmax(w):
        def drain(it_done):
            tDQ.begin_read(it_done)
            yield ("delay", S.jitter(2, 10))
            tDQ.end_read(it_done)
            dq_empty.arrive()

        for it in range(n_iter):
            yield ("wait", sdp_full, it & 1)
            tS.begin_read(it)
            yield ("delay", S.jitter(10, 40))   # TMEM loads + math into registers
            tS.end_read(it)
            if it > 0 and mutate != "no_dq_full_wait":
                yield ("wait", dq_full, (it - 1) & 1)
            # each warp writes its own rows of the shared P/dS tiles: model the tile write as the union of the warps
            if w == 0:
                sPdS.begin_write(it)
            yield ("delay", S.jitter(2, 8))
            wrote[it] = wrote.get(it, 0) + 1
            if wrote[it] == n_warps:
                sPdS.end_write(it)
            pds_full.arrive()
            if it > 0:
                yield from drain(it - 1)
        if n_iter > 0:
            yield ("wait", dq_full, (n_iter - 1) & 1)
            yield from drain(n_iter - 1)

    wrote = {}
    S.spawn("tma", tma())
    S.spawn("mma", mma())
    for w in range(n_warps):
        S.spawn(f"softmax{w}", softmax(w))
    S.run()
    return S.now


# ------------------------------------------------------------------------------------------------
# FlashAttention forward with K/V multicast across a CTA pair (fa2_fwd_wg2c_kernel)
# ------------------------------------------------------------------------------------------------
def sim_fwd_cluster(n_kv, slots=3, n_warps=8, seed=0, mutate=None):
    """n_kv = (tiles of CTA 0, tiles of CTA 1); tiles below min(n_kv) are shared (each CTA multicasts one half).
    ``mutate="local_release"`` (checker self-test) frees a shared slot with a CTA-local commit on a count-1 barrier, so a
    CTA may multicast into its peer's slot while the peer still reads it."""
    S = Sim(seed)
    n_shared = min(n_kv)
    HALF = 1
    cta = []
    for c in range(2):
        cta.append(dict(
            kv_full=[Barrier(f"c{c}.kv_full{s}", 1) for s in range(slots)],
            kv_empty=[Barrier(f"c{c}.kv_empty{s}", 1 if mutate == "local_release" else 2) for s in range(slots)],
            s_full=Barrier(f"c{c}.s_full", 1), p_full=Barrier(f"c{c}.p_full", n_warps), o_full=Barrier(f"c{c}.o_full", 1),
            slot=[Buffer(f"c{c}.slot{s}") for s in range(slots)], halves={},
            tS=Buffer(f"c{c}.tmemS"), sP=Buffer(f"c{c}.sP"), tO=Buffer(f"c{c}.tmemO"), wroteP={}))

    def slot_fill(c, s, use, nbytes):
        """`nbytes` of use #`use` landed in slot s of CTA c (two halves make the tile)."""
        st = cta[c]
        got = st["halves"].get((s, use), 0) + nbytes
        st["halves"][(s, use)] = got
        if got == 2 * HALF:
            st["slot"][s].end_write(use)
        st["kv_full"][s].complete_tx(nbytes)

    def tma(c):
        st = cta[c]
        slot, phase, use = 0, 0, [0] * slots
        for j in range(n_kv[c]):
            for kv in range(2):
                yield ("wait", st["kv_empty"][slot], phase ^ 1)
                st["kv_full"][slot].expect_tx(2 * HALF)
                u = use[slot]
                use[slot] += 1
                if j < n_shared:
                    for d in range(2):  # multicast: our half lands in both CTAs
                        def land(d=d, s=slot, u=u):
                            dst = cta[d]
                            if dst["halves"].get((s, u), 0) == 0 and dst["slot"][s].writing is None:
                                dst["slot"][s].begin_write(u)
                            slot_fill(d, s, u, HALF)
                        S.at(S.jitter(5, 50), land)
                else:
                    def land(s=slot, u=u, c=c):  # solo tile: the whole box comes from our own (non-multicast) load
                        cta[c]["slot"][s].begin_write(u)
                        cta[c]["slot"][s].end_write(u)
                        cta[c]["kv_full"][s].complete_tx(2 * HALF)
                    S.at(S.jitter(5, 50), land)
                slot += 1
                if slot == slots:
                    slot, phase = 0, phase ^ 1
                yield ("delay", S.jitter(0, 3))

    def mma(c):
        st = cta[c]
        if n_kv[c] == 0:
            return
        state = dict(slot=0, phase=0, jq=0, use=[0] * slots)

        def release(tile):
            s = state["slot"]
            if tile < n_shared and mutate != "local_release":
                S.commit(cta[0]["kv_empty"][s])
                S.commit(cta[1]["kv_empty"][s])
            else:
                S.commit(st["kv_empty"][s])

        def advance():
            state["use"][state["slot"]] += 1
            state["slot"] += 1
            if state["slot"] == slots:
                state["slot"], state["phase"] = 0, state["phase"] ^ 1

        def issue_qk():
            s = state["slot"]
            yield ("wait", st["kv_full"][s], state["phase"])
            S.mma(S.jitter(4, 8), reads=[(st["slot"][s], state["use"][s])], writes=[(st["tS"], state["jq"])])
            release(state["jq"])
            state["jq"] += 1
            S.commit(st["s_full"])
            advance()

        yield from issue_qk()
        for j in range(n_kv[c]):
            yield ("wait", st["p_full"], j & 1)
            s = state["slot"]
            yield ("wait", st["kv_full"][s], state["phase"])
            S.mma(S.jitter(4, 8), reads=[(st["sP"], j), (st["slot"][s], state["use"][s])], writes=[(st["tO"], j)])
            release(j)
            S.commit(st["o_full"])
            advance()
            if j + 1 < n_kv[c]:
                yield from issue_qk()

    def softmax(c, w):
        st = cta[c]
        for j in range(n_kv[c]):
            yield ("wait", st["s_full"], j & 1)
            st["tS"].begin_read(j)
            yield ("delay", S.jitter(10, 30))
            if w == 0:
                st["sP"].begin_write(j)
            yield ("delay", S.jitter(2, 6))
            st["tS"].end_read(j)
            st["wroteP"][j] = st["wroteP"].get(j, 0) + 1
            if st["wroteP"][j] == n_warps:
                st["sP"].end_write(j)
            st["p_full"].arrive()
            yield ("wait", st["o_full"], j & 1)
            st["tO"].begin_read(j)
            yield ("delay", S.jitter(2, 6))
            st["tO"].end_read(j)

    for c in range(2):
        S.spawn(f"tma{c}", tma(c))
        S.spawn(f"mma{c}", mma(c))
        for w in range(n_warps):
            S.spawn(f"softmax{c}.{w}", softmax(c, w))
    S.run()
    return S.now


# ------------------------------------------------------------------------------------------------
# FlashAttention forward with O resident in tensor memory and lazy rescaling (fa2_fwd_wg2o_kernel)
# ------------------------------------------------------------------------------------------------
def sim_fwd_o_in_tmem(n_kv, n_warps=8, seed=0, p_correct=0.5, mutate=None):
    """Each softmax warp owns its TMEM lanes of O (modelled as one buffer per warp).  PV_j accumulates into all of them;
    a warp whose rows moved their reference maximum rescales its lanes (read-modify-write) after PV_{j-1} and before it
    arrives on p_full(j), which is what releases PV_j.  ``mutate="no_o_full_wait"`` drops the explicit wait for PV_{j-1}:
    the model shows it is redundant (s_full(j) is committed after PV_{j-1} on the in-order tensor pipe), the kernel
    keeps it as cheap insurance."""
    S = Sim(seed)
    s_full, p_full, o_full = Barrier("s_full", 1), Barrier("p_full", n_warps), Barrier("o_full", 1)
    tS, sP = Buffer("tmem S"), Buffer("sP")
    tO = [Buffer(f"tmem O[{w}]") for w in range(n_warps)]
    ver = [0] * n_warps   # next version number per O buffer
    wrote = {}

    def mma():
        if n_kv == 0:
            return
        S.mma(S.jitter(4, 8), writes=[(tS, 0)])
        S.commit(s_full)
        for j in range(n_kv):
            yield ("wait", p_full, j & 1)
            # PV_j: reads P_j, accumulates into every warp's lanes of O (a write of a new version over the current one)
            start = max(S.now, S.tensor_busy_until)
            dur = S.jitter(4, 8)
            S.tensor_busy_until = start + dur

            def begin(j=j):
                sP.begin_read(j)
                for w in range(n_warps):
                    assert tO[w].writing is None and all(n == 0 for n in tO[w].reads_open.values()), \
                        f"PV_{j} starts while O[{w}] is being rescaled / read"
                    tO[w].begin_write(ver[w])

            def end(j=j):
                sP.end_read(j)
                for w in range(n_warps):
                    tO[w].end_write(ver[w])
                    ver[w] += 1

            S.at(start - S.now, begin)
            S.at(start + dur - S.now, end)
            S.commit(o_full)
            if j + 1 < n_kv:
                S.mma(S.jitter(4, 8), writes=[(tS, j + 1)])
                S.commit(s_full)

    def softmax(w):
        for j in range(n_kv):
            yield ("wait", s_full, j & 1)
            tS.begin_read(j)
            yield ("delay", S.jitter(10, 30))
            if w == 0:
                sP.begin_write(j)
            yield ("delay", S.jitter(2, 6))
            tS.end_read(j)
            wrote[j] = wrote.get(j, 0) + 1
            if wrote[j] == n_warps:
                sP.end_write(j)
            if j > 0 and S.rng.random() < p_correct:
                if mutate != "no_o_full_wait":
                    yield ("wait", o_full, (j - 1) & 1)
                cur = tO[w].written
                tO[w].begin_read(cur)
                yield ("delay", S.jitter(2, 5))
                tO[w].end_read(cur)
                tO[w].begin_write(ver[w])
                yield ("delay", S.jitter(1, 3))
                tO[w].end_write(ver[w])
                ver[w] += 1
            p_full.arrive()
        if n_kv > 0:
            yield ("wait", o_full, (n_kv - 1) & 1)
            cur = tO[w].written
            tO[w].begin_read(cur)
            yield ("delay", S.jitter(2, 5))
            tO[w].end_read(cur)

    S.spawn("mma", mma())
    for w in range(n_warps):
        S.spawn(f"softmax{w}", softmax(w))
    S.run()
    return S.now


# ------------------------------------------------------------------------------------------------
# Cross-rank flag barrier of the NVLink kernels (csrc/comm_adamw.cu: zbarrier; comm_allreduce.cu: cross_rank_barrier)
# ------------------------------------------------------------------------------------------------
def sim_cross_rank_barrier(ws=4, blocks=6, kernels=5, resident=None, seed=0, shuffled_launch=False):
    """Every rank runs the same sequence of kernels; block b of rank r synchronises with block b of every other rank at
    the start and at the end of each kernel through 1-deep flag channels (sender CAS 0→1 on the peer's word, receiver
    CAS 1→0 on its own).  ``resident`` bounds how many blocks of a rank can be on the GPU at once; with
    ``shuffled_launch`` the block scheduler of each rank picks the next block in its own random order — the case in
    which a grid that is not fully co-resident can deadlock."""
    S = Sim(seed)
    flag = {}  # (dst_rank, block, src_rank) -> 0/1

    class Chan:
        def __init__(self, key):
            self.key, self.name = key, f"flag{key}"

    def send(dst, b, src):
        while flag.get((dst, b, src), 0) != 0:
            yield ("delay", 1)
        flag[(dst, b, src)] = 1

    def recv(me, b, src):
        while flag.get((me, b, src), 0) != 1:
            yield ("delay", 1)
        flag[(me, b, src)] = 0

    slots = {r: (resident or blocks) for r in range(ws)}
    done_blocks = {(r, k): 0 for r in range(ws) for k in range(kernels)}
    reads = {}

    def block(r, k, b):
        # kernel k starts on rank r only after kernel k-1 finished there (same stream)
        while k > 0 and done_blocks[(r, k - 1)] < blocks:
            yield ("delay", 1)
        while slots[r] == 0:
            yield ("delay", 1)
        slots[r] -= 1
        for phase in range(2):
            for peer in range(ws):
                yield from send(peer, b, r)
            for peer in range(ws):
                yield from recv(r, b, peer)
            if phase == 0:
                reads[(r, k, b)] = S.now       # after the first barrier every rank has entered kernel k
                yield ("delay", S.jitter(1, 20))
        slots[r] += 1
        done_blocks[(r, k)] += 1

    for r in range(ws):
        for k in range(kernels):
            order = list(range(blocks))
            if shuffled_launch:
                S.rng.shuffle(order)
            for b in order:
                S.spawn(f"r{r}k{k}b{b}", block(r, k, b))
    S.run(limit=200_000)
    return S.now


def main():
    for n in range(0, 7):
        for seed in range(20):
            sim_bwd_reordered(n, seed=seed)
    print("backward, re-ordered: ok (n_iter 0..6, 20 seeds)")
    for a, b in [(1, 1), (2, 1), (1, 2), (4, 4), (5, 4), (4, 5), (9, 8), (8, 9), (16, 16)]:
        for seed in range(20):
            sim_fwd_cluster((a, b), seed=seed)
    print("forward, cluster multicast: ok")
    for n in range(0, 9):
        for seed in range(20):
            sim_fwd_o_in_tmem(n, seed=seed)
    print("forward, O in tensor memory with lazy rescale: ok")
    for seed in range(10):
        sim_cross_rank_barrier(ws=4, blocks=6, kernels=4, seed=seed)
        sim_cross_rank_barrier(ws=3, blocks=8, kernels=3, resident=3, seed=seed)  # in-order launch: partial residency is fine
    print("cross-rank flag barrier: ok")


if __name__ == "__main__":
    main()
